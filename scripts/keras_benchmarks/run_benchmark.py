#!/usr/bin/env python
"""Small-model epoch benchmarks (counterpart of the reference's deprecated second project,
scripts/keras_benchmarks: K/run_benchmark.py, K/models/*_benchmark.py, K/data_generator.py).

Three models on synthetic data, timed per epoch with the first epoch discarded as warm-up
(K/models/timehistory.py:5-14):

* ``mnist_mlp``   784-512-512-10 MLP with dropout 0.2, RMSprop, batch 128
* ``cifar10_cnn`` conv32-conv32-pool-conv64-conv64-pool-FC512-FC10 with dropout, RMSprop(1e-4)
* ``lstm``        LSTM(128) over 40 one-hot steps of 60 symbols + Dense(60), RMSprop(0.01)

The Keras/TF/Theano/CNTK back-ends of the original are replaced by PyTorch (these tiny models
are launch-bound; they are not part of the B200 kernel work). Results go to a JSON line per
model (the original uploaded the same fields to BigQuery, K/upload_benchmarks_bq.py:7-80).

    python scripts/keras_benchmarks/run_benchmark.py --model mnist_mlp --device cpu
"""
import argparse
import json
import time

import torch
import torch.nn as nn


def synthetic(model, n, device):
    g = torch.Generator().manual_seed(0)
    if model == "mnist_mlp":
        return torch.rand(n, 784, generator=g).to(device), torch.randint(0, 10, (n,), generator=g).to(device)
    if model == "cifar10_cnn":
        return torch.rand(n, 3, 32, 32, generator=g).to(device), torch.randint(0, 10, (n,), generator=g).to(device)
    x = torch.zeros(n, 40, 60)
    idx = torch.randint(0, 60, (n, 40), generator=g)
    x.scatter_(2, idx.unsqueeze(-1), 1.0)
    return x.to(device), torch.randint(0, 60, (n,), generator=g).to(device)


class LstmNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.lstm = nn.LSTM(60, 128, batch_first=True)
        self.fc = nn.Linear(128, 60)

    def forward(self, x):
        out, _ = self.lstm(x)
        return self.fc(out[:, -1])


def build(model):
    if model == "mnist_mlp":
        return nn.Sequential(nn.Linear(784, 512), nn.ReLU(), nn.Dropout(0.2), nn.Linear(512, 512),
                             nn.ReLU(), nn.Dropout(0.2), nn.Linear(512, 10)), 128, 1e-3
    if model == "cifar10_cnn":
        return nn.Sequential(
            nn.Conv2d(3, 32, 3, padding=1), nn.ReLU(), nn.Conv2d(32, 32, 3), nn.ReLU(),
            nn.MaxPool2d(2), nn.Dropout(0.25), nn.Conv2d(32, 64, 3, padding=1), nn.ReLU(),
            nn.Conv2d(64, 64, 3), nn.ReLU(), nn.MaxPool2d(2), nn.Dropout(0.25), nn.Flatten(),
            nn.Linear(64 * 6 * 6, 512), nn.ReLU(), nn.Dropout(0.5), nn.Linear(512, 10)), 32, 1e-4
    return LstmNet(), 128, 0.01


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="mnist_mlp", choices=["mnist_mlp", "cifar10_cnn", "lstm"])
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--num_samples", type=int, default=1000)
    a = ap.parse_args()
    net, batch, lr = build(a.model)
    net = net.to(a.device)
    opt = torch.optim.RMSprop(net.parameters(), lr=lr)
    x, y = synthetic(a.model, a.num_samples, a.device)
    times = []
    for _ in range(a.epochs):
        t0 = time.time()
        for i in range(0, a.num_samples, batch):
            opt.zero_grad()
            loss = nn.functional.cross_entropy(net(x[i:i + batch]), y[i:i + batch])
            loss.backward()
            opt.step()
        if a.device.startswith("cuda"):
            torch.cuda.synchronize()
        times.append(time.time() - t0)
    total = sum(times[1:]) if len(times) > 1 else times[0]   # epoch 1 is warm-up
    print(json.dumps({"test_name": a.model, "total_time": total, "epochs": a.epochs,
                      "batch_size": batch, "backend": "pytorch", "device": a.device,
                      "sample_type": "synthetic", "num_samples": a.num_samples}))


if __name__ == "__main__":
    main()
