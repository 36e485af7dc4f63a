#!/usr/bin/env python
"""Small-model epoch benchmarks (counterpart of the reference's deprecated second project,
scripts/keras_benchmarks: K/run_benchmark.py, K/models/*_benchmark.py, K/data_generator.py).

Three models on synthetic data, timed per epoch with the first epoch discarded as warm-up
(K/models/timehistory.py:5-14):

* ``mnist_mlp``   784-512-512-10 MLP with dropout 0.2, RMSprop, batch 128
* ``cifar10_cnn`` conv32-conv32-pool-conv64-conv64-pool-FC512-FC10 with dropout, RMSprop(1e-4)
* ``lstm``        LSTM(128) over 40 one-hot steps of 60 symbols + Dense(60), RMSprop(0.01)

The Keras/TF/Theano/CNTK back-ends of the original are replaced by PyTorch (these tiny models
are launch-bound; they are not part of the B200 kernel work).

Like the original (K/run_benchmark.py:19-70) ``--mode cpu_config|gpu_config|multi_gpu_config``
picks a machine description from ``config.json`` next to this file and runs all three models;
``multi_gpu_config`` replicates the model over ``gpus`` devices and splits each batch
(K/gpu_mode.py:7-42, ``keras.utils.multi_gpu_model``). Each model yields one record with the
columns of the original's BigQuery table ``keras_benchmarks.benchmarks``
(K/upload_benchmarks_bq.py:7-80); there is no BigQuery here, so records are printed as JSON
lines and appended to ``--output`` when given. ``null`` machine facts in ``config.json`` are
filled in from the running host.

    python scripts/keras_benchmarks/run_benchmark.py --mode cpu_config
    python scripts/keras_benchmarks/run_benchmark.py --model mnist_mlp --device cpu
"""
import argparse
import json
import os
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


MODELS = ["mnist_mlp", "cifar10_cnn", "lstm"]


def load_config(mode):
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "config.json")) as f:
        cfg = json.load(f)[mode]
    if cfg.get("cpu_num_cores") is None:
        cfg["cpu_num_cores"] = os.cpu_count()
    if cfg.get("cpu_memory") is None:
        try:
            cfg["cpu_memory"] = round(os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES") /
                                      2 ** 30, 2)
        except (ValueError, OSError):
            cfg["cpu_memory"] = 0.0
    return cfg


def run_model(model, device, epochs, num_samples, gpus=0):
    """Train `model` for `epochs` epochs on synthetic data; returns (total_time, batch_size) with
    the first epoch left out as warm-up (K/models/timehistory.py:5-14)."""
    net, batch, lr = build(model)
    net = net.to(device)
    if gpus > 1:
        if torch.cuda.device_count() < gpus:
            raise SystemExit("multi_gpu_config asks for %d GPUs, %d visible" %
                             (gpus, torch.cuda.device_count()))
        net = nn.DataParallel(net, device_ids=list(range(gpus)))
        batch *= gpus                      # the original scales the batch with the GPU count
    opt = torch.optim.RMSprop(net.parameters(), lr=lr)
    x, y = synthetic(model, num_samples, device)
    times = []
    for _ in range(epochs):
        t0 = time.time()
        for i in range(0, num_samples, batch):
            opt.zero_grad()
            loss = nn.functional.cross_entropy(net(x[i:i + batch]), y[i:i + batch])
            loss.backward()
            opt.step()
        if str(device).startswith("cuda"):
            torch.cuda.synchronize()
        times.append(time.time() - t0)
    total = sum(times[1:]) if len(times) > 1 else times[0]   # epoch 1 is warm-up
    return total, batch


def metrics_record(model, total, epochs, batch, cfg, test_id=None):
    """One row of the original's results table (K/upload_benchmarks_bq.py:58-75)."""
    return {"test_id": int(time.time() * 1e3) if test_id is None else test_id,
            "test_name": model, "metrics_totaltime": total, "metrics_epochs": epochs,
            "metrics_batch_size": batch, "metrics_sampletype": "synthetic",
            "keras_backend_type": "pytorch", "keras_backend_version": torch.__version__,
            "keras_version": "n/a",
            "cpu_info_numcores": cfg["cpu_num_cores"], "cpu_info_memory": cfg["cpu_memory"],
            "cpu_info_memory_units": cfg["cpu_memory_info"],
            "platform_info_type": cfg["platform_type"],
            "platform_info_machine_type": cfg["platform_machine_type"],
            "gpu_info_count": cfg["gpus"], "gpu_info_platform": cfg["gpu_platform"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", choices=["cpu_config", "gpu_config", "multi_gpu_config"],
                    help="run all models on the machine described in config.json")
    ap.add_argument("--model", default=None, choices=MODELS)
    ap.add_argument("--device", default=None)
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--num_samples", type=int, default=1000)
    ap.add_argument("--output", default=None, help="append the records to this JSONL file")
    a = ap.parse_args()
    if a.mode:
        cfg = load_config(a.mode)
        device = a.device or ("cuda" if cfg["gpus"] > 0 else "cpu")
        for model in ([a.model] if a.model else MODELS):
            total, batch = run_model(model, device, a.epochs, a.num_samples, gpus=cfg["gpus"])
            rec = metrics_record(model, total, a.epochs, batch, cfg)
            print(json.dumps(rec), flush=True)
            if a.output:
                with open(a.output, "a") as f:
                    f.write(json.dumps(rec) + "\n")
        return
    device = a.device or ("cuda" if torch.cuda.is_available() else "cpu")
    model = a.model or "mnist_mlp"
    total, batch = run_model(model, device, a.epochs, a.num_samples)
    print(json.dumps({"test_name": model, "total_time": total, "epochs": a.epochs,
                      "batch_size": batch, "backend": "pytorch", "device": device,
                      "sample_type": "synthetic", "num_samples": a.num_samples}))


if __name__ == "__main__":
    main()
