"""End-to-end update semantics on CPU, checked against an independent re-simulation.

Follows the reference's strongest test idea (VariableUpdateTest, T/benchmark_cnn_test.py:
1236-1365 with the oracle of T/test_util.py:365-443): run the real driver, then recompute
every step by hand - per-rank gradients from the reference executor, the mode's reduction
rule (replicated sums, parameter_server / KungFu S-SGD average, SMA averages models,
PairAveraging blends with the peer), the TF optimizer formulas - and compare the variables.
Multi-rank cases run as real processes over gloo (ranks sharing this box), like the
reference's spawn-N-local-processes harness (T/benchmark_cnn_distributed_test.py:161-256).
"""
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

from kf_benchmarks_b200 import benchmark, datasets, params as P
from kf_benchmarks_b200.data import inputs
from kf_benchmarks_b200.models import model_config
from kf_benchmarks_b200.nn.executor_ref import RefExecutor
from kf_benchmarks_b200.nn.params import ParameterStore
from kf_benchmarks_b200.ops import optim_ref
from kf_benchmarks_b200.parallel.strategy import UpdateEngine
from kf_benchmarks_b200.parallel.comm import SingleComm
from kf_benchmarks_b200.utils import log_util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = dict(model="trivial", data_name="cifar10", device="cpu", data_format="NHWC", batch_size=4,
            num_batches=3, num_warmup_batches=0, display_every=1, weight_decay=1e-3,
            init_learning_rate=0.05)


class _Sim:
    """Independent simulation of `world` replicas."""

    def __init__(self, kw, world):
        self.kw, self.world = kw, world
        self.p = P.make_params(**kw)
        ds = datasets.create_dataset(None, "cifar10")
        self.model = model_config.get_model_config(self.p.model, ds, self.p)
        self.model.set_batch_size(self.p.batch_size)
        self.g = self.model.build_graph(ds.num_classes, batch_size=self.p.batch_size)
        per_rank_seed = self.p.variable_update in ("kungfu", "horovod")
        self.stores, self.srcs, self.slots = [], [], []
        for r in range(world):
            st = ParameterStore(self.g, "cpu", fp16_shadow=False)
            st.initialize(self.p.tf_random_seed + (r if per_rank_seed else 0))
            self.stores.append(st)
            seed = self.p.tf_random_seed + (r if per_rank_seed else 0)
            self.srcs.append(inputs.SyntheticHostSource(
                self.p.batch_size, 32, 3, ds.num_classes, "cpu", seed * 1000003 + r,
                dtype=torch.float32))
            rms_init = 1.0 if self.p.optimizer == "rmsprop" else 0.0      # TF: rms slot = ones
            self.slots.append((torch.zeros_like(st.w32), torch.full_like(st.w32, rms_init)))
        # every mode starts from rank 0's variables (broadcast at start)
        for st in self.stores[1:]:
            st.w32.copy_(self.stores[0].w32)
            st.state.copy_(self.stores[0].state)
        self.engine = UpdateEngine(self.p, self.stores[0], SingleComm(), self.model)
        self.nclass = ds.num_classes

    def grads(self, r):
        ex = RefExecutor(self.g, self.stores[r], self.nclass)
        ex.set_inputs(*self.srcs[r].next_batch())
        ex.forward(True)
        ex.loss_backward(1.0)
        return self.stores[r].grad.clone()

    def step(self, k):
        p, n = self.p, self.world
        h = self.engine._hyper(p.init_learning_rate)
        h.adam_t = float(k + 1)
        h.grad_scale = 1.0
        # the reference puts num_towers * wd * l2 on one tower: summed gradients carry N * wd * w
        h.weight_decay = p.weight_decay * (n if p.variable_update in ("replicated", "horovod")
                                           else 1)
        gs = [self.grads(r) for r in range(n)]
        vu, opt = p.variable_update, p.kungfu_option
        if vu == "independent" or n == 1:
            red = gs
        elif vu in ("replicated", "horovod"):
            red = [sum(gs)] * n
        elif vu == "parameter_server" or (vu == "kungfu" and opt == "sync_sgd"):
            red = [sum(gs) / n] * n
        else:
            red = gs
        if vu == "kungfu" and opt == "sma" and n > 1:
            avg = sum(st.w32 for st in self.stores) / n
            for st in self.stores:
                st.w32.mul_(1 - p.sma_alpha).add_(avg, alpha=p.sma_alpha)
        if vu == "kungfu" and opt == "async_sgd" and n == 2:
            w0, w1 = self.stores[0].w32.clone(), self.stores[1].w32.clone()
            self.stores[0].w32.copy_(0.5 * (w0 + w1))
            self.stores[1].w32.copy_(0.5 * (w0 + w1))
        for r in range(n):
            m, v = self.slots[r]
            optim_ref.apply_update(h, self.stores[r].w32, red[r], m, v)


def _run_single(kw):
    lines = []
    old = log_util.log_fn
    log_util.log_fn = lines.append
    try:
        bench = benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw)))
        bench.print_info()
        stats = bench.run()
    finally:
        log_util.log_fn = old
    return bench, stats, lines


@pytest.mark.parametrize("extra", [
    dict(optimizer="sgd"), dict(optimizer="momentum"), dict(optimizer="rmsprop"),
    dict(optimizer="adam"), dict(optimizer="momentum", gradient_clip=1e-3),
    dict(optimizer="sgd", variable_update="replicated"),
    dict(optimizer="sgd", variable_update="independent"),
])
def test_single_process_matches_simulation(extra):
    kw = dict(BASE)
    kw.update(extra)
    bench, stats, lines = _run_single(kw)
    sim = _Sim(kw, 1)
    for k in range(kw["num_batches"]):
        sim.step(k)
    assert torch.allclose(bench.store.w32, sim.stores[0].w32, rtol=1e-5, atol=1e-6)
    # log format: header, 3 step lines, totals
    steps = [l for l in lines if re.match(r"^\d+\timages/sec: ", l)]
    assert len(steps) == 3 and any(l.startswith("total images/sec: ") for l in lines)
    assert "Step\tImg/sec\ttotal_loss" in lines and "Model:       trivial" in lines
    assert stats["num_steps"] == 3 and stats["images_per_sec"] > 0


def _spawn(kw, world, tmp_path, launcher_np=None):
    out = str(tmp_path)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", str(world), "--",
           os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw), out]
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    env.pop("RANK", None)
    env.pop("WORLD_SIZE", None)
    for attempt in range(2):       # one retry: the rendezvous port is picked, released, re-bound
        res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
        if res.returncode == 0:
            break
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return [np.load(os.path.join(out, "rank%d.npz" % r)) for r in range(world)]


@pytest.mark.parametrize("extra", [
    dict(variable_update="replicated", num_gpus=2, optimizer="momentum"),
    dict(variable_update="parameter_server", num_gpus=2, optimizer="sgd"),
    dict(variable_update="independent", num_gpus=2, optimizer="sgd"),
    dict(variable_update="kungfu", kungfu_option="sync_sgd", optimizer="momentum"),
    dict(variable_update="kungfu", kungfu_option="sma", optimizer="sgd"),
    dict(variable_update="kungfu", kungfu_option="async_sgd", optimizer="sgd"),
    dict(variable_update="horovod", optimizer="sgd"),
])
def test_two_ranks_match_simulation(extra, tmp_path):
    kw = dict(BASE)
    kw.update(extra)
    results = _spawn(kw, 2, tmp_path)
    sim = _Sim(kw, 2)
    for k in range(kw["num_batches"]):
        sim.step(k)
    for r in range(2):
        assert np.allclose(results[r]["w"], sim.stores[r].w32.numpy(), rtol=2e-5, atol=2e-6), \
            "rank %d diverges from the simulation" % r
    synced = extra["variable_update"] in ("replicated", "parameter_server", "horovod") or \
        extra.get("kungfu_option") == "sync_sgd"
    same = np.array_equal(results[0]["w"], results[1]["w"])
    assert same == synced      # synchronous modes keep replicas bit-identical, others do not


def test_tf_cluster_flags_start_workers_without_a_launcher(tmp_path):
    """Reference-style launch (T/benchmark_cnn.py:1354-1406): one process per worker with
    --job_name=worker --task_index=i --worker_hosts=...; no torchrun / launcher environment."""
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    hosts = "127.0.0.1:%d,127.0.0.1:%d" % (port, port + 1)
    out = str(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    procs = []
    for i in range(2):
        kw = dict(BASE, variable_update="distributed_replicated", optimizer="sgd",
                  job_name="worker", task_index=i, worker_hosts=hosts, ps_hosts="127.0.0.1:1")
        procs.append(subprocess.Popen(
            [sys.executable, os.path.join(ROOT, "tests", "helpers", "rank_worker.py"),
             json.dumps(kw), out], cwd=ROOT, env=env, stdout=subprocess.PIPE,
            stderr=subprocess.STDOUT, text=True))
    for pr in procs:
        o, _ = pr.communicate(timeout=600)
        assert pr.returncode == 0, o[-3000:]
    w0 = np.load(os.path.join(out, "rank0.npz"))["w"]
    w1 = np.load(os.path.join(out, "rank1.npz"))["w"]
    assert np.array_equal(w0, w1)          # averaged-gradient sync keeps the replicas identical
    sim = _Sim(dict(BASE, variable_update="parameter_server", num_gpus=2, optimizer="sgd"), 2)
    for k in range(BASE["num_batches"]):
        sim.step(k)
    assert np.allclose(w0, sim.stores[0].w32.numpy(), rtol=2e-5, atol=2e-6)


def test_resume_from_checkpoint(tmp_path):
    kw = dict(BASE, optimizer="momentum", train_dir=str(tmp_path / "ckpt"), num_batches=2)
    bench, _, lines = _run_single(kw)
    # the reference's testSaveLoadModel looks for exactly this file (T/benchmark_cnn_test.py:74-142)
    assert os.path.exists(os.path.join(kw["train_dir"], "model.ckpt-2.index"))
    # --num_batches is the total from global step 0 (T/benchmark_cnn.py:2242-2250): resuming
    # with the same value has nothing left to do, with a larger one runs the difference
    bench2, stats2, lines2 = _run_single(kw)
    assert any("Restored checkpoint" in l for l in lines2)
    assert bench2.global_step == 2 and stats2["num_steps"] == 0
    bench3, stats3, _ = _run_single(dict(kw, num_batches=5))
    assert bench3.global_step == 5 and stats3["num_steps"] == 3
    assert os.path.exists(os.path.join(kw["train_dir"], "model.ckpt-5.index"))
    assert os.path.exists(os.path.join(kw["train_dir"], "model.ckpt-5.data-00000-of-00001"))


def test_eval_and_eval_during_training(tmp_path):
    kw = dict(BASE, train_dir=str(tmp_path / "ckpt"), num_batches=2,
              eval_during_training_every_n_steps=1, num_eval_batches=2)
    _, _, lines = _run_single(kw)
    acc = [l for l in lines if l.startswith("Accuracy @ 1 = ")]
    assert len(acc) == 2 and re.match(
        r"^Accuracy @ 1 = \d\.\d{4} Accuracy @ 5 = \d\.\d{4} \[8 examples\]$", acc[0])
    kw = dict(BASE, train_dir=str(tmp_path / "ckpt"), eval=True, num_batches=3,
              eval_dir=str(tmp_path / "eval"))
    kw.pop("init_learning_rate")
    _, res, lines = _run_single(kw)
    assert res["num_examples"] == 12 and 0.0 <= res["top_1_accuracy"] <= res["top_5_accuracy"]
    # --eval_dir receives the scalars of the reference's eval summary
    recs = [json.loads(l) for l in open(str(tmp_path / "eval" / "summaries.jsonl"))]
    assert len(recs) == 1 and recs[0]["step"] == 2
    assert recs[0]["eval/Accuracy@1"] == res["top_1_accuracy"]
    assert recs[0]["eval/Accuracy@5"] == res["top_5_accuracy"]


def test_forward_only_and_loss_reporting():
    _, stats, lines = _run_single(dict(BASE, forward_only=True))
    assert any(l.endswith("top_1_accuracy\ttop_5_accuracy") for l in lines)
    b1, s1, _ = _run_single(dict(BASE, loss_type_to_report="base_loss", num_batches=1))
    b2, s2, _ = _run_single(dict(BASE, loss_type_to_report="total_loss", num_batches=1))
    l2 = 0.5 * float((b2.store.w32.double() ** 2).sum()) * BASE["weight_decay"]
    assert s2["last_average_loss"] == pytest.approx(s1["last_average_loss"] + l2, rel=1e-4)


def test_relaxed_consistency_applies_gradients_one_step_late(tmp_path):
    """variable_consistency=relaxed (T/benchmark_cnn_test.py:1368-1476: the result lags one
    step): after K steps the variables equal a strong run's after K-1 updates computed from
    gradients taken at the stale weights - checked with an explicit re-simulation."""
    kw = dict(BASE, variable_update="replicated", num_gpus=2, optimizer="sgd",
              variable_consistency="relaxed", num_batches=3)
    results = _spawn(kw, 2, tmp_path)
    sim = _Sim(dict(kw, variable_consistency="strong"), 2)
    pending = None
    for k in range(3):
        h = sim.engine._hyper(kw["init_learning_rate"])
        h.grad_scale = 1.0
        h.weight_decay = kw["weight_decay"] * 2        # summed gradients carry N * wd * w
        gs = [sim.grads(r) for r in range(2)]
        if pending is not None:
            for r in range(2):
                optim_ref.apply_update(h, sim.stores[r].w32, pending, None, None)
        pending = sum(gs)
    for r in range(2):
        assert np.allclose(results[r]["w"], sim.stores[r].w32.numpy(), rtol=2e-5, atol=2e-6)


def _numpy_trivial_losses(kw, steps):
    """Independent float64 NumPy re-simulation of the driver on the trivial CIFAR model
    (flatten -> FC1+ReLU -> FC4096+ReLU -> FC11, softmax cross-entropy + wd * l2): the printed
    loss of every step. Counterpart of the reference's `manually_compute_losses`
    (T/test_util.py:365-443, T/benchmark_cnn_test.py:1236-1365); nothing of the framework's
    executors or update kernels is used - only the initial variables and the input batches."""
    p = P.make_params(**kw)
    ds = datasets.create_dataset(None, "cifar10")
    model = model_config.get_model_config(p.model, ds, p)
    g = model.build_graph(ds.num_classes, batch_size=p.batch_size)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    st.initialize(p.tf_random_seed)
    tf = st.to_tf_arrays()
    names = ["v0/cg/affine%d/%s" % (i, k) for i in range(3) for k in ("weights", "biases")]
    w = [tf[n].astype(np.float64) for n in names]          # [in,out] weights, [out] biases
    slots = [(np.zeros_like(x), np.full_like(x, 1.0 if p.optimizer == "rmsprop" else 0.0))
             for x in w]
    src = inputs.SyntheticHostSource(p.batch_size, 32, 3, ds.num_classes, "cpu",
                                     p.tf_random_seed * 1000003, dtype=torch.float32)
    lr, wd = p.init_learning_rate, p.weight_decay or 0.0
    losses = []
    for k in range(steps):
        images, labels = src.next_batch()
        x0 = images.numpy().astype(np.float64).reshape(p.batch_size, -1)
        y = labels.numpy().astype(np.int64)
        z1 = x0 @ w[0] + w[1]
        a1 = np.maximum(z1, 0)
        z2 = a1 @ w[2] + w[3]
        a2 = np.maximum(z2, 0)
        z3 = a2 @ w[4] + w[5]
        z3s = z3 - z3.max(1, keepdims=True)
        logp = z3s - np.log(np.exp(z3s).sum(1, keepdims=True))
        base = -logp[np.arange(len(y)), y].mean()
        losses.append(base + wd * sum(0.5 * (v ** 2).sum() for v in w))
        d3 = np.exp(logp)
        d3[np.arange(len(y)), y] -= 1
        d3 /= len(y)
        grads = [None] * 6
        grads[4], grads[5] = a2.T @ d3, d3.sum(0)
        d2 = (d3 @ w[4].T) * (z2 > 0)
        grads[2], grads[3] = a1.T @ d2, d2.sum(0)
        d1 = (d2 @ w[2].T) * (z1 > 0)
        grads[0], grads[1] = x0.T @ d1, d1.sum(0)
        for i in range(6):
            gi = grads[i] + wd * w[i]
            if p.gradient_clip:
                gi = np.clip(gi, -p.gradient_clip, p.gradient_clip)
            m, v = slots[i]
            if p.optimizer == "sgd":
                w[i] = w[i] - lr * gi
            elif p.optimizer == "momentum":                     # ApplyMomentum, use_nesterov
                m[:] = p.momentum * m + gi
                w[i] = w[i] - lr * (gi + p.momentum * m)
            elif p.optimizer == "rmsprop":
                v[:] = p.rmsprop_decay * v + (1 - p.rmsprop_decay) * gi * gi
                m[:] = p.rmsprop_momentum * m + lr * gi / np.sqrt(v + p.rmsprop_epsilon)
                w[i] = w[i] - m
            else:                                               # adam
                b1, b2, t = p.adam_beta1, p.adam_beta2, k + 1
                m[:] = b1 * m + (1 - b1) * gi
                v[:] = b2 * v + (1 - b2) * gi * gi
                w[i] = w[i] - lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t) * m / \
                    (np.sqrt(v) + p.adam_epsilon)
    return losses


@pytest.mark.parametrize("extra", [
    dict(optimizer="sgd"), dict(optimizer="momentum"), dict(optimizer="rmsprop"),
    dict(optimizer="adam"), dict(optimizer="momentum", gradient_clip=0.01),
    dict(optimizer="sgd", weight_decay=0.0),
])
def test_printed_losses_match_an_independent_numpy_simulation(extra):
    kw = dict(BASE, num_batches=5, init_learning_rate=1e-7, **extra)
    kw.update(extra)
    _, _, lines = _run_single(kw)
    printed = [float(l.split("\t")[-1]) for l in lines if re.match(r"^\d+\timages/sec: ", l)]
    want = _numpy_trivial_losses(kw, 5)
    assert len(printed) == 5
    assert np.allclose(printed, want, rtol=2e-4, atol=2e-3), (printed, want)
    assert printed[0] != printed[-1]                    # the variables really move


def test_early_stop_and_eval_schedule_with_mocked_accuracies(monkeypatch):
    """T/benchmark_cnn_test.py:1005-1169: evaluations happen exactly at the requested steps
    and training stops at the first evaluation whose top-1 reaches --stop_at_top_1_accuracy."""
    accs = iter([0.1, 0.4, 0.95, 0.99])
    calls = []

    def fake_eval(self):
        calls.append(self.global_step)
        a = next(accs)
        return {"top_1_accuracy": a, "top_5_accuracy": a, "num_examples": 4,
                "eval_average_loss": 0.0, "images_per_sec": 1.0}
    monkeypatch.setattr(benchmark.BenchmarkCNN, "_eval_during_training", fake_eval)
    kw = dict(BASE, num_batches=10, eval_during_training_at_specified_steps=["2", "3", "5", "8"],
              stop_at_top_1_accuracy=0.9)
    bench, stats, lines = _run_single(kw)
    assert calls == [2, 3, 5]                       # stopped after the third evaluation
    assert stats["num_steps"] == 5
    assert any("top_1_accuracy" in l and "0.95" in l or "Stopping" in l or "stop" in l.lower()
               for l in lines)
    # every-n-steps schedule without early stop evaluates floor(10 / 4) times
    accs2 = iter([0.0] * 10)
    calls.clear()
    monkeypatch.setattr(benchmark.BenchmarkCNN, "_eval_during_training",
                        lambda self: (calls.append(self.global_step), {
                            "top_1_accuracy": next(accs2), "top_5_accuracy": 0.0, "num_examples": 4,
                            "eval_average_loss": 0.0, "images_per_sec": 1.0})[1])
    _, stats, _ = _run_single(dict(BASE, num_batches=10, eval_during_training_every_n_steps=4))
    assert calls == [4, 8] and stats["num_steps"] == 10


@pytest.mark.parametrize("opt", ["momentum", "rmsprop", "adam"])
def test_resume_continues_exactly_including_optimizer_slots(tmp_path, opt):
    """2 steps + resume for 2 more == 4 steps in one run (variables and slots are restored)."""
    base = dict(BASE, optimizer=opt, init_learning_rate=1e-7)
    one, _, _ = _run_single(dict(base, num_batches=4, train_dir=str(tmp_path / "a")))
    _run_single(dict(base, num_batches=2, train_dir=str(tmp_path / "b")))
    # the synthetic source alternates two batches: after 2 steps the resumed run starts on the
    # same batch the uninterrupted run would use for step 3
    two, _, lines = _run_single(dict(base, num_batches=4, train_dir=str(tmp_path / "b")))
    assert any("Restored checkpoint" in l for l in lines)
    assert two.global_step == one.global_step == 4
    assert torch.allclose(one.store.w32, two.store.w32, rtol=1e-6, atol=1e-9)
    from kf_benchmarks_b200 import ckpt
    a = ckpt.load_arrays(ckpt.latest_checkpoint(str(tmp_path / "a"))[0])
    b = ckpt.load_arrays(ckpt.latest_checkpoint(str(tmp_path / "b"))[0])
    assert set(a) == set(b) and any("/" + {"momentum": "Momentum", "rmsprop": "RMSProp",
                                           "adam": "Adam_1"}[opt] in k for k in a)
    for k in a:           # variables and optimizer slots, logical (checkpoint) layout
        assert np.allclose(a[k], b[k], rtol=1e-5, atol=1e-12), k


def test_parameter_server_checkpoints_use_the_v_scope(tmp_path):
    """T/variable_mgr.py:212-214: one shared copy `v/...` with a parameter server, `v0/...`
    (tower 0) otherwise; a checkpoint of one mode restores into the other."""
    from kf_benchmarks_b200 import ckpt
    d = str(tmp_path / "ps")
    ps, _, _ = _run_single(dict(BASE, variable_update="parameter_server", optimizer="momentum",
                                num_batches=2, train_dir=d))
    arrays = ckpt.load_arrays(ckpt.latest_checkpoint(d)[0])
    assert "v/cg/affine0/weights" in arrays and "v/cg/affine0/weights/Momentum" in arrays
    assert not any(k.startswith("v0/") for k in arrays)
    rep, _, lines = _run_single(dict(BASE, variable_update="replicated", optimizer="momentum",
                                     num_batches=2, train_dir=d))
    assert any("Restored checkpoint" in l for l in lines)
    assert torch.equal(rep.store.w32, ps.store.w32)        # nothing left to run: same variables
    arrays = ckpt.load_arrays(ckpt.latest_checkpoint(d)[0])
    assert "v0/cg/affine0/weights" in arrays


@pytest.mark.parametrize("extra", [
    dict(all_reduce_spec="pscpu:2k:xring", agg_small_grads_max_bytes=1024,
         agg_small_grads_max_group=10),
    dict(all_reduce_spec="nccl", compact_gradient_transfer=False, gradient_repacking=2),
    dict(hierarchical_copy=True, gradient_repacking=3),
])
def test_gradient_aggregation_matrix(extra, tmp_path):
    """Counterpart of VariableMgrLocalReplicatedTest (T/benchmark_cnn_test.py:1368-1476): whatever
    the packing / algorithm flags say about HOW gradients travel (repacking, small-tensor
    aggregation, compact transfer, hierarchical copy, all_reduce_spec), the applied gradient of
    every variable is the plain sum over replicas."""
    kw = dict(BASE)
    kw.update(variable_update="replicated", num_gpus=2, optimizer="sgd")
    kw.update(extra)
    results = _spawn(kw, 2, tmp_path)
    sim = _Sim(dict(BASE, variable_update="replicated", num_gpus=2, optimizer="sgd"), 2)
    for k in range(kw["num_batches"]):
        sim.step(k)
    for r in range(2):
        assert np.allclose(results[r]["w"], sim.stores[r].w32.numpy(), rtol=2e-5, atol=2e-6)
    assert np.array_equal(results[0]["w"], results[1]["w"])


def test_model_can_keep_variables_out_of_the_l2_term():
    """The SSD model decays no batch-norm variable (T/models/ssd_model.py:91-94). The engine
    turns the model's ``l2_decay_filter`` into a per-128-element-block mask that the update
    (CPU restatement here, the CUDA kernels through ``kfb_hyper.decay_mask``) multiplies the
    weight-decay coefficient with, and total_loss leaves the same variables out."""
    import torch
    from kf_benchmarks_b200 import datasets, params as params_lib
    from kf_benchmarks_b200.models import model_config
    from kf_benchmarks_b200.nn.params import ParameterStore
    from kf_benchmarks_b200.parallel.comm import SingleComm
    from kf_benchmarks_b200.parallel.strategy import UpdateEngine
    p = params_lib.make_params(model="resnet20", data_name="cifar10", device="cpu",
                               optimizer="sgd", weight_decay=0.1, variable_update="independent")
    ds = datasets.create_dataset(None, "cifar10")
    model = model_config.get_model_config("resnet20", ds, p)
    g = model.build_graph(ds.num_classes, batch_size=2)
    engines = {}
    for name, flt in (("all", None), ("no_bn", lambda v: "batchnorm" not in v)):
        store = ParameterStore(g, "cpu", fp16_shadow=False)
        store.initialize(7)
        if flt is not None:
            model.l2_decay_filter = flt
        eng = UpdateEngine(p, store, SingleComm(), model)
        store.grad.zero_()                       # only the L2 term moves the weights
        eng.begin_step(1.0)
        w0 = store.w32.clone()
        eng.finish_step()
        engines[name] = (store, w0, eng)
    store, w0, eng = engines["all"]
    assert eng.decay_mask is None
    assert torch.allclose(store.w32, w0 * 0.9)
    store, w0, eng = engines["no_bn"]
    assert eng.decay_mask is not None and int(eng.decay_mask.sum()) < eng.decay_mask.numel()
    for slot in store.order:
        a = store.view(slot.spec.name)
        b = w0[slot.offset:slot.offset + slot.numel].view_as(a)
        if "batchnorm" in slot.spec.name:
            assert torch.equal(a, b), slot.spec.name
        else:
            assert torch.allclose(a, b * 0.9), slot.spec.name
    from kf_benchmarks_b200.models import ssd_model
    assert ssd_model.SSD300Model.l2_decay_filter(None, "ssd300/conv3/batchnorm3/gamma") is False
    assert ssd_model.SSD300Model.l2_decay_filter(None, "ssd300/conv3/conv2d/kernel") is True
