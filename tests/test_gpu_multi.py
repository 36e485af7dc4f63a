"""Multi-GPU: fused P2P kernels and the overlapped all-reduce training step. Needs >= 2 GPUs
(skipped on a single-GPU box); ranks are real processes, one per GPU."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _launch(n, args, timeout=600):
    env = dict(os.environ, PYTHONPATH=ROOT)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", str(n), "--"] + args
    return subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_p2p_kernels_match_host_simulation():
    n = min(_ngpu(), 8)
    res = _launch(n, [os.path.join(ROOT, "tools", "p2p_check.py")])
    assert res.returncode == 0 and "P2P_CHECK ALL PASS" in res.stdout, res.stdout[-4000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("mode", [
    ["--variable_update=replicated", "--num_gpus=2"],
    ["--variable_update=kungfu", "--kungfu_option=sync_sgd"],
    ["--variable_update=kungfu", "--kungfu_option=sma"],
    ["--variable_update=kungfu", "--kungfu_option=async_sgd"],
    ["--variable_update=replicated", "--num_gpus=2", "--variable_consistency=relaxed"],
    ["--variable_update=parameter_server", "--num_gpus=2", "--fp16_enable_auto_loss_scale"],
    ["--variable_update=replicated", "--num_gpus=2", "--comm_impl=nccl"],   # comparison baseline
])
def test_two_gpu_training_modes_run(mode):
    args = ["-m", "kf_benchmarks_b200.cli", "--model=resnet20", "--data_name=cifar10",
            "--batch_size=64", "--use_fp16", "--optimizer=momentum", "--num_batches=12",
            "--num_warmup_batches=2", "--display_every=4", "--init_learning_rate=0.05"] + mode
    res = _launch(2, args)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "total images/sec:" in res.stdout
    losses = [float(l.split("\t")[-1]) for l in res.stdout.splitlines()
              if "images/sec:" in l and l[0].isdigit()]
    assert losses and all(x == x and x < 20 for x in losses), res.stdout[-2000:]


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_replicas_stay_bit_identical_and_match_nccl_baseline(tmp_path):
    """Synchronous modes keep every replica's variables bit-identical, and the fused P2P path
    lands on (numerically) the same variables as the NCCL + separate-update baseline."""
    import json
    import numpy as np
    outs = {}
    for impl, compact in (("p2p", False), ("nccl", False), ("p2p", True)):
        key = impl + ("_compact" if compact else "")
        d = tmp_path / key
        d.mkdir()
        kw = dict(model="resnet20", data_name="cifar10", batch_size=32, use_fp16=True,
                  optimizer="momentum", init_learning_rate=0.05, num_batches=6,
                  num_warmup_batches=0, variable_update="replicated", num_gpus=2, comm_impl=impl,
                  compact_gradient_transfer=compact)
        res = _launch(2, [os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw),
                          str(d)])
        assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
        outs[key] = [np.load(str(d / ("rank%d.npz" % r)))["w"] for r in range(2)]
        assert np.array_equal(outs[key][0], outs[key][1])
    a, b = outs["p2p"][0], outs["nccl"][0]
    cos = float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b)))
    assert cos > 0.9999 and np.abs(a - b).max() < 5e-2
    # compact gradient transfer (the default, as in the reference): the gradients cross NVLink as
    # fp16, so six steps at this learning rate drift from the fp32 exchange by fp16 rounding
    c = outs["p2p_compact"][0]
    cos = float((a * c).sum() / (np.linalg.norm(a) * np.linalg.norm(c)))
    assert cos > 0.999 and np.isfinite(c).all()


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("model,data,batch", [("resnet20", "cifar10", 32), ("vgg11", "imagenet", 8)])
def test_fused_wgrad_reduce_scatter_is_bit_identical_to_the_pull_path(tmp_path, model, data, batch):
    """--comm_impl=p2p_fused: the weight-gradient kernels push their tiles to the owning rank
    (tcgen05 epilogue for single-split GEMMs such as the VGG FC layers, the split-K reduction
    otherwise); owners sum their staging slots in rank order, so the trained variables must be
    bit-identical to the pull-based all-reduce."""
    import json
    import numpy as np
    outs = {}
    for impl in ("p2p", "p2p_fused"):
        d = tmp_path / impl
        d.mkdir()
        kw = dict(model=model, data_name=data, batch_size=batch, use_fp16=True,
                  optimizer="momentum", init_learning_rate=1e-4, num_batches=5,
                  num_warmup_batches=0, variable_update="replicated", num_gpus=2, comm_impl=impl,
                  compact_gradient_transfer=False)   # both paths with fp32 on the wire
        res = _launch(2, [os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw),
                          str(d)])
        assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
        outs[impl] = [np.load(str(d / ("rank%d.npz" % r)))["w"] for r in range(2)]
        assert np.isfinite(outs[impl][0]).all()
        assert np.array_equal(outs[impl][0], outs[impl][1])
    assert np.array_equal(outs["p2p"][0], outs["p2p_fused"][0])


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("mode", ["replicated", "replicated_compact", "kungfu_sma"])
def test_nvls_all_reduce_matches_the_pull_kernels(tmp_path, mode):
    """--comm_impl=nvls (multimem.ld_reduce / multimem.st on multicast-bound arenas): replicas
    bit-identical to each other, variables equal to the pull path up to summation order. With
    the compact (fp16) gradient transfer the switch and the pull kernel may round a sum
    differently by one fp16 ulp, which six training steps amplify: the comparison is looser."""
    import json
    import numpy as np
    outs = {}
    kung = mode == "kungfu_sma"
    compact = mode == "replicated_compact"
    for impl in ("p2p", "nvls"):
        d = tmp_path / impl
        d.mkdir()
        kw = dict(model="resnet20", data_name="cifar10", batch_size=32, use_fp16=True,
                  optimizer="momentum", init_learning_rate=0.05, num_batches=6,
                  num_warmup_batches=0, comm_impl=impl, compact_gradient_transfer=compact,
                  variable_update="kungfu" if kung else "replicated",
                  num_gpus=1 if kung else 2)
        if kung:
            kw["kungfu_option"] = "sma"
        res = _launch(2, [os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw),
                          str(d)])
        assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
        outs[impl] = [np.load(str(d / ("rank%d.npz" % r)))["w"] for r in range(2)]
        assert np.isfinite(outs[impl][0]).all()
        if not kung:
            assert np.array_equal(outs[impl][0], outs[impl][1])
    a, b = outs["p2p"][0], outs["nvls"][0]
    cos = float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b)))
    if compact:
        assert cos > 0.999
    else:
        assert cos > 0.9999 and np.abs(a - b).max() < 5e-2


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_multi_rank_cuda_graph_matches_the_eager_step(tmp_path):
    """--cuda_graph on 2 ranks (device-resident sequence offset for the P2P flag protocol):
    same trained variables as the eager step, replicas bit-identical."""
    import json
    import numpy as np
    outs = {}
    for graph in (False, True):
        d = tmp_path / ("graph" if graph else "eager")
        d.mkdir()
        kw = dict(model="resnet20", data_name="cifar10", batch_size=32, use_fp16=True,
                  optimizer="momentum", init_learning_rate=0.05, num_batches=8,
                  num_warmup_batches=0, variable_update="replicated", num_gpus=2,
                  cuda_graph=graph)
        res = _launch(2, [os.path.join(ROOT, "tests", "helpers", "rank_worker.py"), json.dumps(kw),
                          str(d)])
        assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
        outs[graph] = [np.load(str(d / ("rank%d.npz" % r)))["w"] for r in range(2)]
        assert np.isfinite(outs[graph][0]).all()
        assert np.array_equal(outs[graph][0], outs[graph][1])
    a, b = outs[False][0], outs[True][0]
    cos = float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b)))
    assert cos > 0.9999 and np.abs(a - b).max() < 5e-2
