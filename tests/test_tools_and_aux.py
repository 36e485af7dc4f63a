"""all_reduce_benchmark, trace / profile outputs, benchmark logger (reference tests:
T/all_reduce_benchmark_test.py:28-47, T/benchmark_cnn_test.py:656-675)."""
import json
import os
import re
import subprocess
import sys

from kf_benchmarks_b200 import benchmark, params as P
from kf_benchmarks_b200.utils import log_util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_all_reduce_benchmark_two_cpu_ranks():
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", "2", "--", "-m",
           "kf_benchmarks_b200.all_reduce_benchmark", "--model=trivial", "--data_name=cifar10",
           "--device=cpu", "--data_format=NHWC", "--num_batches=2", "--num_warmup_batches=1",
           "--iters_per_step=2"]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = [l for l in res.stdout.splitlines() if l.startswith("Average time per step:")]
    assert lines and re.match(r"^Average time per step: [0-9.e-]+$", lines[-1])


def test_trace_profile_and_benchmark_log(tmp_path):
    kw = dict(model="trivial", data_name="cifar10", device="cpu", data_format="NHWC", batch_size=4,
              num_batches=11, num_warmup_batches=3, trace_file=str(tmp_path / "trace.json"),
              tfprof_file=str(tmp_path / "prof.txt"), benchmark_log_dir=str(tmp_path / "blog"),
              benchmark_test_id="t1")
    lines = []
    old = log_util.log_fn
    log_util.log_fn = lines.append
    try:
        benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw))).run()
    finally:
        log_util.log_fn = old
    assert any(l.startswith("Dumping trace to") for l in lines)
    with open(kw["trace_file"]) as f:
        trace = json.load(f)
    assert "traceEvents" in trace and len(trace["traceEvents"]) > 0
    assert os.path.getsize(kw["tfprof_file"]) > 0
    with open(os.path.join(kw["benchmark_log_dir"], "benchmark_run.log")) as f:
        info = json.load(f)
    assert info["model_name"] == "trivial" and info["test_id"] == "t1"
    with open(os.path.join(kw["benchmark_log_dir"], "metric.log")) as f:
        metrics = [json.loads(l) for l in f]
    assert {m["name"] for m in metrics} >= {"average_examples_per_sec", "last_average_loss",
                                            "current_examples_per_sec"}
    cur = [m for m in metrics if m["name"] == "current_examples_per_sec"]
    assert [m["global_step"] for m in cur] == [1, 10] and cur[0]["unit"] == "examples/sec"


def test_cli_entry_point_and_help():
    env = dict(os.environ, PYTHONPATH=ROOT)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    res = subprocess.run([sys.executable, "-m", "kf_benchmarks_b200.cli", "--model=trivial",
                          "--device=cpu", "--data_format=NHWC", "--data_name=cifar10",
                          "--batch_size=4", "--num_batches=2", "--num_warmup_batches=0",
                          "--display_every=1"], cwd=ROOT, env=env, capture_output=True, text=True,
                         timeout=300)
    assert res.returncode == 0, res.stderr
    assert re.search(r"^total images/sec: \d+\.\d\d$", res.stdout, re.M)
    res = subprocess.run([sys.executable, "-m", "kf_benchmarks_b200.cli", "--model", "trivial",
                          "stray"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=60)
    assert res.returncode == 2 and "positional" in res.stderr


def test_summaries_jsonl(tmp_path):
    kw = dict(model="trivial", data_name="cifar10", device="cpu", data_format="NHWC", batch_size=4,
              num_batches=4, num_warmup_batches=0, train_dir=str(tmp_path), summary_verbosity=3,
              save_summaries_steps=2)
    old = log_util.log_fn
    log_util.log_fn = lambda s: None
    try:
        benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw))).run()
    finally:
        log_util.log_fn = old
    recs = [json.loads(l) for l in open(os.path.join(str(tmp_path), "summaries.jsonl"))]
    assert [r["step"] for r in recs] == [2, 4]
    assert "learning_rate" in recs[0] and "log_gradients" in recs[0]
    assert "cg/affine0/weights" in recs[0]["variables"]
    # the same summaries as a TensorBoard event file (what tf.summary.FileWriter leaves behind)
    from kf_benchmarks_b200.utils import tfevents
    path = tfevents.latest_event_file(str(tmp_path))
    assert os.path.basename(path).startswith("events.out.tfevents.")
    events = list(tfevents.read_events(path))                  # CRC-checked TFRecord framing
    assert events[0]["file_version"] == "brain.Event:2"
    assert [e["step"] for e in events[1:]] == [2, 4]
    e = events[1]
    assert abs(e["scalars"]["learning_rate"] - recs[0]["learning_rate"]) < 1e-6
    assert abs(e["scalars"]["total_loss"] - recs[0]["loss"]) < 1e-3 * max(1.0, abs(recs[0]["loss"]))
    h = e["histograms"]["cg/affine0/weights"]
    assert h["num"] > 0 and len(h["bucket"]) == len(h["bucket_limit"]) == 30
    assert abs(sum(h["bucket"]) - h["num"]) < 1e-6 and h["min"] <= h["max"]
    assert "cg/affine0/weights/gradients" in e["histograms"] and "log_gradients" in e["histograms"]


def test_tfevents_encoding_against_hand_assembled_bytes():
    """The Event encoder against bytes assembled by hand from the protobuf wire format."""
    import struct
    from kf_benchmarks_b200.utils import tfevents
    got = tfevents.encode_event(3, 1.5, scalars={"a": 0.25})
    value = b"\x0a\x01a" + b"\x15" + struct.pack("<f", 0.25)          # tag = 1, simple_value = 2
    summary = b"\x0a" + bytes([len(value)]) + value                     # Summary.value = 1
    want = b"\x09" + struct.pack("<d", 1.5) + b"\x10\x03" + b"\x2a" + bytes([len(summary)]) + summary
    assert got == want
    assert tfevents.parse_event(got) == {"wall_time": 1.5, "step": 3, "scalars": {"a": 0.25},
                                         "histograms": {}}
    h = tfevents.histogram_of([1.0, 2.0, 2.0, 4.0], bins=3)
    assert h["num"] == 4 and h["sum"] == 9 and h["sum_squares"] == 25 and h["bucket"] == [1, 2, 1]
    back = tfevents.parse_event(tfevents.encode_event(-1, 0.0, histograms={"h": h}))
    assert back["step"] == -1 and back["histograms"]["h"] == h


def test_launcher_kills_group_when_a_rank_dies(tmp_path):
    """One rank exits at step 1: the launcher reports it, stops the survivor (which would
    otherwise wait forever in the all-reduce) and returns the failing exit code."""
    import time
    env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS="2", KFB_FAULT_INJECT="1:1:exit")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    cmd = [sys.executable, "-m", "kf_benchmarks_b200.cli", "--model=trivial", "--data_name=cifar10",
           "--device=cpu", "--data_format=NHWC", "--batch_size=4", "--num_batches=50",
           "--num_warmup_batches=0", "--num_gpus=2", "--variable_update=replicated"]
    t0 = time.time()
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=240)
    assert res.returncode == 17, (res.returncode, res.stderr[-500:])
    assert "rank 1 exited with 17" in res.stderr
    assert time.time() - t0 < 200


def test_export_tool_lists_reference_names(tmp_path):
    kw = dict(model="trivial", data_name="cifar10", device="cpu", data_format="NHWC", batch_size=4,
              num_batches=1, num_warmup_batches=0, train_dir=str(tmp_path), optimizer="momentum",
              checkpoint_format="npz")
    old = log_util.log_fn
    log_util.log_fn = lambda s: None
    try:
        benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw))).run()
    finally:
        log_util.log_fn = old
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "export_tf_checkpoint.py"),
                          "--list", os.path.join(str(tmp_path), "model.ckpt-1.npz")],
                         capture_output=True, text=True, timeout=60)
    assert res.returncode == 0
    assert "v/cg/affine0/weights" in res.stdout and "v/cg/affine0/weights/Momentum" in res.stdout  # default variable_update=parameter_server -> outer scope v/
    assert "global_step" in res.stdout
    # npz -> TensorFlow bundle -> resume training from the bundle
    out = tmp_path / "tf" / "model.ckpt-1"
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "export_tf_checkpoint.py"),
                          os.path.join(str(tmp_path), "model.ckpt-1.npz"), str(out)],
                         capture_output=True, text=True, timeout=60)
    assert res.returncode == 0 and os.path.exists(str(out) + ".index"), res.stderr
    with open(str(tmp_path / "tf" / "checkpoint"), "w") as f:
        f.write('model_checkpoint_path: "model.ckpt-1"\nall_model_checkpoint_paths: '
                '"model.ckpt-1"\n')
    lines = []
    log_util.log_fn = lines.append
    try:
        kw2 = dict(kw, train_dir=str(tmp_path / "tf"), num_batches=2, checkpoint_format="tf")
        benchmark.BenchmarkCNN(benchmark.setup(P.make_params(**kw2))).run()
    finally:
        log_util.log_fn = old
    assert any("Restored checkpoint" in l and "global_step 1" in l for l in lines)
    assert os.path.exists(str(tmp_path / "tf" / "model.ckpt-2.index"))


def test_platform_hooks_and_cluster_manager():
    from kf_benchmarks_b200 import params as P
    from kf_benchmarks_b200.platforms import util as platforms_util
    assert os.path.isdir(platforms_util.get_test_output_dir())
    cmd = platforms_util.get_command_to_run_python_module("cli")
    assert cmd[-1] == "kf_benchmarks_b200.cli"
    p = P.make_params(job_name="worker", task_index=1, worker_hosts="a:1,b:2", ps_hosts="c:3")
    cm = platforms_util.get_cluster_manager(p)
    assert cm.num_workers() == 2 and cm.num_ps() == 1 and cm.get_target() == "a:1"
    assert cm.get_cluster_spec() == {"worker": ["a:1", "b:2"], "ps": ["c:3"]}
    assert cm.join_server() is None
    platforms_util.initialize(p)


def test_keras_style_epoch_benchmark_runs():
    import json
    res = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "keras_benchmarks",
                                                       "run_benchmark.py"), "--model", "mnist_mlp",
                          "--device", "cpu", "--epochs", "2", "--num_samples", "256"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    d = json.loads(res.stdout.strip().splitlines()[-1])
    assert d["test_name"] == "mnist_mlp" and d["epochs"] == 2 and d["total_time"] > 0


def test_keras_style_benchmark_mode_writes_the_reference_record_schema(tmp_path):
    """--mode runs all three models and emits the columns of the original's results table
    (K/upload_benchmarks_bq.py:58-75) to a JSONL sink."""
    import json
    out = str(tmp_path / "records.jsonl")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "keras_benchmarks",
                                                       "run_benchmark.py"), "--mode", "cpu_config",
                          "--num_samples", "128", "--output", out],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    recs = [json.loads(l) for l in open(out)]
    assert [r["test_name"] for r in recs] == ["mnist_mlp", "cifar10_cnn", "lstm"]
    cols = {"test_id", "test_name", "metrics_totaltime", "metrics_epochs", "metrics_batch_size",
            "metrics_sampletype", "keras_backend_type", "keras_backend_version", "keras_version",
            "cpu_info_numcores", "cpu_info_memory", "cpu_info_memory_units", "platform_info_type",
            "platform_info_machine_type", "gpu_info_count", "gpu_info_platform"}
    for r in recs:
        assert set(r) == cols and r["metrics_totaltime"] > 0 and r["gpu_info_count"] == 0
        assert r["cpu_info_numcores"] >= 1 and r["metrics_sampletype"] == "synthetic"


def test_fake_imagenet_generator_and_pipeline_bench(tmp_path):
    """tools/make_fake_imagenet.py (counterpart of T/test_data/tfrecord_image_generator.py) and
    the host-side input benchmark built on it."""
    d = str(tmp_path / "imnet")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "make_fake_imagenet.py"), d,
                          "--images", "8", "--shards", "2", "--size", "60x80"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert sorted(os.listdir(d)) == ["train-00000-of-00002", "train-00001-of-00002",
                                     "validation-00000-of-00002", "validation-00001-of-00002"]
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "input_pipeline_bench.py"), d,
                          "--seconds", "1", "--batch_size", "4", "--image_size", "32",
                          "--threads", "2", "--shuffle_buffer", "4"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "images/sec" in res.stdout and "2 decode threads" in res.stdout


def test_bench_reference_arm_and_clock_sampler(tmp_path, monkeypatch):
    """bench.py contract pieces that need no GPU: `--impl reference` prints one JSON line with
    `unavailable` and exits 0; the nvidia-smi sampler parses clocks / throttle reasons."""
    import json
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--gpus", "1", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    d = json.loads(res.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and d["unavailable"]

    sys.path.insert(0, ROOT)
    import bench
    s = bench.ClockSampler(0)
    s.proc = type("P", (), {"terminate": lambda self: None, "wait": lambda self, timeout=None: 0,
                            "kill": lambda self: None})()
    s.lines = ["0, 1965, 1965, 650.12, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
               "0, 1905, 1965, 980.50, 0x0000000000000004, Not Active, Not Active, Not Active, Active",
               "0, 1935, 1965, 900.00, 0x0000000000000000, Not Active, Not Active, Not Active, Not Active",
               "garbage line"]
    out = s.stop()
    assert out["sm_mhz"] == 1935.0 and out["sm_max_mhz"] == 1965.0 and out["samples"] == 3
    assert out["reasons"] == ["sw_power_cap"] and out["power_w_max"] == 980.5
    s.lines = ["0, 1200, 1965, 300.0, 0x8, Active, Active, Not Active, Not Active"]
    assert bench.ClockSampler.stop(s)["reasons"] == ["hw_slowdown", "hw_thermal_slowdown"]


def test_native_launcher(tmp_path):
    """kfb-run (csrc/host/launcher/kfb_run.cpp): rank environment, output prefixes, per-rank log
    files named like kungfu-run's, first failure stops the job, timeout, and a real 2-rank
    gloo job started through it."""
    import time
    from kf_benchmarks_b200 import build as build_lib
    exe = build_lib.build_launcher()
    code = "import os;print('hi', os.environ['RANK'], os.environ['WORLD_SIZE'], " \
           "os.environ['LOCAL_WORLD_SIZE'], os.environ['MASTER_ADDR'])"
    res = subprocess.run([exe, "-np", "3", "-logdir", str(tmp_path / "logs"), "--", sys.executable,
                          "-c", code], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert sorted(res.stdout.splitlines()) == ["[rank 1] hi 1 3 3 127.0.0.1",
                                               "[rank 2] hi 2 3 3 127.0.0.1", "hi 0 3 3 127.0.0.1"]
    logs = sorted(os.listdir(str(tmp_path / "logs")))
    assert logs == ["127.0.0.1.1000%d.%s.log" % (r, k) for r in range(3) for k in ("stderr", "stdout")]
    assert open(str(tmp_path / "logs" / "127.0.0.1.10002.stdout.log")).read() == "hi 2 3 3 127.0.0.1\n"
    # the first failing rank ends the job quickly and sets the exit status
    code = "import os,sys,time\nif os.environ['RANK']=='1': sys.exit(7)\ntime.sleep(60)"
    t0 = time.time()
    res = subprocess.run([exe, "-np", "3", "-grace", "1", "--", sys.executable, "-c", code],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 7 and time.time() - t0 < 30
    assert "rank 1 exited with 7" in res.stderr
    res = subprocess.run([exe, "-np", "2", "-timeout", "1", "-grace", "1", "--", "sleep", "30"],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 124 and "timeout" in res.stderr
    assert subprocess.run([exe, "-np", "0", "--", "true"], capture_output=True).returncode == 2
    # a real job: two CPU replicas over gloo, started by the native launcher
    res = subprocess.run([sys.executable, "-m", "kf_benchmarks_b200.parallel.launcher", "-np", "2",
                          "--native", "--", "-m", "kf_benchmarks_b200.cli", "--model=trivial",
                          "--device=cpu", "--data_format=NHWC", "--num_gpus=2", "--batch_size=4",
                          "--variable_update=replicated", "--num_batches=3",
                          "--num_warmup_batches=0", "--display_every=1"],
                         cwd=ROOT, env=dict(os.environ, PYTHONPATH=ROOT), capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "total images/sec:" in res.stdout


def test_analytic_roofline_of_the_headline_step():
    """utils/roofline.py: FLOPs / bytes / roof per kernel class from the graph alone; ResNet-50
    at batch 256 needs 5.9 TFLOP of GEMM work per step (3 x 2 x 3.86 GMAC x 256 minus the stem's
    input gradient) and its roof is a memory-bound ~13 ms at the B200 peaks with the round-2 pass
    structure (batch-norm statistics from the conv epilogue, packed ReLU masks, shortcut gradient
    added by the block's first convolution; 14.1 ms with the round-1 structure)."""
    from kf_benchmarks_b200 import datasets
    from kf_benchmarks_b200.models import model_config
    from kf_benchmarks_b200.utils import roofline
    p = P.make_params(model="resnet50", batch_size=256, use_fp16=True, data_format="NHWC")
    ds = datasets.create_dataset(None, None)
    m = model_config.get_model_config("resnet50", ds, p)
    m.set_batch_size(256)
    g = m.build_graph(ds.num_classes, batch_size=256)
    costs = roofline.model_costs(g, 6550.7e9, 1399.7e12)
    flops = sum(v[0] for v in costs.values())
    assert 5.7e12 < flops < 6.1e12
    fprop = costs["conv fprop (W1)"][0]
    # ResNet-50 v1 (stride in the first 1x1 of a block, as in the reference): 3.86 GMAC / image
    assert abs(fprop - 2 * 256 * 3.86e9) / fprop < 0.02
    roof_ms = sum(v[2] for v in costs.values()) * 1e3
    assert 12.0 < roof_ms < 14.0
    assert sum(v[2] for v in costs.layers.values()) <= sum(v[2] for v in costs.values())
    assert roofline.class_of("void kfb::gemm_tcgen05_kernel<128, 2, true>(...)") == "conv fprop (W1)"
    assert roofline.class_of("void kfb::gemm_tcgen05_kernel<64, 4, true>(...)") == "conv wgrad (W3)"
    assert roofline.class_of("void kfb::bn_bwd_apply_kernel(...)") == "BN bwd apply (W5)"
    assert roofline.class_of("void kfb::bn_bwd_fused_kernel<3, 4, 2>(...)") == "BN bwd apply (W5)"
    assert roofline.class_of("void kfb::bn_fwd_fused_kernel<4, 2>(...)") == \
        "BN fwd apply+ReLU+add (W5)"
    assert roofline.class_of("at::native::vectorized_elementwise_kernel") is None


def test_helper_scripts_parse_and_reference_existing_files():
    """tools/*.sh are only ever run on a GPU box: at least make sure they parse, that every
    repository path they mention exists and that the pytest -k expressions they use select
    something."""
    import glob
    import re
    scripts = sorted(glob.glob(os.path.join(ROOT, "tools", "*.sh")))
    assert len(scripts) >= 6
    test_src = "".join(open(p).read() for p in glob.glob(os.path.join(ROOT, "tests", "test_*.py")))
    for path in scripts:
        assert subprocess.run(["bash", "-n", path], capture_output=True).returncode == 0, path
        text = open(path).read()
        for rel in set(re.findall(r"\b((?:tools|tests|kf_benchmarks_b200)/[\w/]+\.(?:py|sh))\b", text)):
            assert os.path.exists(os.path.join(ROOT, rel)), "%s mentions %s" % (path, rel)
        for mod in set(re.findall(r"-m (kf_benchmarks_b200(?:\.\w+)+)", text)):
            assert os.path.exists(os.path.join(ROOT, mod.replace(".", "/") + ".py")), (path, mod)
        for expr in re.findall(r"-k (\w+)", text):
            assert ("def test_" in test_src) and re.search(r"def test_\w*%s" % expr, test_src), \
                "%s: -k %s selects nothing" % (path, expr)
