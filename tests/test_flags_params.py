"""Flag surface, parsing and validation (reference: T/flags.py, T/benchmark_cnn.py:114-636,
959-1034,1268-1352; tests T/benchmark_cnn_test.py:1179-1234)."""
import pytest

from kf_benchmarks_b200 import benchmark, flags, params as P


def test_all_reference_flags_present_with_defaults():
    assert P.NUM_REFERENCE_FLAGS == 129
    d = P.make_params()
    assert d.model == "trivial" and d.num_gpus == 1 and d.display_every == 10
    assert d.variable_update == "parameter_server" and d.kungfu_option == "sync_sgd"
    assert d.optimizer == "sgd" and d.momentum == 0.9 and d.weight_decay == 4e-5
    assert d.rmsprop_epsilon == 1.0 and d.adam_epsilon == 1e-8
    assert d.data_format == "NCHW" and d.device == "gpu" and d.use_fp16 is False
    assert d.fp16_inc_loss_scale_every_n == 1000 and d.max_ckpts_to_keep == 5
    assert d.eval_dir == "/tmp/tf_cnn_benchmarks/eval" and d.tf_random_seed == 1234
    assert d.compact_gradient_transfer is True and d.variable_consistency == "strong"
    assert d.trt_max_workspace_size_bytes == 4 << 30


def test_parse_styles_and_aliases():
    p = P.make_params_from_flags(["--model=resnet50", "--batch_size", "256", "--use_fp16",
                                  "--nodistortions", "--kungfu_strategy=sma",
                                  "--variable_update=kungfu", "--gradient_clip=1.5",
                                  "--eval_during_training_at_specified_steps=3,7"])
    assert (p.model, p.batch_size, p.use_fp16, p.distortions) == ("resnet50", 256, True, False)
    assert p.kungfu_option == "sma" and p.gradient_clip == 1.5
    assert p.eval_during_training_at_specified_steps == ["3", "7"]
    p = P.make_params_from_flags(["--use_fp16=false", "--xla=true"])
    assert p.use_fp16 is False and p.xla is True


def test_positional_arguments_are_rejected():
    with pytest.raises(flags.FlagError):
        P.make_params_from_flags(["--distortions", "False"])
    with pytest.raises(flags.FlagError):
        P.make_params_from_flags(["--no_such_flag=1"])


def test_bounds_and_enums():
    with pytest.raises(ValueError):
        P.make_params(gradient_repacking=-1)
    with pytest.raises(ValueError):
        P.make_params(gpu_memory_frac_for_testing=1.5)
    with pytest.raises(ValueError):
        P.make_params(optimizer="lamb")
    with pytest.raises(ValueError):
        P.make_params(variable_update="magic")
    with pytest.raises(ValueError):
        P.make_params(not_a_param=1)
    assert P.make_params(kungfu_option="ada_sgd").kungfu_option == "ada_sgd"


def test_remove_param_fields():
    p = P.remove_param_fields(P.make_params(), ["model", "xla"])
    assert not hasattr(p, "model") and hasattr(p, "batch_size")


@pytest.mark.parametrize("kw", [
    dict(device="cpu", data_format="NCHW"),
    dict(num_epochs_per_decay=2.0),
    dict(minimum_learning_rate=0.1),
    dict(variable_update="horovod", num_gpus=2),
    dict(variable_update="kungfu", num_gpus=2),
    dict(use_fp16=True, fp16_vars=True, gradient_repacking=2),
    dict(use_fp16=True, fp16_enable_auto_loss_scale=True, variable_update="kungfu"),
    dict(hierarchical_copy=True, num_gpus=1),
    dict(save_model_secs=10, save_model_steps=5),
    dict(eval_during_training_every_n_steps=1, eval_during_training_every_n_epochs=1.0),
    dict(eval_during_training_every_n_steps=1, forward_only=True),
    dict(stop_at_top_1_accuracy=0.5),
    dict(staged_vars=True, variable_update="replicated"),
    dict(trt_mode="FP16"),
    dict(debugger="nonsense"),
    dict(all_reduce_spec="bogus_alg"),
])
def test_invalid_combinations_raise(kw):
    base = dict(device="cpu", data_format="NHWC")
    base.update(kw)
    with pytest.raises(ValueError):
        benchmark.BenchmarkCNN(P.make_params(**base))


def test_unsupported_features_fail_loudly_in_setup():
    for kw in (dict(xla=True), dict(job_name="worker"), dict(trt_mode="FP16")):
        with pytest.raises(ValueError):
            benchmark.setup(P.make_params(**kw))


def test_ps_and_controller_roles_exit_cleanly():
    for role in ("ps", "controller"):
        with pytest.raises(SystemExit) as e:
            benchmark.setup(P.make_params(job_name=role, ps_hosts="a:1", worker_hosts="b:2"))
        assert e.value.code == 0


def test_warmup_default_and_batch_sizes():
    b = benchmark.BenchmarkCNN(P.make_params(device="cpu", data_format="NHWC", model="resnet50"))
    assert b.num_warmup_batches == 10 and b.batch_size_per_device == 64 and b.num_batches == 100
    b = benchmark.BenchmarkCNN(P.make_params(device="cpu", data_format="NHWC", autotune_threshold=3,
                                             batch_size=8, num_gpus=1))
    assert b.num_warmup_batches == 45 and b.batch_size == 8


def test_every_flag_is_either_used_or_declared_without_effect():
    """No flag is swallowed silently: each one is read somewhere in the package, rejected with
    an error, or listed in IGNORED_ON_B200 (a note is logged when it is set)."""
    import os
    import re
    from kf_benchmarks_b200 import params as P
    root = os.path.dirname(os.path.abspath(P.__file__))
    src = ""
    for d, _, fs in os.walk(root):
        for f in fs:
            if f.endswith(".py") and f != "params.py":
                with open(os.path.join(d, f)) as fh:
                    src += fh.read()
    declared = set(P.IGNORED_ON_B200) | set(P.UNSUPPORTED_WHEN_SET)
    silent = [f for f in P.make_params()._fields
              if f not in declared and not re.search(r"\b%s\b" % re.escape(f), src)]
    assert silent == []
    assert not (set(P.IGNORED_ON_B200) - set(P.make_params()._fields))


def test_ctypes_mirrors_match_the_built_library():
    """The structures that cross the C ABI have the size the library was compiled with
    (the library loads without a GPU)."""
    import ctypes
    from kf_benchmarks_b200 import build as build_lib
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.parallel.strategy import Hyper, PushTarget
    if not native.available():
        build_lib.build()
    lib = native.load()
    assert lib.kfb_sizeof(0) == ctypes.sizeof(native.ConvDesc)
    assert lib.kfb_sizeof(1) == ctypes.sizeof(Hyper)
    assert lib.kfb_sizeof(2) == ctypes.sizeof(PushTarget)
    assert lib.kfb_sizeof(99) == -1


def test_every_native_call_site_names_an_exported_function_with_the_right_arity():
    """Static check of the Python -> C ABI boundary: every ``native.call("name", ...)`` in the
    package and the tools names a function the built library exports, and passes as many
    arguments as the prototype in include/kfb200.h declares (sites with a starred argument are
    checked for the minimum only)."""
    import ast
    import os
    import re
    from kf_benchmarks_b200 import build as build_lib
    from kf_benchmarks_b200.ops import native
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not native.available():
        build_lib.build()
    lib = native.load()
    header = open(os.path.join(root, "include", "kfb200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    arity = {}
    for m in re.finditer(r"\b(?:int|long long|const char\*)\s+kfb_(\w+)\s*\(([^;{]*?)\)\s*;", header):
        params = m.group(2).strip()
        arity[m.group(1)] = 0 if params in ("", "void") else params.count(",") + 1
    assert len(arity) > 60
    sites = 0
    for top in ("kf_benchmarks_b200", "tools"):
        for d, _, fs in os.walk(os.path.join(root, top)):
            for f in fs:
                if not f.endswith(".py"):
                    continue
                tree = ast.parse(open(os.path.join(d, f)).read())
                for node in ast.walk(tree):
                    if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute)
                            and node.func.attr == "call" and isinstance(node.func.value, ast.Name)
                            and node.func.value.id == "native" and node.args
                            and isinstance(node.args[0], ast.Constant)
                            and isinstance(node.args[0].value, str)):
                        continue
                    name = node.args[0].value
                    where = "%s:%d native.call(%r)" % (f, node.lineno, name)
                    assert hasattr(lib, "kfb_" + name), where + ": not exported"
                    assert name in arity, where + ": no prototype in include/kfb200.h"
                    given = len(node.args) - 1
                    starred = any(isinstance(a, ast.Starred) for a in node.args)
                    if starred:
                        assert given - 1 <= arity[name], where
                    else:
                        assert given == arity[name], "%s: %d arguments, prototype has %d" % (
                            where, given, arity[name])
                    sites += 1
    assert sites > 80
    # direct calls through the ctypes handle (lib.kfb_xxx(...)) name exported functions too
    host = None
    for top in ("kf_benchmarks_b200", "tools", "tests"):
        for d, _, fs in os.walk(os.path.join(root, top)):
            for f in fs:
                if not f.endswith(".py") or f == "test_flags_params.py":
                    continue
                for name in set(re.findall(r"\.kfb_(\w+)\b", open(os.path.join(d, f)).read())):
                    if hasattr(lib, "kfb_" + name):
                        continue
                    if host is None:
                        from kf_benchmarks_b200.data import native_reader
                        build_lib.build_host()
                        host = native_reader._load()
                    assert hasattr(host, "kfb_" + name), "%s: kfb_%s is not exported" % (f, name)


def test_native_call_argument_types_match_the_prototypes():
    """Static type check of the Python -> C boundary: where a call site builds an argument with
    a recognisable constructor (i32 / i64 / f32 / u64 / c_uint / ptr / byref / None ...), its
    kind must be the kind of the parameter in include/kfb200.h - an int passed for a long long
    (or the reverse) leaves half a register undefined."""
    import ast
    import os
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "kfb200.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)

    def kind_of_param(p):
        p = p.strip()
        if "*" in p:
            return "ptr"
        for prefix, kind in (("unsigned long long", "u64"), ("long long", "i64"),
                             ("unsigned", "u32"), ("float", "f32"), ("int", "i32")):
            if p.startswith(prefix):
                return kind
        return "?"

    protos = {}
    for m in re.finditer(r"\b(?:int|long long|const char\*)\s+kfb_(\w+)\s*\(([^;{]*?)\)\s*;", header):
        ps = m.group(2).strip()
        protos[m.group(1)] = [] if ps in ("", "void") else [kind_of_param(x) for x in ps.split(",")]
    ctor = {"i32": "i32", "i64": "i64", "f32": "f32", "u64": "u64", "c_uint": "u32",
            "c_int": "i32", "c_longlong": "i64", "c_ulonglong": "u64", "c_float": "f32",
            "ptr": "ptr", "byref": "ptr", "c_void_p": "ptr", "_s": "ptr", "stream_ptr": "ptr",
            "_w16": "ptr", "_w32": "ptr", "_g32": "ptr", "_ptr_at": "ptr"}

    def kind_of_expr(e):
        if isinstance(e, ast.Constant) and e.value is None:
            return "ptr"
        if isinstance(e, ast.Call):
            f = e.func
            name = f.id if isinstance(f, ast.Name) else (f.attr if isinstance(f, ast.Attribute)
                                                         else None)
            return ctor.get(name)
        if isinstance(e, ast.Attribute) and e.attr == "p":
            return "ptr"
        if isinstance(e, ast.IfExp):
            a, b = kind_of_expr(e.body), kind_of_expr(e.orelse)
            return a if a == b else (a or b)
        return None

    bad, checked = [], 0
    for top in ("kf_benchmarks_b200", "tools"):
        for d, _, fs in os.walk(os.path.join(root, top)):
            for f in fs:
                if not f.endswith(".py"):
                    continue
                tree = ast.parse(open(os.path.join(d, f)).read())
                for node in ast.walk(tree):
                    if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute)
                            and node.func.attr == "call" and isinstance(node.func.value, ast.Name)
                            and node.func.value.id == "native" and node.args
                            and isinstance(node.args[0], ast.Constant)):
                        continue
                    name = node.args[0].value
                    if name not in protos or any(isinstance(a, ast.Starred) for a in node.args):
                        continue
                    for i, (a, want) in enumerate(zip(node.args[1:], protos[name])):
                        got = kind_of_expr(a)
                        if got is None:
                            continue
                        checked += 1
                        if got != want:
                            bad.append("%s:%d %s arg %d: passes %s, prototype wants %s" %
                                       (f, node.lineno, name, i, got, want))
    assert checked > 600
    assert not bad, "\n".join(bad)


def test_fp32_on_the_gpu_is_refused_not_relabelled():
    """--use_fp16 defaults to false in the reference (T/benchmark_cnn.py:464-466); the sm_100a
    kernels are fp16-storage only, so the GPU path refuses the fp32 label instead of silently
    computing in fp16 (the CPU path does full fp32)."""
    import pytest
    from kf_benchmarks_b200 import benchmark
    p = P.make_params(model="trivial", device="gpu", num_batches=1)
    with pytest.raises(ValueError, match="use_fp16=false is not available with --device=gpu"):
        benchmark.BenchmarkCNN(benchmark.setup(p))
    with pytest.raises(ValueError, match="variable_consistency=relaxed"):
        benchmark.BenchmarkCNN(benchmark.setup(P.make_params(
            model="trivial", device="cpu", data_format="NHWC", use_fp16=True,
            fp16_enable_auto_loss_scale=True, variable_consistency="relaxed",
            variable_update="replicated")))
