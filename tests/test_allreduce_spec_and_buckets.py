"""all_reduce_spec grammar and bucket planning (T/allreduce.py:35-104,420-454;
tests T/allreduce_test.py:32-300)."""
import pytest

from kf_benchmarks_b200.parallel import allreduce_spec as A
from kf_benchmarks_b200.parallel.strategy import plan_buckets, reduction_semantics
from kf_benchmarks_b200 import params as P


def test_parse_general_int():
    assert A.parse_general_int("64") == 64
    assert A.parse_general_int("4k") == 4096 and A.parse_general_int("4K") == 4096
    assert A.parse_general_int("2M") == 2 << 20 and A.parse_general_int("1G") == 1 << 30
    with pytest.raises(ValueError):
        A.parse_general_int("4q")


def test_parse_all_reduce_spec():
    assert A.parse_all_reduce_spec("nccl") == [A.AllReduceSpecTuple("nccl", 1, -1)]
    assert A.parse_all_reduce_spec("xring#2") == [A.AllReduceSpecTuple("xring", 2, -1)]
    spec = A.parse_all_reduce_spec("pscpu:32k:xring")
    assert spec == [A.AllReduceSpecTuple("pscpu", 1, 32768), A.AllReduceSpecTuple("xring", 1, -1)]
    spec = A.parse_all_reduce_spec("psgpu#4:1M:nccl/xring#2:64M:collective")
    assert [s.alg for s in spec] == ["psgpu", "nccl/xring", "collective"]
    assert [s.limit for s in spec] == [1 << 20, 64 << 20, -1]
    for bad in ("nccl:abc:xring", "nccl:64k", "foo", "nccl:2M:xring:1M:pscpu"):
        with pytest.raises(ValueError):
            A.parse_all_reduce_spec(bad)


def test_extract_ranges():
    assert A.extract_ranges([]) == ([], [])
    assert A.extract_ranges([1, 3, 4, 6, 7, 8, 9]) == ([[3, 4], [6, 9]], [1])
    r, s = A.extract_ranges(list(range(10)), range_size_limit=3)
    assert r == [[0, 4], [5, 9]] and s == []
    groups, singles = A.small_tensor_groups([4, 4, 4000, 4, 4, 4, 9000, 4], 64, 10)
    assert groups == [[0, 1], [3, 4, 5]] and singles == [7]


def test_plan_buckets():
    b = plan_buckets(1000, 256)
    assert b[0] == (0, 256) and b[-1][1] == 1000 and all(lo % 128 == 0 for lo, _ in b)
    b = plan_buckets(10_000, 999999, num_buckets=4)
    assert len(b) == 4 and b[-1][1] == 10_000
    assert plan_buckets(0, 128) == []
    cover = []
    for lo, hi in plan_buckets(25_600_000, 2 << 20):
        cover.append(hi - lo)
    assert sum(cover) == 25_600_000


@pytest.mark.parametrize("vu,opt,expect", [
    ("independent", "sync_sgd", ("none", False)),
    ("replicated", "sync_sgd", ("allreduce", False)),     # sums, T/batch_allreduce.py:24-29
    ("horovod", "sync_sgd", ("allreduce", False)),        # hvd.allreduce(average=False)
    ("parameter_server", "sync_sgd", ("allreduce", True)),
    ("kungfu", "sync_sgd", ("allreduce", True)),
    ("kungfu", "sma", ("sma", True)),
    ("kungfu", "async_sgd", ("pair", False)),
    ("kungfu", "ada_sgd", ("ada", True)),
])
def test_reduction_semantics(vu, opt, expect):
    assert reduction_semantics(P.make_params(variable_update=vu, kungfu_option=opt)) == expect


def test_fused_push_buckets_cut_only_between_variables():
    """--comm_impl=p2p_fused: every weight gradient must lie in exactly one bucket so that its
    owner map is one (bucket start, shard length) pair (parallel/strategy.py)."""
    import types
    from kf_benchmarks_b200 import datasets
    from kf_benchmarks_b200.models import model_config
    from kf_benchmarks_b200.nn.params import ParameterStore
    from kf_benchmarks_b200.parallel.strategy import PushTarget, UpdateEngine
    ds = datasets.create_dataset(None, "cifar10")
    g = model_config.get_model_config("resnet20", ds, None).build_graph(11, batch_size=2)
    st = ParameterStore(g, "cpu", fp16_shadow=False)
    buckets = UpdateEngine._plan_buckets_on_slots(types.SimpleNamespace(store=st), 40000)
    assert buckets[0][0] == 0 and buckets[-1][1] == st.num_trainable_padded
    assert all(a[1] == b[0] for a, b in zip(buckets, buckets[1:]))
    starts = {s.offset for s in st.order} | {st.num_trainable_padded}
    assert all(lo in starts and hi in starts for lo, hi in buckets)
    assert all(hi - lo >= 40000 for lo, hi in buckets[:-1]) and len(buckets) > 2
    for s in st.order:       # no variable straddles a bucket boundary
        assert sum(1 for lo, hi in buckets if lo <= s.offset < hi) == 1
        lo, hi = next(b for b in buckets if b[0] <= s.offset < b[1])
        assert s.offset + s.numel <= hi
    t = PushTarget()
    t.nranks, t.per, t.rel = 8, 1024, 256
    assert ctypes_sizeof(t) >= 4 + 8 + 8 + 8 * 8


def ctypes_sizeof(obj):
    import ctypes
    return ctypes.sizeof(obj)


def test_nvls_symmetric_allocation_bookkeeping(monkeypatch):
    """--comm_impl=nvls: peer pointers / multicast alias follow the tensor's offset inside the
    symmetric block, and a fabric without multicast support is a loud error (torch's symmetric
    memory and the device are faked; the kernels are covered by tools/nvls_check.py)."""
    import types

    import torch
    import torch.distributed._symmetric_memory as symm_mem
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.parallel.comm import P2PComm

    class FakeTensor:
        def __init__(self, n):
            self.n = n

        def zero_(self):
            return self

        def data_ptr(self):
            return 0x7000_0000_1000 + 0x40          # 64 bytes into the block

        def __getitem__(self, sl):
            return self

    class FakeHandle:
        buffer_ptrs = [0x7000_0000_1000, 0x7100_0000_0000]
        multicast_ptr = 0x7F00_0000_0000

    monkeypatch.setattr(symm_mem, "empty", lambda n, dtype, device: FakeTensor(n))
    monkeypatch.setattr(symm_mem, "rendezvous", lambda t, group: FakeHandle())
    monkeypatch.setattr(torch.cuda, "synchronize", lambda dev=None: None)
    fake_dist = types.SimpleNamespace(group=types.SimpleNamespace(WORLD=object()),
                                      barrier=lambda: None)
    comm = types.SimpleNamespace(rank=0, size=2, device=torch.device("cpu"), dist=fake_dist,
                                 _by_ptr={}, _mc_by_ptr={}, _symm_keep=[])
    t = P2PComm._symm_alloc_multicast(comm, 1000, torch.float32)
    assert comm._by_ptr[t.data_ptr()] == [0x7000_0000_1040, 0x7100_0000_0040]
    assert comm._mc_by_ptr[t.data_ptr()] == 0x7F00_0000_0040
    assert P2PComm.mc_ptr(comm, t) == 0x7F00_0000_0040

    FakeHandle.multicast_ptr = 0
    with pytest.raises(native.KfbError, match="NVLS"):
        P2PComm._symm_alloc_multicast(comm, 1000, torch.float32)


def test_sequence_numbers_stay_consecutive_across_graph_replays():
    """Multi-rank CUDA graphs: a captured step replays the same by-value sequence numbers, the
    kernels add a device counter that the step's last node advances. Model the device side and
    check that the effective numbers (what the flag protocol compares) increase by one per
    collective through eager calls, capture, replays and eager calls again."""
    import types
    from kf_benchmarks_b200.parallel.comm import P2PComm
    comm = types.SimpleNamespace(seq=0, _dev_base=0)
    device_counter = 0
    effective = []

    def launch(rel):                      # what a kernel does with the number it is given
        effective.append((rel + device_counter) & 0xFFFFFFFF)

    for _ in range(5):                    # warm-up steps, eager
        launch(P2PComm.next_seq(comm))
    seq0 = comm.seq
    captured = [P2PComm.next_seq(comm) for _ in range(3)]     # capture: nothing executes
    k = comm.seq - seq0
    comm.seq = seq0
    for _ in range(4):                    # replays
        for rel in captured:
            launch(rel)
        device_counter += k               # kfb_seq_advance, last node of the graph
        P2PComm.graph_replayed(comm, k)
    for _ in range(2):                    # e.g. the barrier at exit
        launch(P2PComm.next_seq(comm))
    assert effective == list(range(1, 5 + 4 * 3 + 2 + 1))
    # wrap-around of the 32-bit counter keeps the differences intact
    comm = types.SimpleNamespace(seq=2 ** 32 - 2, _dev_base=7)
    a, b, c = (P2PComm.next_seq(comm) for _ in range(3))
    assert ((b - a) & 0xFFFFFFFF) == 1 and ((c - b) & 0xFFFFFFFF) == 1


def test_graphed_step_control_flow_with_two_ranks(monkeypatch):
    """`BenchmarkCNN._step_graphed` for a multi-rank job, driven with stand-ins for the device
    (no CUDA here): two eager steps, one capture, replays - the collectives' sequence numbers
    and the optimizer step count stay in step with what an eager run would have used."""
    import contextlib
    import types

    import torch
    from kf_benchmarks_b200 import benchmark
    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.parallel.comm import P2PComm

    calls = []
    monkeypatch.setattr(native, "call", lambda name, *a: calls.append(name))
    monkeypatch.setattr(native, "stream_ptr", lambda: None)
    monkeypatch.setattr(native, "ptr", lambda t: None)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda dev=None: None)
    monkeypatch.setattr(torch.cuda, "CUDAGraph", lambda: types.SimpleNamespace(
        replay=lambda: calls.append("replay")))
    monkeypatch.setattr(torch.cuda, "graph", lambda g, **kw: contextlib.nullcontext())

    comm = types.SimpleNamespace(seq=0, _dev_base=0, barrier=lambda: None)
    comm.enable_device_seq = lambda: "counter"
    comm.graph_replayed = lambda k: P2PComm.graph_replayed(comm, k)
    launched = []

    class Engine:
        step_count = 0
        loss_scale = 1.0

        def loss_scale_kwargs(self):
            return dict(loss_scale=1.0, dyn_scale=None)

        def learning_rate(self, *a):
            return 0.1

        def enable_dynamic_hyper(self):
            pass

        def push_dynamic_hyper(self, lr):
            self.pushed_t = self.step_count + 1

        def begin_step(self, lr, push_dyn=True):
            self.begin_push = getattr(self, "begin_push", []) + [push_dyn]

        def on_grads_ready(self, lo, hi):
            pass

        def finish_step(self):
            for _ in range(3):                                   # three buckets per step
                launched.append(P2PComm.next_seq(comm))
            self.step_count += 1

    ex = types.SimpleNamespace(set_inputs=lambda *a: None, forward=lambda train: None,
                               loss_backward=lambda **kw: None,
                               push_step_values=lambda: None, step_index=0)
    b = types.SimpleNamespace(executor=ex, engine=Engine(), comm=comm, world=2, global_step=0,
                              device=torch.device("cpu"), dataset=types.SimpleNamespace(
                                  num_examples_per_epoch=lambda s: 1000),
                              input_source=types.SimpleNamespace(next_batch=lambda: ()),
                              _examples_per_step=lambda: 8)
    for _ in range(6):
        benchmark.BenchmarkCNN._step_graphed(b)
    assert b.global_step == 6 and b.engine.step_count == 6
    assert b._graph_collectives == 3 and calls.count("replay") == 4
    assert calls.count("seq_advance") == 1                        # recorded once, in the capture
    assert launched == [1, 2, 3, 4, 5, 6, 7, 8, 9]                # 2 eager steps + the capture
    assert comm.seq == 6 + 4 * 3 and comm._dev_base == 4 * 3      # 4 replays moved the mirror
    assert b.engine.pushed_t == 6                                 # Adam's t of the 6th step
    assert b.engine.begin_push == [True, True, False]             # no push inside the capture
    assert P2PComm.next_seq(comm) == 19 - 12                      # an eager call afterwards


def test_allreduce_call_passes_the_prototypes_argument_lists(monkeypatch):
    """UpdateEngine._allreduce_call builds its argument list with a starred tail: run it with
    stand-ins and compare the lists with the prototypes of include/kfb200.h (pull and NVLS)."""
    import ctypes
    import os
    import re
    import types

    from kf_benchmarks_b200.ops import native
    from kf_benchmarks_b200.parallel.strategy import Hyper, UpdateEngine
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = re.sub(r"/\*.*?\*/", "", open(os.path.join(root, "include", "kfb200.h")).read(),
                    flags=re.S)

    def arity(name):
        m = re.search(r"\bint\s+kfb_%s\s*\(([^;{]*?)\)\s*;" % name, header)
        return m.group(1).count(",") + 1

    seen = {}
    monkeypatch.setattr(native, "call", lambda name, *a: seen.__setitem__(name, a))

    class T:                                    # stand-in for a device tensor
        def __init__(self, addr):
            self.addr = addr

        def data_ptr(self):
            return self.addr

    store = types.SimpleNamespace(w32=T(0x1000), grad=T(0x2000), w16=T(0x3000))
    comm = types.SimpleNamespace(rank=1, size=4, next_seq=lambda: 77)
    eng = types.SimpleNamespace(
        store=store, comm=comm, m=T(0x4000), v=None, dyn=None, reduced=T(0x5000),
        params=types.SimpleNamespace(sma_alpha=0.25), nvls=False, ar_v2=False, compact=False,
        _w_ptrs="W", _grad_ptrs="G", _red_ptrs="R", _flag_ptrs="F",
        _w_mc=ctypes.c_void_p(0x6000), _grad_mc=ctypes.c_void_p(0x7000),
        _red_mc=ctypes.c_void_p(0x8000), _s=lambda stream=None: ctypes.c_void_p(9))
    h = Hyper()
    UpdateEngine._allreduce_call(eng, 0, 128, 4096, h, None, None)
    a = seen["allreduce_update"]
    assert len(a) == arity("allreduce_update") == 20
    assert (a[0].value, a[1].value, a[2].value, a[3], a[4], a[5]) == (0, 1, 4, "G", "R", "F")
    assert (a[6].value, a[7].value, a[8].value) == (128, 4096, 77)
    assert [x.value for x in a[9:14]] == [0x1000, 0x2000, 0x4000, None, 0x3000]   # w g m v w16
    assert a[17].value == 0.0 and a[18].value == 0 and a[19].value == 9
    eng.nvls = True
    UpdateEngine._allreduce_call(eng, 1, 0, 64, h, None, None)
    b = seen["allreduce_update_nvls"]
    assert len(b) == arity("allreduce_update_nvls") == 21
    assert (b[0].value, b[3].value, b[4].value, b[5].value, b[6]) == (1, 0x6000, 0x8000, 0x5000, "F")
    assert b[18].value == 0.25                                    # SMA blend factor, mode 1
    # second-generation entry point: pull form, compact (fp16) transfer of a gradient bucket
    eng.nvls, eng.ar_v2, eng.compact = False, True, True
    eng.g16, eng.r16 = T(0xA000), T(0xB000)
    eng._g16_ptrs, eng._r16_ptrs = "G16", "R16"
    UpdateEngine._allreduce_call(eng, 0, 128, 4096, h, None, None)
    d = seen["allreduce_update2"]
    assert len(d) == arity("allreduce_update2") == 26
    assert (d[3], d[4], d[5], d[6], d[7].value, d[8], d[9].value) == \
        ("G16", "R16", None, None, 0xB000, "F", 1)
    assert (d[10].value, d[11].value, d[12].value, d[13].value, d[14].value) == \
        (0x2000, 0xA000, 128, 4096, 77)
    # NVLS form of SMA: fp32 weights through the multicast aliases, no compact transfer
    eng.nvls = True
    UpdateEngine._allreduce_call(eng, 1, 0, 64, h, None, None)
    d = seen["allreduce_update2"]
    assert (d[3], d[4], d[5].value, d[6].value, d[7].value, d[9].value) == \
        (None, None, 0x6000, 0x8000, 0x5000, 0)
    assert d[10] is None and d[11] is None and d[23].value == 0.25
    # ranges that are not multiples of 8 elements stay on the first-generation kernels
    UpdateEngine._allreduce_call(eng, 0, 4, 60, h, None, None)
    assert seen["allreduce_update_nvls"][7].value == 4
