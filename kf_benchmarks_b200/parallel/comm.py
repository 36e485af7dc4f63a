"""Communicators: how one rank (= one process = one GPU) talks to its peers.

The reference reaches its peers through four external stacks (TF in-process NCCL /
cross-device copies, TF collective ops over gRPC, Horovod, KungFu - SURVEY.md 2.4). Here
there is one process model - a rank per GPU, rendezvous through ``torch.distributed`` for
the plumbing only - and three interchangeable communicators:

* ``P2PComm``   - the product path. Gradient / weight arenas, landing buffers and a flag
  page are cudaMalloc'ed by the native runtime, exported over CUDA IPC and mapped by every
  peer, so the fused kernels of ``csrc/optim_comm.cu`` read and write peer HBM directly
  over NVLink 5 / NVSwitch.
  With ``nvls=True`` (``--comm_impl=nvls``) the same allocations come from torch's symmetric
  memory instead (VMM allocations bound to an NVSwitch multicast object): peers are still
  addressable one by one, and ``mc_ptr`` gives the multicast alias the ``multimem`` kernels
  reduce / broadcast through.
* ``TorchDistComm`` - ``torch.distributed`` all-reduce (NCCL on GPUs, gloo on CPU ranks)
  followed by a separate update: the "only calls NCCL" baseline and the CPU-replica path.
* ``SingleComm`` - world size 1.
"""
from __future__ import annotations

import ctypes
import os
from typing import List

import torch

from ..ops import native
from ..ops.native import i32, ptr


class _CudaArray:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw device pointer."""

    def __init__(self, addr: int, numel: int, typestr: str):
        self.__cuda_array_interface__ = {
            "shape": (numel,), "typestr": typestr, "data": (addr, False), "version": 3,
            "strides": None,
        }


_TYPESTR = {torch.float32: "<f4", torch.float16: "<f2", torch.int32: "<i4", torch.uint8: "|u1"}


def wrap_device_pointer(addr: int, numel: int, dtype: torch.dtype, device) -> torch.Tensor:
    return torch.as_tensor(_CudaArray(addr, numel, _TYPESTR[dtype]), device=device)


class Comm:
    rank = 0
    size = 1
    kind = "single"

    def barrier(self) -> None:
        pass

    def allgather_object(self, obj):
        return [obj]

    def allocator(self):
        """Arena allocator handed to ParameterStore (None: plain torch allocation)."""
        return None

    def max_over_ranks(self, value: float) -> float:
        return value

    def close(self) -> None:
        pass


class SingleComm(Comm):
    pass


def init_process_group(backend: str) -> None:
    import torch.distributed as dist
    if dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    dist.init_process_group(backend=backend, rank=int(os.environ["RANK"]),
                            world_size=int(os.environ["WORLD_SIZE"]))


class TorchDistComm(Comm):
    """torch.distributed plumbing; ``backend`` nccl (GPU baseline) or gloo (CPU replicas)."""

    def __init__(self, backend: str, device):
        import torch.distributed as dist
        init_process_group(backend)
        self.dist = dist
        self.rank = dist.get_rank()
        self.size = dist.get_world_size()
        self.kind = backend
        self.device = torch.device(device)

    def barrier(self) -> None:
        self.dist.barrier()

    def allgather_object(self, obj):
        out = [None] * self.size
        self.dist.all_gather_object(out, obj)
        return out

    def all_reduce_(self, t: torch.Tensor) -> torch.Tensor:
        self.dist.all_reduce(t)
        return t

    def broadcast_(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        self.dist.broadcast(t, src=src)
        return t

    def max_over_ranks(self, value: float) -> float:
        t = torch.tensor([value], dtype=torch.float64,
                         device=self.device if self.kind == "nccl" else "cpu")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


class P2PComm(Comm):
    """Symmetric memory over CUDA IPC + the fused NVLink kernels.

    Every allocation made through ``allocator()`` is symmetric: all ranks call it in the same
    order with the same size, the handles are exchanged once, and ``peers(tensor)`` returns the
    per-rank device pointers of that allocation as mapped into this process.
    """
    kind = "p2p"

    def __init__(self, device, bootstrap_backend: str = "gloo", nvls: bool = False):
        import torch.distributed as dist
        # the process group is plumbing only (rendezvous, barriers, object exchange on the host).
        # KFB_BOOTSTRAP_BACKEND overrides it, e.g. "cpu:gloo,cuda:nccl" should torch's symmetric
        # memory (nvls) want a CUDA-capable group for its rendezvous.
        init_process_group(os.environ.get("KFB_BOOTSTRAP_BACKEND") or bootstrap_backend)
        self.dist = dist
        self.rank = dist.get_rank()
        self.size = dist.get_world_size()
        if self.size > 8:
            raise ValueError("P2PComm supports up to 8 ranks (one NVSwitch box)")
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.lib = native.load()
        self._allocs = []  # (local_ptr, nbytes, [peer ptrs])
        self._by_ptr = {}
        self.nvls = bool(nvls)
        self._mc_by_ptr = {}   # local pointer -> multicast alias (nvls only)
        self._symm_keep = []   # torch symmetric-memory tensors and handles (nvls only)
        if self.nvls:
            self.kind = "nvls"
        self.seq = 0
        # device-resident part of the sequence numbers (multi-rank CUDA graphs, see
        # enable_device_seq): the kernels add the device counter, the host mirrors it
        self._seq_counter = None
        self._dev_base = 0
        flag_bytes = int(self.lib.kfb_flag_page_bytes())
        self.flags = self.symm_alloc(flag_bytes // 4, torch.int32)
        self._flag_ptrs = self._ptr_array(self.flags)
        self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.barrier()

    # -- symmetric allocation -------------------------------------------------------
    def _symm_alloc_multicast(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        """Symmetric allocation bound to a multicast object (torch symmetric memory does the
        cuMemCreate / cuMulticastBindMem / handle exchange; the kernels are ours)."""
        import torch.distributed._symmetric_memory as symm_mem
        esize = torch.empty((), dtype=dtype).element_size()
        padded = (max(int(numel) * esize, 256) + 15) // 16 * 16 // esize
        t = symm_mem.empty(padded, dtype=dtype, device=self.device)
        t.zero_()
        torch.cuda.synchronize(self.device)
        try:
            hdl = symm_mem.rendezvous(t, self.dist.group.WORLD)
        except Exception as e:
            raise native.KfbError(
                "--comm_impl=nvls: torch symmetric-memory rendezvous failed (%s). If it objects "
                "to the gloo bootstrap group, set KFB_BOOTSTRAP_BACKEND=cpu:gloo,cuda:nccl" % e)
        base = [int(a) for a in hdl.buffer_ptrs]
        if len(base) != self.size:
            raise native.KfbError("symmetric memory rendezvous returned %d peers, expected %d"
                                  % (len(base), self.size))
        # the handle describes the allocation block; the tensor may start inside it
        off = t.data_ptr() - base[self.rank]
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if mc == 0:
            raise native.KfbError(
                "--comm_impl=nvls needs NVSwitch multicast (NVLS) support, which this "
                "driver / fabric does not report; use --comm_impl=p2p")
        peer_ptrs = [a + off for a in base]
        self._by_ptr[t.data_ptr()] = peer_ptrs
        self._mc_by_ptr[t.data_ptr()] = mc + off
        self._symm_keep.append((t, hdl))
        self.dist.barrier()   # every rank has zeroed and mapped before anyone writes
        return t[:int(numel)] if padded != int(numel) else t

    def mc_ptr(self, t: torch.Tensor) -> int:
        """Multicast alias of the symmetric allocation `t` starts at (nvls only)."""
        return self._mc_by_ptr[t.data_ptr()]

    def symm_alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        if self.nvls:
            return self._symm_alloc_multicast(numel, dtype)
        nbytes = max(int(numel) * torch.empty((), dtype=dtype).element_size(), 256)
        p = ctypes.c_void_p()
        native.check(self.lib.kfb_symm_alloc(ctypes.c_longlong(nbytes), ctypes.byref(p)),
                     "symm_alloc")
        hb = int(self.lib.kfb_ipc_handle_bytes())
        handle = ctypes.create_string_buffer(hb)
        native.check(self.lib.kfb_ipc_get_handle(p, handle), "ipc_get_handle")
        handles = self.allgather_object(bytes(handle.raw))
        peer_ptrs = []
        for r, h in enumerate(handles):
            if r == self.rank:
                peer_ptrs.append(p.value)
                continue
            q = ctypes.c_void_p()
            native.check(self.lib.kfb_ipc_open(ctypes.create_string_buffer(h, hb),
                                               ctypes.byref(q)), "ipc_open")
            peer_ptrs.append(q.value)
        t = wrap_device_pointer(p.value, int(numel), dtype, self.device)
        self._allocs.append((p.value, nbytes, peer_ptrs))
        self._by_ptr[p.value] = peer_ptrs
        return t

    def allocator(self):
        return lambda n, dt: self.symm_alloc(n, dt)

    def peers(self, t: torch.Tensor) -> List[int]:
        """Per-rank pointers of the symmetric allocation `t` starts at."""
        return self._by_ptr[t.data_ptr()]

    def _ptr_array(self, t: torch.Tensor):
        arr = (ctypes.c_void_p * 8)()
        for r, a in enumerate(self.peers(t)):
            arr[r] = a
        return arr

    def ptr_array(self, t: torch.Tensor):
        return self._ptr_array(t)

    # -- plumbing ----------------------------------------------------------------------
    def allgather_object(self, obj):
        out = [None] * self.size
        self.dist.all_gather_object(out, obj)
        return out

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        self.dist.barrier()

    def max_over_ranks(self, value: float) -> float:
        t = torch.tensor([value], dtype=torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def next_seq(self) -> int:
        """Sequence number of the next collective as the kernels must be given it: the absolute
        count minus what the device counter already holds (0 unless enable_device_seq)."""
        self.seq += 1
        return (self.seq - self._dev_base) & 0xFFFFFFFF

    def enable_device_seq(self) -> torch.Tensor:
        """Keep part of the sequence numbers in device memory so that a captured CUDA graph,
        which replays the same by-value numbers, still presents fresh ones: every collective
        adds the counter, the captured step ends with ``kfb_seq_advance`` and ``graph_replayed``
        moves the host mirror. Must be called on every rank at the same point."""
        if self._seq_counter is None:
            self._seq_counter = torch.zeros(1, dtype=torch.int32, device=self.device)
            torch.cuda.synchronize(self.device)
            native.check(self.lib.kfb_set_seq_base(ptr(self._seq_counter)), "set_seq_base")
        return self._seq_counter

    def graph_replayed(self, collectives_per_step: int) -> None:
        """Host mirror of the device counter after one replay of a captured step."""
        self.seq += collectives_per_step
        self._dev_base += collectives_per_step

    def device_barrier(self, stream=None) -> None:
        s = ctypes.c_void_p((stream or torch.cuda.current_stream(self.device)).cuda_stream)
        native.call("device_barrier", i32(self.rank), i32(self.size), self._flag_ptrs,
                    ctypes.c_uint(self.next_seq()), s)

    def flag_or(self, flag: torch.Tensor, stream=None) -> None:
        s = ctypes.c_void_p((stream or torch.cuda.current_stream(self.device)).cuda_stream)
        native.call("flag_or", i32(self.rank), i32(self.size), self._flag_ptrs,
                    ctypes.c_uint(self.next_seq()), ptr(flag), s)

    def close(self) -> None:
        try:
            self.barrier()
        except Exception:
            pass
        for local, _, peer_ptrs in self._allocs:
            for r, a in enumerate(peer_ptrs):
                if r != self.rank:
                    self.lib.kfb_ipc_close(ctypes.c_void_p(a))
        self._allocs = []
        self._symm_keep = []
        if self._seq_counter is not None:
            self.lib.kfb_set_seq_base(None)
            self._seq_counter = None


def create_comm(params, device) -> Comm:
    """Pick the communicator from the environment torchrun / our launcher set up."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return SingleComm()
    dev = torch.device(device)
    impl = params.comm_impl
    if dev.type == "cpu":
        return TorchDistComm("gloo", dev)
    if impl == "nccl":
        return TorchDistComm("nccl", dev)
    if impl == "gloo":
        raise ValueError("--comm_impl=gloo needs --device=cpu")
    if impl == "auto":
        # the in-switch reduction pays from 4 ranks up; every rank takes the same decision and,
        # should the multicast bind fail anywhere, all of them fall back to the pull kernels
        if world >= 4 and os.environ.get("KFB_AUTO_NVLS", "1") != "0":
            try:
                return P2PComm(dev, nvls=True)
            except Exception as e:  # noqa: BLE001 - any bind / rendezvous failure
                import torch.distributed as dist
                print("comm_impl=auto: NVLS unavailable (%s); using the pull kernels"
                      % str(e).splitlines()[0][:160], flush=True)
                if not dist.is_initialized():
                    raise
        return P2PComm(dev, nvls=False)
    return P2PComm(dev, nvls=impl == "nvls")      # p2p, p2p_fused, nvls
