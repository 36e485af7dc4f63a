#!/usr/bin/env python
"""Multi-rank check of the fused NVLink kernels against a host simulation of all ranks.

Run under the launcher / torchrun, one rank per GPU:
    python -m kf_benchmarks_b200.parallel.launcher -np 2 -- tools/p2p_check.py
Every rank prints PASS/FAIL lines; rank 0 prints the summary. Exit code != 0 on failure.
"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from kf_benchmarks_b200.ops import native, optim_ref  # noqa: E402
from kf_benchmarks_b200.ops.native import f32, i32, i64, ptr  # noqa: E402
from kf_benchmarks_b200.parallel.comm import P2PComm  # noqa: E402
from kf_benchmarks_b200.parallel.strategy import Hyper  # noqa: E402


def make_hyper(opt, lr=0.1, scale=1.0, wd=1e-4, clip=0.0):
    h = Hyper()
    h.opt, h.lr, h.momentum, h.decay, h.epsilon = opt, lr, 0.9, 0.9, 1.0
    h.weight_decay, h.grad_scale, h.clip, h.adam_t = wd, scale, clip, 1.0
    return h


def rank_data(rank, n, seed):
    g = torch.Generator().manual_seed(seed * 100 + rank)
    return torch.randn(n, generator=g)


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    # KFB_P2P_NVLS=1: same checks through the NVLS (multimem) form of the all-reduce kernel
    nvls = os.environ.get("KFB_P2P_NVLS") == "1"
    comm = P2PComm(dev, nvls=nvls)
    ok = True
    n = int(os.environ.get("KFB_P2P_N", "3000064"))
    grad = comm.symm_alloc(n, torch.float32)
    w = comm.symm_alloc(n, torch.float32)
    red = comm.symm_alloc(n, torch.float32)
    w16 = torch.zeros(n, dtype=torch.float16, device=dev)
    m = torch.zeros(n, device=dev)
    s = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    gp, wp, rp, fp = (comm.ptr_array(grad), comm.ptr_array(w), comm.ptr_array(red),
                      comm.ptr_array(comm.flags))

    # KFB_P2P_V1=1: the first-generation entry points (kfb_allreduce_update / _nvls); default is
    # the second-generation kernel the engine uses (kfb_allreduce_update2), fp32 on the wire here
    v1 = os.environ.get("KFB_P2P_V1") == "1"

    def ar(mode, src, *tail):
        lo_, cnt_ = tail[0].value, tail[1].value
        if not v1 and lo_ % 8 == 0 and cnt_ % 8 == 0:
            native.call("allreduce_update2", i32(mode), i32(rank), i32(world),
                        None if nvls else (gp if src is grad else wp), None if nvls else rp,
                        ctypes.c_void_p(comm.mc_ptr(src)) if nvls else None,
                        ctypes.c_void_p(comm.mc_ptr(red)) if nvls else None, ptr(red), fp,
                        i32(0), None, None, *tail)
        elif nvls:
            native.call("allreduce_update_nvls", i32(mode), i32(rank), i32(world),
                        ctypes.c_void_p(comm.mc_ptr(src)), ctypes.c_void_p(comm.mc_ptr(red)),
                        ptr(red), fp, *tail)
        else:
            native.call("allreduce_update", i32(mode), i32(rank), i32(world),
                        gp if src is grad else wp, rp, fp, *tail)

    def check(name, got, want, tol=1e-5):
        nonlocal ok
        err = float((got.cpu() - want).abs().max())
        good = err <= tol * (1 + float(want.abs().max()))
        ok &= good
        print("[rank %d] %s %s (max err %.3e)" % (rank, "PASS" if good else "FAIL", name, err),
              flush=True)

    # ---- 1. gradient all-reduce (sum) + Nesterov momentum, two steps, several buckets ----
    w0 = rank_data(99, n, 1)  # identical initial weights
    w.copy_(w0)
    ref_w, ref_m = w0.clone(), torch.zeros(n)
    for step in range(2):
        grad.copy_(rank_data(rank, n, 10 + step))
        comm.barrier()
        h = make_hyper(1, scale=1.0 / 128)
        for lo, hi in [(0, 1_000_064), (1_000_064, 2_500_096), (2_500_096, n)] if n == 3000064 else [(0, n)]:
            ar(0, grad, i64(lo),
                        i64(hi - lo), ctypes.c_uint(comm.next_seq()), ptr(w), ptr(grad), ptr(m),
                        None, ptr(w16), ctypes.byref(h), None, None, f32(0.0), i32(0), s)
        torch.cuda.synchronize()
        gsum = sum(rank_data(r, n, 10 + step) for r in range(world))
        optim_ref.apply_update(h, ref_w, gsum, ref_m, None)
        check("allreduce_update step %d (weights)" % step, w, ref_w, 2e-5)
        check("allreduce_update step %d (fp16 shadow)" % step, w16.float(), ref_w, 2e-3)
    # ---- 1b. compact gradient transfer: fp16 on the wire, fp32 sum, two buckets ----
    if not v1:
        g16 = comm.symm_alloc(n, torch.float16)
        r16 = comm.symm_alloc(n, torch.float16)
        g16p, r16p = comm.ptr_array(g16), comm.ptr_array(r16)
        w.copy_(w0)
        m.zero_()
        grad.copy_(rank_data(rank, n, 15))
        comm.barrier()
        h = make_hyper(1, scale=1.0 / 128)
        cut = (n // 3) // 128 * 128
        for lo, hi in [(0, cut), (cut, n)]:
            native.call("allreduce_update2", i32(0), i32(rank), i32(world),
                        None if nvls else g16p, None if nvls else r16p,
                        ctypes.c_void_p(comm.mc_ptr(g16)) if nvls else None,
                        ctypes.c_void_p(comm.mc_ptr(r16)) if nvls else None, ptr(r16), fp, i32(1),
                        ptr(grad), ptr(g16), i64(lo), i64(hi - lo),
                        ctypes.c_uint(comm.next_seq()), ptr(w), ptr(grad), ptr(m), None, ptr(w16),
                        ctypes.byref(h), None, None, f32(0.0), i32(0), s)
        torch.cuda.synchronize()
        gsum = sum(rank_data(r, n, 15).half().float() for r in range(world)).half().float()
        cw, cm = w0.clone(), torch.zeros(n)
        optim_ref.apply_update(h, cw, gsum, cm, None)
        # the switch may round the fp32 sum once more than the host model does: 1 fp16 ulp of
        # the summed gradient times lr * scale
        check("allreduce_update2 compact (fp16 wire, fp32 sum)", w, cw, 2e-5)
        all_w = comm.allgather_object(w.cpu()[:: max(n // 65536, 1)].clone())
        same = all(torch.equal(all_w[0], x) for x in all_w)
        ok &= same
        print("[rank %d] %s compact: replicas bit-identical" % (rank, "PASS" if same else "FAIL"),
              flush=True)
    # ---- 2. SMA: average of the weight arenas, blend, local SGD ----
    wr = rank_data(rank, n, 20)
    w.copy_(wr)
    grad.copy_(rank_data(rank, n, 21))
    comm.barrier()
    h = make_hyper(0, scale=1.0, wd=0.0)
    ar(1, w, i64(0), i64(n),
                ctypes.c_uint(comm.next_seq()), ptr(w), ptr(grad), None, None, ptr(w16),
                ctypes.byref(h), None, None, f32(0.1), i32(0), s)
    torch.cuda.synchronize()
    avg = sum(rank_data(r, n, 20) for r in range(world)) / world
    want = 0.9 * wr + 0.1 * avg
    want = want - 0.1 * rank_data(rank, n, 21)
    check("sma_update", w, want, 2e-5)
    # ---- 3. broadcast from rank 0 ----
    w.copy_(rank_data(rank, n, 30))
    comm.barrier()
    native.call("broadcast_pull", i32(rank), i32(world), wp, fp, i32(0), i64(0), i64(n),
                ctypes.c_uint(comm.next_seq()), ptr(w), ptr(w16), i32(0), s)
    torch.cuda.synchronize()
    check("broadcast_pull", w, rank_data(0, n, 30), 0.0)
    # ---- 4. barrier + flag OR ----
    comm.device_barrier()
    flag = torch.tensor([1 if rank == world - 1 else 0], dtype=torch.int32, device=dev)
    comm.flag_or(flag)
    torch.cuda.synchronize()
    good = int(flag.item()) == 1
    ok &= good
    print("[rank %d] %s flag_or" % (rank, "PASS" if good else "FAIL"), flush=True)
    # ---- 5. pair averaging: publish, then blend with the neighbour's published model ----
    pub = comm.symm_alloc(3 * n, torch.float32)
    ver = comm.symm_alloc(64, torch.int32)
    torn = torch.zeros(1, dtype=torch.int32, device=dev)
    pull = torch.empty(n, device=dev)
    psync = torch.zeros(4, dtype=torch.int32, device=dev)
    peer = (rank + 1) % world
    h = make_hyper(0, scale=1.0, wd=0.0)
    w.copy_(rank_data(rank, n, 40))
    grad.zero_()
    comm.barrier()
    # step 1: nothing published yet -> plain local step (zero grad) + publish version 1
    native.call("pair_average_update", ctypes.c_void_p(comm.peers(pub)[peer]),
                ctypes.c_void_p(comm.peers(ver)[peer]), ptr(pub), ptr(ver), i64(n), i64(n),
                ptr(w), ptr(grad), None, None, ptr(w16), ctypes.byref(h), None, ctypes.c_uint(1),
                i32(1), ptr(torn), ptr(pull), ptr(psync), i32(0), s)
    torch.cuda.synchronize()
    comm.barrier()
    # step 2: every rank now sees its peer's version 1
    native.call("pair_average_update", ctypes.c_void_p(comm.peers(pub)[peer]),
                ctypes.c_void_p(comm.peers(ver)[peer]), ptr(pub), ptr(ver), i64(n), i64(n),
                ptr(w), ptr(grad), None, None, ptr(w16), ctypes.byref(h), None, ctypes.c_uint(2),
                i32(1), ptr(torn), ptr(pull), ptr(psync), i32(0), s)
    torch.cuda.synchronize()
    want = 0.5 * (rank_data(rank, n, 40) + rank_data(peer, n, 40))
    check("pair_average_update", w, want, 1e-6)
    good = int(torn.item()) == 0
    ok &= good
    comm.barrier()
    # ---- 6. all-reduce micro timing ----
    for blocks in [int(b) for b in os.environ.get("KFB_AR_BLOCKS_SWEEP", "0").split(",")]:
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h = make_hyper(1, scale=1.0)
        comm.barrier()
        e0.record()
        for _ in range(reps):
            ar(0, grad, i64(0),
                        i64(n), ctypes.c_uint(comm.next_seq()), ptr(w), ptr(grad), ptr(m), None,
                        ptr(w16), ctypes.byref(h), None, None, f32(0.0), i32(blocks), s)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps
        us = comm.max_over_ranks(us)
        if rank == 0:
            link = 2.0 * (world - 1) / world * n * 4 / (us * 1e-6) / 1e9
            print("allreduce_update %.1f MB fp32 on %d ranks, %d blocks: %.1f us  (%.1f GB/s "
                  "per-GPU link traffic)" % (n * 4 / 1e6, world, blocks, us, link), flush=True)
    # ---- 7. protocol stress: many back-to-back calls on random bucket partitions; replicas must
    #         stay bit-identical and match the host model ----
    iters = int(os.environ.get("KFB_P2P_STRESS", "200"))
    import random
    rng = random.Random(4242)  # same partition sequence on every rank
    m.zero_()
    w.copy_(rank_data(99, n, 50))
    grad.copy_(rank_data(rank, n, 51))
    comm.barrier()
    h = make_hyper(0, lr=1e-3, scale=1.0, wd=0.0)
    for it in range(iters):
        cuts = sorted({0, n} | {rng.randrange(0, n // 128) * 128 for _ in range(rng.randint(0, 4))})
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            ar(0, grad, i64(lo),
                        i64(hi - lo), ctypes.c_uint(comm.next_seq()), ptr(w), ptr(grad), None,
                        None, ptr(w16), ctypes.byref(h), None, None, f32(0.0),
                        i32(rng.choice([0, 8, 33, 148])), s)
    torch.cuda.synchronize()
    gsum = sum(rank_data(r, n, 51) for r in range(world))
    want = rank_data(99, n, 50) - iters * 1e-3 * gsum
    check("stress: %d iterations" % iters, w, want, 2e-4)
    mine = w.cpu()
    all_w = comm.allgather_object(mine[:: max(n // 65536, 1)].clone())
    same = all(torch.equal(all_w[0], x) for x in all_w)
    ok &= same
    print("[rank %d] %s stress: replicas bit-identical" % (rank, "PASS" if same else "FAIL"),
          flush=True)
    # ---- 8. device-resident sequence offset (multi-rank CUDA graphs): the same by-value
    #         numbers are presented three times, the device counter moves them forward ----
    if os.environ.get("KFB_P2P_DEVSEQ") == "1":
        counter = comm.enable_device_seq()
        w.copy_(rank_data(99, n, 60))
        grad.copy_(rank_data(rank, n, 61))
        comm.barrier()
        h = make_hyper(0, lr=1e-3, scale=1.0, wd=0.0)
        seq0 = comm.seq
        fixed = [ctypes.c_uint(comm.next_seq()) for _ in range(2)]   # what a capture would bake in
        k = comm.seq - seq0
        comm.seq = seq0
        half = (n // 2) // 128 * 128
        for _ in range(3):                                           # three "replays"
            ar(0, grad, i64(0), i64(half), fixed[0], ptr(w), ptr(grad), None, None, ptr(w16),
               ctypes.byref(h), None, None, f32(0.0), i32(0), s)
            ar(0, grad, i64(half), i64(n - half), fixed[1], ptr(w), ptr(grad), None, None,
               ptr(w16), ctypes.byref(h), None, None, f32(0.0), i32(0), s)
            native.call("seq_advance", ptr(counter), ctypes.c_uint(k), s)
            comm.graph_replayed(k)
        comm.device_barrier()                                        # an eager call afterwards
        torch.cuda.synchronize()
        gsum = sum(rank_data(r, n, 61) for r in range(world))
        check("device sequence offset: 3 replays", w, rank_data(99, n, 60) - 3 * 1e-3 * gsum, 2e-5)
        check("device counter value", counter.float(), torch.tensor([3.0 * k]), 0.0)
    allok = all(comm.allgather_object(bool(ok)))
    if rank == 0:
        print("P2P_CHECK %s" % ("ALL PASS" if allok else "FAILED"), flush=True)
    comm.close()
    return 0 if allok else 1


if __name__ == "__main__":
    sys.exit(main())
