// Gradient-path kernels for sm_100a: fused optimizers on the flat parameter arena and the
// peer-to-peer (NVLink 5 / NVSwitch) collectives that replace the reference's gradient
// replication stack:
//
//   * kfb_optimizer_step       - unscale + weight decay + clip + finite-skip + SGD / Nesterov
//                                momentum / RMSProp / Adam on the fp32 master, fp16 re-emit
//                                (T/benchmark_cnn.py:1172-1190, 2787-2808, 3110-3120)
//   * kfb_allreduce_update     - ONE kernel per bucket: two-shot all-reduce over mapped peer
//                                arenas (reduce-scatter by P2P loads, all-gather by P2P
//                                stores) fused with the optimizer epilogue above
//                                (replaces T/batch_allreduce.py + T/allreduce.py:280-324,
//                                hvd.allreduce T/benchmark_cnn.py:3122-3130, KungFu
//                                SynchronousSGDOptimizer T/benchmark_cnn.py:1193-1195)
//   * kfb_sma_update           - KungFu SynchronousAveraging: all-reduce-average the *model*,
//                                blend, then apply the local gradient (:1199-1201)
//   * kfb_pair_average_update  - KungFu PairAveraging / AD-PSGD: pull one peer's published
//                                model straight from its HBM, blend 1/2, apply the local
//                                gradient, publish (:1196-1198)
//   * kfb_broadcast_pull       - initial variable broadcast from rank 0 (T/variable_mgr.py:342-356)
//   * kfb_device_barrier       - flag barrier in symmetric memory (kungfu.run_barrier)
//   * kfb_check_finite / kfb_loss_scale_update - auto loss-scale state machine
//                                (T/variable_mgr_util.py:51-139)
//
// Synchronisation protocol: every rank owns a flag page in symmetric memory. Flags only ever
// grow (per-call sequence numbers), are written with st.release.sys after a system fence and
// polled with ld.acquire.sys, so there is no reset race. Waits are bounded (trap after 20 s).
#include <stdlib.h>

#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

constexpr int kMaxRanks = 8;
// flag page layout (uint32 words): [kind][src_rank], kinds below; one extra block counter row
enum FlagKind : int { FLAG_READY = 0, FLAG_REDUCED = 1, FLAG_DONE = 2, FLAG_BARRIER = 3,
                      FLAG_KINDS = 4 };
constexpr int kCounterWord = FLAG_KINDS * kMaxRanks;  // grid arrival counters live after flags

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

struct HyperParams {
  int opt;             // 0 sgd, 1 momentum (Nesterov, TF ApplyMomentum), 2 rmsprop, 3 adam
  float lr;
  float momentum;      // momentum / rmsprop momentum / adam beta1
  float decay;         // rmsprop decay / adam beta2
  float epsilon;
  float weight_decay;  // L2 coefficient: g += wd * w
  float grad_scale;    // multiplies the (reduced) gradient: 1/loss_scale, 1/N for means
  float clip;          // elementwise clip (<= 0: off)
  float adam_t;        // step count for Adam bias correction
  // optional: one byte per 128-element arena block (slots are 128-aligned), 0 = this block is
  // excluded from the L2 term (the SSD model decays no batch-norm variable,
  // T/models/ssd_model.py:91-94); indexed like the kernel's element index
  const unsigned char* decay_mask;
};

__device__ __forceinline__ float decay_at(const HyperParams& h, long long i) {
  return (h.decay_mask != nullptr && h.decay_mask[i >> 7] == 0) ? 0.f : h.weight_decay;
}

// dyn (optional, device memory): {lr, grad_scale, adam_t, unused}. Lets a captured CUDA graph
// see per-step values.
__device__ __forceinline__ void resolve_dyn(HyperParams& h, const float* dyn) {
  if (dyn != nullptr) {
    h.lr = dyn[0];
    h.grad_scale = dyn[1];
    h.adam_t = dyn[2];
  }
}

__device__ __forceinline__ float apply_update(const HyperParams& h, float w, float g_raw, float* m,
                                              float* v, long long i) {
  float g = g_raw * h.grad_scale + decay_at(h, i) * w;
  if (h.clip > 0.f) g = fminf(fmaxf(g, -h.clip), h.clip);
  switch (h.opt) {
    case 0:
      return w - h.lr * g;
    case 1: {
      const float acc = m[i] * h.momentum + g;
      m[i] = acc;
      return w - h.lr * (g + h.momentum * acc);
    }
    case 2: {
      const float ms = h.decay * v[i] + (1.f - h.decay) * g * g;
      v[i] = ms;
      const float mom = h.momentum * m[i] + h.lr * g * rsqrtf(ms + h.epsilon);
      m[i] = mom;
      return w - mom;
    }
    default: {
      const float b1 = h.momentum, b2 = h.decay;
      const float mm = b1 * m[i] + (1.f - b1) * g;
      const float vv = b2 * v[i] + (1.f - b2) * g * g;
      m[i] = mm;
      v[i] = vv;
      const float lr_t = h.lr * sqrtf(1.f - __powf(b2, h.adam_t)) / (1.f - __powf(b1, h.adam_t));
      return w - lr_t * mm / (sqrtf(vv) + h.epsilon);
    }
  }
}

// Four consecutive elements at i (i % 4 == 0, all arenas 16-byte aligned).
__device__ __forceinline__ void apply_update4(const HyperParams& h, float* __restrict__ w,
                                              const float4 g4, float* m, float* v,
                                              __half* __restrict__ w16, long long i,
                                              const float4* blend_with, float blend_alpha) {
  float4 w4 = *reinterpret_cast<const float4*>(w + i);
  if (blend_with != nullptr) {
    const float a = blend_alpha;
    w4.x = (1.f - a) * w4.x + a * blend_with->x;
    w4.y = (1.f - a) * w4.y + a * blend_with->y;
    w4.z = (1.f - a) * w4.z + a * blend_with->z;
    w4.w = (1.f - a) * w4.w + a * blend_with->w;
  }
  float wv[4] = {w4.x, w4.y, w4.z, w4.w};
  const float gv[4] = {g4.x, g4.y, g4.z, g4.w};
  float mv[4] = {0.f, 0.f, 0.f, 0.f}, vv[4] = {0.f, 0.f, 0.f, 0.f};
  if (h.opt != 0) {
    const float4 t = *reinterpret_cast<const float4*>(m + i);
    mv[0] = t.x; mv[1] = t.y; mv[2] = t.z; mv[3] = t.w;
  }
  if (h.opt >= 2) {
    const float4 t = *reinterpret_cast<const float4*>(v + i);
    vv[0] = t.x; vv[1] = t.y; vv[2] = t.z; vv[3] = t.w;
  }
  const float wd = decay_at(h, i);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float g = gv[k] * h.grad_scale + wd * wv[k];
    if (h.clip > 0.f) g = fminf(fmaxf(g, -h.clip), h.clip);
    if (h.opt == 0) {
      wv[k] -= h.lr * g;
    } else if (h.opt == 1) {
      mv[k] = mv[k] * h.momentum + g;
      wv[k] -= h.lr * (g + h.momentum * mv[k]);
    } else if (h.opt == 2) {
      vv[k] = h.decay * vv[k] + (1.f - h.decay) * g * g;
      mv[k] = h.momentum * mv[k] + h.lr * g * rsqrtf(vv[k] + h.epsilon);
      wv[k] -= mv[k];
    } else {
      const float b1 = h.momentum, b2 = h.decay;
      mv[k] = b1 * mv[k] + (1.f - b1) * g;
      vv[k] = b2 * vv[k] + (1.f - b2) * g * g;
      const float lr_t = h.lr * sqrtf(1.f - __powf(b2, h.adam_t)) / (1.f - __powf(b1, h.adam_t));
      wv[k] -= lr_t * mv[k] / (sqrtf(vv[k]) + h.epsilon);
    }
  }
  *reinterpret_cast<float4*>(w + i) = make_float4(wv[0], wv[1], wv[2], wv[3]);
  if (h.opt != 0) *reinterpret_cast<float4*>(m + i) = make_float4(mv[0], mv[1], mv[2], mv[3]);
  if (h.opt >= 2) *reinterpret_cast<float4*>(v + i) = make_float4(vv[0], vv[1], vv[2], vv[3]);
  if (w16 != nullptr) {
    uint2 u;
    u.x = pack_half2(wv[0], wv[1]);
    u.y = pack_half2(wv[2], wv[3]);
    *reinterpret_cast<uint2*>(w16 + i) = u;
  }
}

// ---------------------------------------------------------------------------
// Local fused optimizer over [0, n)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    optimizer_step_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m,
                          float* __restrict__ v, __half* __restrict__ w16, long long n,
                          HyperParams h, const float* __restrict__ dyn,
                          const int* __restrict__ skip_flag) {
  if (skip_flag != nullptr && *skip_flag != 0) return;
  resolve_dyn(h, dyn);
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long n4 = n & ~3ll;
  for (long long i = tid * 4; i < n4; i += nthreads * 4)
    apply_update4(h, w, *reinterpret_cast<const float4*>(g + i), m, v, w16, i, nullptr, 0.f);
  for (long long i = n4 + tid; i < n; i += nthreads) {
    const float nw = apply_update(h, w[i], g[i], m, v, i);
    w[i] = nw;
    if (w16 != nullptr) w16[i] = __float2half_rn(nw);
  }
}

// SMA applied from a local copy of the averaged model: w <- (1-a) w + a avg, then the local
// gradient step (the average was all-reduced on the exchange stream while backward ran).
__global__ void __launch_bounds__(256)
    optimizer_step_blend_kernel(float* __restrict__ w, const float* __restrict__ avg,
                                const float* __restrict__ g, float* __restrict__ m,
                                float* __restrict__ v, __half* __restrict__ w16, long long n,
                                HyperParams h, const float* __restrict__ dyn, float alpha) {
  resolve_dyn(h, dyn);
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long n4 = n & ~3ll;
  for (long long i = tid * 4; i < n4; i += nthreads * 4) {
    const float4 a4 = *reinterpret_cast<const float4*>(avg + i);
    apply_update4(h, w, *reinterpret_cast<const float4*>(g + i), m, v, w16, i, &a4, alpha);
  }
  for (long long i = n4 + tid; i < n; i += nthreads) {
    const float blended = (1.f - alpha) * w[i] + alpha * avg[i];
    const float nw = apply_update(h, blended, g[i], m, v, i);
    w[i] = nw;
    if (w16 != nullptr) w16[i] = __float2half_rn(nw);
  }
}

// flag[0] |= any non-finite in g[0..n)
__global__ void __launch_bounds__(256)
    check_finite_kernel(const float* __restrict__ g, long long n, int* __restrict__ flag) {
  int bad = 0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float x = g[i];
    bad |= !(fabsf(x) <= 3.4028234e38f);
  }
  if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// state = {loss_scale, normal_steps, skipped steps so far, -}; dyn[1] (grad_scale) is refreshed
// to base_grad_scale/scale and the flag is cleared for the next step.
// Skip flag set: halve, reset counter. Else count up to inc_every clean steps, then double
// (same transitions as T/variable_mgr_util.py:51-88).
__global__ void loss_scale_update_kernel(float* state, int* flag, int inc_every, float* dyn,
                                         float base_grad_scale) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  float scale = state[0];
  float steps = state[1];
  if (*flag != 0) {
    scale *= 0.5f;
    steps = 0.f;
    state[2] += 1.f;
    *flag = 0;
  } else if (steps < static_cast<float>(inc_every)) {
    steps += 1.f;
  } else {
    scale *= 2.f;
    steps = 0.f;
  }
  state[0] = scale;
  state[1] = steps;
  if (dyn != nullptr) dyn[1] = base_grad_scale / scale;
}

// dyn = {lr, grad_scale, adam_t, -}: the entries selected by `mask` (bit i = entry i) are set
// from the arguments. Written by a kernel so that neither a pinned staging buffer (which a host
// running ahead of the device could overwrite before its copy executes) nor a synchronisation
// is needed, and so that a device-maintained entry (grad_scale under automatic loss scaling)
// can be left alone.
__global__ void set_hyper_dyn_kernel(float* dyn, float lr, float grad_scale, float adam_t,
                                     int mask) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  if (mask & 1) dyn[0] = lr;
  if (mask & 2) dyn[1] = grad_scale;
  if (mask & 4) dyn[2] = adam_t;
}

// ---------------------------------------------------------------------------
// Peer context (passed by value)
// ---------------------------------------------------------------------------
struct PeerCtx {
  int rank, nranks;
  float* src[kMaxRanks];      // per-rank source arena (gradients, or weights for SMA)
  float* reduced[kMaxRanks];  // per-rank landing buffer for reduced shards
  uint32_t* flags[kMaxRanks]; // per-rank flag page
  // Optional device-resident offset added to every call's sequence number (kfb_set_seq_base):
  // a captured CUDA graph replays the same by-value `seq`, the counter moves it forward.
  const uint32_t* seq_base;
};

__device__ __forceinline__ uint32_t seq_offset(const PeerCtx& c) {
  return c.seq_base != nullptr ? *reinterpret_cast<const volatile uint32_t*>(c.seq_base) : 0u;
}

__device__ __forceinline__ void wait_flags(const PeerCtx& c, int kind, uint32_t seq) {
  // thread t < nranks polls the flag written by rank t into MY page
  if (threadIdx.x < c.nranks) {
    const uint32_t* f = c.flags[c.rank] + kind * kMaxRanks + threadIdx.x;
    const uint64_t t0 = globaltimer_ns();
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(f) - seq) < 0) {
      if ((++spins & 0xfff) == 0 && globaltimer_ns() - t0 > 20000000000ull) {
        printf("kfb: rank %d timed out waiting for flag kind %d from rank %d (want %u have %u)\n",
               c.rank, kind, (int)threadIdx.x, seq, ld_acquire_sys(f));
        __trap();
      }
    }
  }
  __syncthreads();
}

// All threads have issued their stores; publish `seq` of `kind` to every rank once the whole
// grid has arrived. counter_slot selects one of the arrival counters in my flag page.
__device__ __forceinline__ void grid_arrive_and_signal(const PeerCtx& c, int kind, uint32_t seq,
                                                       int counter_slot) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* counter = c.flags[c.rank] + kCounterWord + counter_slot;
    const uint32_t prev = atomicAdd(counter, 1u);
    if (prev == gridDim.x - 1) {
      *counter = 0;  // all blocks have arrived; nobody touches it again in this launch
      __threadfence_system();
      for (int r = 0; r < c.nranks; ++r)
        st_release_sys(c.flags[r] + kind * kMaxRanks + c.rank, seq);
    }
  }
}

struct PushDst {
  float* base[kMaxRanks];
};

// MODE 0: gradient all-reduce + optimizer. MODE 1: SMA (source = weights).
template <int MODE>
__global__ void __launch_bounds__(512)
    allreduce_update_kernel(PeerCtx c, long long offset, long long count, uint32_t seq,
                            float* __restrict__ w, const float* __restrict__ g_local,
                            float* __restrict__ m, float* __restrict__ v,
                            __half* __restrict__ w16, HyperParams h,
                            const float* __restrict__ dyn, const int* __restrict__ skip_flag,
                            float sma_alpha, int reduce_only, const float* __restrict__ stage,
                            long long stage_stride) {
  resolve_dyn(h, dyn);
  seq += seq_offset(c);
  const int n = c.nranks;
  // Reduce-scatter sources: the peers' arenas (pull over NVLink), or - when the weight-gradient
  // kernels already pushed their tiles here (kfb_*_rs) - this rank's n local staging slots.
  const float* srcp[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    srcp[r] = stage != nullptr ? stage + static_cast<long long>(r) * stage_stride
                               : (r < n ? c.src[r] : nullptr);
  // --- announce: my source data for [offset, offset+count) is final ---
  if (blockIdx.x == 0 && threadIdx.x < n)
    st_release_sys(c.flags[threadIdx.x] + FLAG_READY * kMaxRanks + c.rank, seq);
  wait_flags(c, FLAG_READY, seq);

  // --- phase 1: reduce my shard from every peer, push the result to every rank ---
  const long long per = ((count + n - 1) / n + 3) & ~3ll;  // shard length, multiple of 4
  const long long s_begin = offset + per * c.rank;
  long long s_end = s_begin + per;
  if (s_end > offset + count) s_end = offset + count;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const float inv_n = 1.f / static_cast<float>(n);
  // Two float4 per peer per thread are in flight before the first add (2 * n loads).
  constexpr int kU = 2;
  for (long long i0 = s_begin + tid * 4; i0 < s_end; i0 += nthreads * 4 * kU) {
    float4 acc[kU];
    bool ok[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      ok[u] = i + 4 <= s_end;
      acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r) {
      if (r < n) {
        float4 x[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u)
          if (ok[u]) x[u] = *reinterpret_cast<const float4*>(srcp[r] + i0 + u * nthreads * 4);
#pragma unroll
        for (int u = 0; u < kU; ++u)
          if (ok[u]) {
            acc[u].x += x[u].x;
            acc[u].y += x[u].y;
            acc[u].z += x[u].z;
            acc[u].w += x[u].w;
          }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      if (ok[u]) {
        if (MODE == 1) {
          acc[u].x *= inv_n; acc[u].y *= inv_n; acc[u].z *= inv_n; acc[u].w *= inv_n;
        }
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < n) *reinterpret_cast<float4*>(c.reduced[r] + i) = acc[u];
      } else if (i < s_end) {
        for (long long j = i; j < s_end; ++j) {
          float a1 = 0.f;
          for (int r = 0; r < n; ++r) a1 += srcp[r][j];
          if (MODE == 1) a1 *= inv_n;
          for (int r = 0; r < n; ++r) c.reduced[r][j] = a1;
        }
      }
    }
  }
  grid_arrive_and_signal(c, FLAG_REDUCED, seq, 0);
  wait_flags(c, FLAG_REDUCED, seq);

  // --- phase 2: every rank applies the identical update to its full replica ---
  const bool skip = (skip_flag != nullptr && *skip_flag != 0) || reduce_only;
  if (!skip) {
    const float* red = c.reduced[c.rank];
    const long long end = offset + count;
    const long long end4 = offset + (count & ~3ll);  // offset is a multiple of 4
    for (long long i = offset + tid * 4; i < end4; i += nthreads * 4) {
      const float4 r4 = *reinterpret_cast<const float4*>(red + i);
      if (MODE == 0) {
        apply_update4(h, w, r4, m, v, w16, i, nullptr, 0.f);
      } else {
        // SMA: w <- (1-a) w + a * avg(w); then the local gradient step
        apply_update4(h, w, *reinterpret_cast<const float4*>(g_local + i), m, v, w16, i, &r4,
                      sma_alpha);
      }
    }
    for (long long i = end4 + tid; i < end; i += nthreads) {
      float nw;
      if (MODE == 0) {
        nw = apply_update(h, w[i], red[i], m, v, i);
      } else {
        const float blended = (1.f - sma_alpha) * w[i] + sma_alpha * red[i];
        nw = apply_update(h, blended, g_local[i], m, v, i);
      }
      w[i] = nw;
      if (w16 != nullptr) w16[i] = __float2half_rn(nw);
    }
  }
  // No trailing barrier: a peer can only touch this range again (write my landing buffer, read
  // my gradients or weights) after it has seen my READY flag of the NEXT call on this range,
  // and that flag is sent by a later launch on this stream, i.e. after this kernel has ended.
}

// ---------------------------------------------------------------------------
// NVLS variant (--comm_impl=nvls): the same two-shot all-reduce + update, with the reduction
// and the broadcast done by the NVSwitch. The source arena and the landing buffer are bound to
// a multicast object (parallel/comm.py, torch symmetric memory); one
// multimem.ld_reduce on the multicast address returns the sum over all ranks' copies of that
// address (added inside the switch), one multimem.st writes every rank's copy. Per bucket of
// S bytes each GPU's NVLink port carries about S in and S out (S/N reduced shard in, the
// other ranks' broadcasts in, its own copies read out by the switch, S/N broadcast out) -
// half of the pull kernel's 2 (N-1)/N S per direction - and the N loads + N stores per element
// become one instruction each. The flag protocol is the one of allreduce_update_kernel.
// The switch's summation order is fixed per address but is not the rank order of the pull
// kernel, so results match it to rounding, not bit for bit; replicas stay bit-identical to
// each other because every element is reduced once and broadcast.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ float multimem_ld_reduce_add1(const float* mc) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st4(float* mc, const float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void multimem_st1(float* mc, const float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}

// c.src is unused (the switch reads every rank's copy through src_mc); c.reduced[c.rank] is
// this rank's landing buffer, reduced_mc its multicast alias.
template <int MODE>
__global__ void __launch_bounds__(512)
    allreduce_update_nvls_kernel(PeerCtx c, const float* __restrict__ src_mc,
                                 float* __restrict__ reduced_mc, long long offset,
                                 long long count, uint32_t seq, float* __restrict__ w,
                                 const float* __restrict__ g_local, float* __restrict__ m,
                                 float* __restrict__ v, __half* __restrict__ w16, HyperParams h,
                                 const float* __restrict__ dyn,
                                 const int* __restrict__ skip_flag, float sma_alpha,
                                 int reduce_only) {
  resolve_dyn(h, dyn);
  seq += seq_offset(c);
  const int n = c.nranks;
  // --- announce: my source data for [offset, offset+count) is final ---
  if (blockIdx.x == 0 && threadIdx.x < n)
    st_release_sys(c.flags[threadIdx.x] + FLAG_READY * kMaxRanks + c.rank, seq);
  wait_flags(c, FLAG_READY, seq);

  // --- phase 1: in-switch reduction of my shard, in-switch broadcast of the result ---
  const long long per = ((count + n - 1) / n + 3) & ~3ll;
  const long long s_begin = offset + per * c.rank;
  long long s_end = s_begin + per;
  if (s_end > offset + count) s_end = offset + count;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const float inv_n = 1.f / static_cast<float>(n);
  constexpr int kU = 4;  // reductions in flight per thread before the first store
  for (long long i0 = s_begin + tid * 4; i0 < s_end; i0 += nthreads * 4 * kU) {
    float4 acc[kU];
    bool ok[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      ok[u] = i + 4 <= s_end;
      if (ok[u]) acc[u] = multimem_ld_reduce_add4(src_mc + i);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      if (ok[u]) {
        if (MODE == 1) {
          acc[u].x *= inv_n; acc[u].y *= inv_n; acc[u].z *= inv_n; acc[u].w *= inv_n;
        }
        multimem_st4(reduced_mc + i, acc[u]);
      } else if (i < s_end) {
        for (long long j = i; j < s_end; ++j) {  // at most 3 elements, end of the last shard
          float a1 = multimem_ld_reduce_add1(src_mc + j);
          if (MODE == 1) a1 *= inv_n;
          multimem_st1(reduced_mc + j, a1);
        }
      }
    }
  }
  grid_arrive_and_signal(c, FLAG_REDUCED, seq, 0);
  wait_flags(c, FLAG_REDUCED, seq);

  // --- phase 2: every rank applies the identical update to its full replica ---
  const bool skip = (skip_flag != nullptr && *skip_flag != 0) || reduce_only;
  if (!skip) {
    const float* red = c.reduced[c.rank];
    const long long end = offset + count;
    const long long end4 = offset + (count & ~3ll);
    for (long long i = offset + tid * 4; i < end4; i += nthreads * 4) {
      const float4 r4 = *reinterpret_cast<const float4*>(red + i);
      if (MODE == 0) {
        apply_update4(h, w, r4, m, v, w16, i, nullptr, 0.f);
      } else {
        apply_update4(h, w, *reinterpret_cast<const float4*>(g_local + i), m, v, w16, i, &r4,
                      sma_alpha);
      }
    }
    for (long long i = end4 + tid; i < end; i += nthreads) {
      float nw;
      if (MODE == 0) {
        nw = apply_update(h, w[i], red[i], m, v, i);
      } else {
        const float blended = (1.f - sma_alpha) * w[i] + sma_alpha * red[i];
        nw = apply_update(h, blended, g_local[i], m, v, i);
      }
      w[i] = nw;
      if (w16 != nullptr) w16[i] = __float2half_rn(nw);
    }
  }
}

// ---------------------------------------------------------------------------
// Second generation of the fused all-reduce + update (kfb_allreduce_update2). Same two-shot
// protocol and flags as above; what changed, and why (profiles/README.md, all_reduce_benchmark):
//   * phase 2 (the optimizer over the whole bucket) keeps 4 x 16 bytes per array in flight per
//     thread - with one vector per thread the few CTAs this kernel is allowed (the backward pass
//     wants the SMs) reached ~2.3 TB/s and phase 2 was half of the kernel's time;
//   * phase 1 keeps 4 vectors per peer in flight;
//   * HALF = compact gradient transfer (--compact_gradient_transfer, T/batch_allreduce.py:429-453,
//     T/variable_mgr.py:311-313): the sources are fp16 copies of the gradients and the reduced
//     shard travels as fp16; the sum itself is taken in fp32 (by the owner, or inside the switch
//     with .acc::f32), which the reference's fp16 NCCL all-reduce does not do;
//   * NVLS = reduction and broadcast done by the NVSwitch on multicast addresses.
// Requires offset and count to be multiples of 8 elements (arena slots are 128-aligned).
// ---------------------------------------------------------------------------
struct Ar2Args {
  const void* src[kMaxRanks];  // pull: per-rank source arena (fp32 or fp16), element 0
  void* red[kMaxRanks];        // pull: per-rank landing buffer, element 0
  const void* src_mc;          // NVLS: multicast alias of the source arena
  void* red_mc;                // NVLS: multicast alias of the landing buffer
  const void* red_local;       // this rank's landing buffer
  long long offset, count;
  uint32_t seq;
  float* w;
  const float* g_local;
  float* m;
  float* v;
  __half* w16;
  const float* dyn;
  const int* skip_flag;
  float sma_alpha;
  int reduce_only;
};

template <bool HALF>
__device__ __forceinline__ void vec_unpack(const uint4& q, float* f) {
  if (HALF) {
    const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __half22float2(h[i]);
      f[2 * i] = t.x;
      f[2 * i + 1] = t.y;
    }
  } else {
    f[0] = __uint_as_float(q.x);
    f[1] = __uint_as_float(q.y);
    f[2] = __uint_as_float(q.z);
    f[3] = __uint_as_float(q.w);
  }
}
template <bool HALF>
__device__ __forceinline__ uint4 vec_pack(const float* f) {
  uint4 q;
  if (HALF) {
    q.x = pack_half2(f[0], f[1]);
    q.y = pack_half2(f[2], f[3]);
    q.z = pack_half2(f[4], f[5]);
    q.w = pack_half2(f[6], f[7]);
  } else {
    q.x = __float_as_uint(f[0]);
    q.y = __float_as_uint(f[1]);
    q.z = __float_as_uint(f[2]);
    q.w = __float_as_uint(f[3]);
  }
  return q;
}
template <bool HALF>
__device__ __forceinline__ uint4 multimem_ld_reduce_vec(const void* mc) {
  uint4 q;
  if (HALF) {
    asm volatile(
        "multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
        : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
        : "l"(mc)
        : "memory");
  } else {
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w)
                 : "l"(mc)
                 : "memory");
  }
  return q;
}
template <bool HALF>
__device__ __forceinline__ void multimem_st_vec(void* mc, const uint4& q) {
  if (HALF) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(mc),
                 "r"(q.x), "r"(q.y), "r"(q.z), "r"(q.w)
                 : "memory");
  } else {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc),
                 "r"(q.x), "r"(q.y), "r"(q.z), "r"(q.w)
                 : "memory");
  }
}

// One optimizer step on E consecutive elements held in registers.
template <int E>
__device__ __forceinline__ void update_regs(const HyperParams& h, float wd, float (&wv)[E],
                                            const float (&gv)[E], float (&mv)[E],
                                            float (&vv)[E]) {
#pragma unroll
  for (int k = 0; k < E; ++k) {
    float g = gv[k] * h.grad_scale + wd * wv[k];
    if (h.clip > 0.f) g = fminf(fmaxf(g, -h.clip), h.clip);
    if (h.opt == 0) {
      wv[k] -= h.lr * g;
    } else if (h.opt == 1) {
      mv[k] = mv[k] * h.momentum + g;
      wv[k] -= h.lr * (g + h.momentum * mv[k]);
    } else if (h.opt == 2) {
      vv[k] = h.decay * vv[k] + (1.f - h.decay) * g * g;
      mv[k] = h.momentum * mv[k] + h.lr * g * rsqrtf(vv[k] + h.epsilon);
      wv[k] -= mv[k];
    } else {
      const float b1 = h.momentum, b2 = h.decay;
      mv[k] = b1 * mv[k] + (1.f - b1) * g;
      vv[k] = b2 * vv[k] + (1.f - b2) * g * g;
      const float lr_t = h.lr * sqrtf(1.f - __powf(b2, h.adam_t)) / (1.f - __powf(b1, h.adam_t));
      wv[k] -= lr_t * mv[k] / (sqrtf(vv[k]) + h.epsilon);
    }
  }
}

template <int E>
__device__ __forceinline__ void load_f32(const float* p, float (&f)[E]) {
#pragma unroll
  for (int k = 0; k < E; k += 4) {
    const float4 t = *reinterpret_cast<const float4*>(p + k);
    f[k] = t.x; f[k + 1] = t.y; f[k + 2] = t.z; f[k + 3] = t.w;
  }
}
template <int E>
__device__ __forceinline__ void store_f32(float* p, const float (&f)[E]) {
#pragma unroll
  for (int k = 0; k < E; k += 4)
    *reinterpret_cast<float4*>(p + k) = make_float4(f[k], f[k + 1], f[k + 2], f[k + 3]);
}

template <int MODE, bool NVLS, bool HALF>
__global__ void __launch_bounds__(512)
    allreduce_update2_kernel(PeerCtx c, Ar2Args a, HyperParams h) {
  constexpr int E = HALF ? 8 : 4;             // elements per 16-byte vector on the wire
  constexpr int ES = HALF ? 2 : 4;            // bytes per element on the wire
  resolve_dyn(h, a.dyn);
  const uint32_t seq = a.seq + seq_offset(c);
  const int n = c.nranks;
  // --- announce: my source data for [offset, offset+count) is final ---
  if (blockIdx.x == 0 && threadIdx.x < n)
    st_release_sys(c.flags[threadIdx.x] + FLAG_READY * kMaxRanks + c.rank, seq);
  wait_flags(c, FLAG_READY, seq);

  // --- phase 1: reduce my shard (pull from the peers, or in the switch), hand the result to
  //     every rank ---
  const long long per = ((a.count + n - 1) / n + 7) & ~7ll;  // shard length, multiple of 8
  const long long s_begin = a.offset + per * c.rank;
  long long s_end = s_begin + per;
  if (s_end > a.offset + a.count) s_end = a.offset + a.count;
  const long long nvec = s_end > s_begin ? (s_end - s_begin) / E : 0;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const float inv_n = 1.f / static_cast<float>(n);
  constexpr int kU = 4;
  for (long long v0 = tid; v0 < nvec; v0 += nthreads * kU) {
    uint4 out[kU];
    bool ok[kU];
    long long byte_off[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long vi = v0 + u * nthreads;
      ok[u] = vi < nvec;
      byte_off[u] = (s_begin + (ok[u] ? vi : v0) * E) * ES;
    }
    if (NVLS) {
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (ok[u])
          out[u] = multimem_ld_reduce_vec<HALF>(static_cast<const char*>(a.src_mc) + byte_off[u]);
      if (MODE == 1) {
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          float f[E];
          vec_unpack<HALF>(out[u], f);
#pragma unroll
          for (int k = 0; k < E; ++k) f[k] *= inv_n;
          out[u] = vec_pack<HALF>(f);
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (ok[u]) multimem_st_vec<HALF>(static_cast<char*>(a.red_mc) + byte_off[u], out[u]);
    } else {
      float acc[kU][E];
#pragma unroll
      for (int u = 0; u < kU; ++u)
#pragma unroll
        for (int k = 0; k < E; ++k) acc[u][k] = 0.f;
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < n) {
          uint4 x[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u)
            x[u] = *reinterpret_cast<const uint4*>(static_cast<const char*>(a.src[r]) +
                                                   byte_off[u]);
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            float f[E];
            vec_unpack<HALF>(x[u], f);
#pragma unroll
            for (int k = 0; k < E; ++k) acc[u][k] += f[k];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (MODE == 1) {
#pragma unroll
          for (int k = 0; k < E; ++k) acc[u][k] *= inv_n;
        }
        out[u] = vec_pack<HALF>(acc[u]);
      }
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < n) {
#pragma unroll
          for (int u = 0; u < kU; ++u)
            if (ok[u])
              *reinterpret_cast<uint4*>(static_cast<char*>(a.red[r]) + byte_off[u]) = out[u];
        }
      }
    }
  }
  grid_arrive_and_signal(c, FLAG_REDUCED, seq, 0);
  wait_flags(c, FLAG_REDUCED, seq);

  // --- phase 2: every rank applies the identical update to its full replica ---
  const bool skip = (a.skip_flag != nullptr && *a.skip_flag != 0) || a.reduce_only;
  if (skip) return;
  constexpr int kU2 = HALF ? 2 : 4;           // 4 x 16 bytes of w / m in flight per thread
  const long long nv2 = a.count / E;
  const char* red = static_cast<const char*>(a.red_local);
  for (long long v0 = tid; v0 < nv2; v0 += nthreads * kU2) {
    uint4 rq[kU2];
    float wv[kU2][E], mv[kU2][E], vv[kU2][E], gl[kU2][E];
    long long i[kU2];
    bool ok[kU2];
#pragma unroll
    for (int u = 0; u < kU2; ++u) {
      const long long vi = v0 + u * nthreads;
      ok[u] = vi < nv2;
      i[u] = a.offset + (ok[u] ? vi : v0) * E;
      rq[u] = __ldcg(reinterpret_cast<const uint4*>(red + i[u] * ES));
    }
#pragma unroll
    for (int u = 0; u < kU2; ++u) load_f32<E>(a.w + i[u], wv[u]);
    if (h.opt != 0) {
#pragma unroll
      for (int u = 0; u < kU2; ++u) load_f32<E>(a.m + i[u], mv[u]);
    }
    if (h.opt >= 2) {
#pragma unroll
      for (int u = 0; u < kU2; ++u) load_f32<E>(a.v + i[u], vv[u]);
    }
    if (MODE == 1) {
#pragma unroll
      for (int u = 0; u < kU2; ++u) load_f32<E>(a.g_local + i[u], gl[u]);
    }
#pragma unroll
    for (int u = 0; u < kU2; ++u) {
      float r[E];
      vec_unpack<HALF>(rq[u], r);
      if (MODE == 1) {
        // SMA: w <- (1-a) w + a * avg(w); then the local gradient step
#pragma unroll
        for (int k = 0; k < E; ++k)
          wv[u][k] = (1.f - a.sma_alpha) * wv[u][k] + a.sma_alpha * r[k];
        update_regs<E>(h, decay_at(h, i[u]), wv[u], gl[u], mv[u], vv[u]);
      } else {
        update_regs<E>(h, decay_at(h, i[u]), wv[u], r, mv[u], vv[u]);
      }
      if (!ok[u]) continue;
      store_f32<E>(a.w + i[u], wv[u]);
      if (h.opt != 0) store_f32<E>(a.m + i[u], mv[u]);
      if (h.opt >= 2) store_f32<E>(a.v + i[u], vv[u]);
      if (a.w16 != nullptr) {
#pragma unroll
        for (int k = 0; k < E; k += 4) {
          uint2 q;
          q.x = pack_half2(wv[u][k], wv[u][k + 1]);
          q.y = pack_half2(wv[u][k + 2], wv[u][k + 3]);
          *reinterpret_cast<uint2*>(a.w16 + i[u] + k) = q;
        }
      }
    }
  }
}

// fp32 -> fp16 copy of a gradient range into the (symmetric) compact-transfer arena
__global__ void __launch_bounds__(256)
    cast_range_f16_kernel(const float* __restrict__ in, __half* __restrict__ out, long long n) {
  const long long nv = n / 8;
  for (long long v0 = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < nv;
       v0 += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float4 a0 = *reinterpret_cast<const float4*>(in + v0 * 8);
    const float4 a1 = *reinterpret_cast<const float4*>(in + v0 * 8 + 4);
    uint4 q;
    q.x = pack_half2(a0.x, a0.y);
    q.y = pack_half2(a0.z, a0.w);
    q.z = pack_half2(a1.x, a1.y);
    q.w = pack_half2(a1.z, a1.w);
    *reinterpret_cast<uint4*>(out + v0 * 8) = q;
  }
}

// Pull `count` floats at `offset` from rank `root`'s arena into mine (+ fp16 shadow).
__global__ void __launch_bounds__(512)
    broadcast_pull_kernel(PeerCtx c, int root, long long offset, long long count, uint32_t seq,
                          float* __restrict__ dst, __half* __restrict__ dst16) {
  seq += seq_offset(c);
  if (blockIdx.x == 0 && threadIdx.x < c.nranks)
    st_release_sys(c.flags[threadIdx.x] + FLAG_READY * kMaxRanks + c.rank, seq);
  wait_flags(c, FLAG_READY, seq);
  const float* src = c.src[root];
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  // 16-byte peer loads, four in flight per thread (offset is 4-aligned; scalar tail)
  const long long end = offset + count, end4 = offset + (count & ~3ll);
  const bool vec = (offset & 3) == 0;
  constexpr int kU = 4;
  for (long long i0 = offset + tid * 4; vec && i0 < end4; i0 += nthreads * 4 * kU) {
    float4 x[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      if (i < end4) x[u] = *reinterpret_cast<const float4*>(src + i);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * nthreads * 4;
      if (i >= end4) continue;
      if (c.rank != root) *reinterpret_cast<float4*>(dst + i) = x[u];
      if (dst16 != nullptr) {
        uint2 q;
        q.x = pack_half2(x[u].x, x[u].y);
        q.y = pack_half2(x[u].z, x[u].w);
        *reinterpret_cast<uint2*>(dst16 + i) = q;
      }
    }
  }
  for (long long i = (vec ? end4 : offset) + tid; i < end; i += nthreads) {
    const float x = src[i];
    if (c.rank != root) dst[i] = x;
    if (dst16 != nullptr) dst16[i] = __float2half_rn(x);
  }
  grid_arrive_and_signal(c, FLAG_DONE, seq, 1);
  wait_flags(c, FLAG_DONE, seq);
}

__global__ void device_barrier_kernel(PeerCtx c, uint32_t seq) {
  seq += seq_offset(c);
  if (threadIdx.x < c.nranks) {
    __threadfence_system();
    st_release_sys(c.flags[threadIdx.x] + FLAG_BARRIER * kMaxRanks + c.rank, seq);
  }
  wait_flags(c, FLAG_BARRIER, seq);
}

// flag <- OR over ranks of flag (auto loss scale: every rank must take the same skip decision,
// T/variable_mgr.py:321-332). Each rank posts its value into every page, then reads all.
__global__ void flag_or_kernel(PeerCtx c, uint32_t seq, int* __restrict__ flag) {
  seq += seq_offset(c);
  // [parity][src_rank] words after the arrival counters: two sets, chosen by the parity of the
  // call's sequence number, so that a rank which already entered the next flag_or cannot
  // overwrite a value a slower rank has not read yet
  const int kPostWord = kCounterWord + 8 + static_cast<int>(seq & 1u) * kMaxRanks;
  if (threadIdx.x < c.nranks) {
    *reinterpret_cast<volatile uint32_t*>(c.flags[threadIdx.x] + kPostWord + c.rank) =
        static_cast<uint32_t>(*flag != 0);
    __threadfence_system();
    st_release_sys(c.flags[threadIdx.x] + FLAG_BARRIER * kMaxRanks + c.rank, seq);
  }
  wait_flags(c, FLAG_BARRIER, seq);
  if (threadIdx.x == 0) {
    uint32_t any = 0;
    for (int r = 0; r < c.nranks; ++r)
      any |= *reinterpret_cast<volatile uint32_t*>(c.flags[c.rank] + kPostWord + r);
    *flag = any ? 1 : 0;
  }
}

// ---------------------------------------------------------------------------
// PairAveraging (AD-PSGD). Each rank publishes its model into a ring of 3 snapshot buffers
// guarded by {begin, end} version words (a seqlock): version k lives in slot k % 3. A reader
// picks the peer's newest complete version, reads it, and re-checks `begin`: the slot can only
// have been overwritten if the writer began version k+3 meanwhile.
// ---------------------------------------------------------------------------
struct PairCtx {
  const float* peer_pub;       // peer's publish ring base (3 * stride floats)
  const uint32_t* peer_ver;    // peer's {begin, end}
  float* my_pub;               // my publish ring
  uint32_t* my_ver;            // my {begin, end}
  long long stride;            // floats per slot
};

// Grid-wide barrier of a fully resident grid (the host clamps the grid): sync[0] counts
// arrivals, sync[1] is the generation.
__device__ __forceinline__ void pair_grid_barrier(unsigned* sync) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned gen = *reinterpret_cast<volatile unsigned*>(sync + 1);
    __threadfence();
    if (atomicAdd(sync, 1u) == gridDim.x - 1) {
      atomicExch(sync, 0u);
      __threadfence();
      atomicAdd(sync + 1, 1u);
    } else {
      const uint64_t t0 = globaltimer_ns();
      while (*reinterpret_cast<volatile unsigned*>(sync + 1) == gen) {
        if (globaltimer_ns() - t0 > 20000000000ull) {
          printf("kfb: pair_average grid barrier timeout (block %d)\n", blockIdx.x);
          __trap();
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// Phase A pulls the peer's newest complete snapshot into a local buffer with 16-byte loads and
// validates it afterwards (seqlock: the slot can only have been overwritten if the writer began
// version k+3 meanwhile); a torn pull is REPEATED with the then-newest version (up to three
// times) before anything is blended, so the model never mixes half-old and half-new weights.
// Phase B blends from the local copy, applies the local gradient and publishes.
// sync: {barrier count, barrier generation, retry flag, -}; torn_counter counts repeated pulls.
__global__ void __launch_bounds__(512)
    pair_average_kernel(PairCtx pc, long long count, float* __restrict__ w,
                        const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                        __half* __restrict__ w16, HyperParams h, const float* __restrict__ dyn,
                        uint32_t my_next_version, int has_peer, int* __restrict__ torn_counter,
                        float* __restrict__ pulled, unsigned* __restrict__ sync) {
  resolve_dyn(h, dyn);
  __shared__ uint32_t s_ver;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long n4 = count & ~3ll;
  bool blend = false;
  for (int attempt = 0; has_peer && attempt < 3; ++attempt) {
    if (threadIdx.x == 0) s_ver = ld_acquire_sys(pc.peer_ver + 1);
    __syncthreads();
    const uint32_t ver = s_ver;
    if (ver == 0) break;  // nothing published yet
    const float* peer = pc.peer_pub + static_cast<long long>(ver % 3u) * pc.stride;
    constexpr int kU = 4;
    for (long long i0 = tid * 4; i0 < n4; i0 += nthreads * 4 * kU) {
      float4 x[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = i0 + u * nthreads * 4;
        if (i < n4) x[u] = *reinterpret_cast<const float4*>(peer + i);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long i = i0 + u * nthreads * 4;
        if (i < n4) *reinterpret_cast<float4*>(pulled + i) = x[u];
      }
    }
    for (long long i = n4 + tid; i < count; i += nthreads) pulled[i] = peer[i];
    // every block validates its own reads; one torn block makes the whole grid pull again
    if (threadIdx.x == 0 && ld_acquire_sys(pc.peer_ver) >= ver + 3u) atomicExch(sync + 2, 1u);
    pair_grid_barrier(sync);
    const bool torn = *reinterpret_cast<volatile unsigned*>(sync + 2) != 0;
    pair_grid_barrier(sync);               // everyone has read the flag
    if (blockIdx.x == 0 && threadIdx.x == 0 && torn) {
      atomicExch(sync + 2, 0u);
      atomicAdd(torn_counter, 1);
    }
    if (!torn) {
      blend = true;
      break;
    }
    pair_grid_barrier(sync);               // flag cleared before the next attempt may set it
  }
  float* mine = pc.my_pub + static_cast<long long>(my_next_version % 3u) * pc.stride;
  // every block announces `begin` before its own first store into the slot
  if (threadIdx.x == 0) st_release_sys(pc.my_ver, my_next_version);
  __syncthreads();
  for (long long i = tid * 4; i < n4; i += nthreads * 4) {
    float4 p4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (blend) p4 = __ldcg(reinterpret_cast<const float4*>(pulled + i));
    apply_update4(h, w, *reinterpret_cast<const float4*>(g + i), m, v, w16, i,
                  blend ? &p4 : nullptr, 0.5f);
    *reinterpret_cast<float4*>(mine + i) = *reinterpret_cast<const float4*>(w + i);
  }
  for (long long i = n4 + tid; i < count; i += nthreads) {
    float wi = w[i];
    if (blend) wi = 0.5f * (wi + pulled[i]);
    const float nw = apply_update(h, wi, g[i], m, v, i);
    w[i] = nw;
    if (w16 != nullptr) w16[i] = __float2half_rn(nw);
    mine[i] = nw;
  }
  // publish: all blocks' stores must be visible before `end` moves
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* counter = pc.my_ver + 2;
    const uint32_t prev = atomicAdd(counter, 1u);
    if (prev == gridDim.x - 1) {
      *counter = 0;
      __threadfence_system();
      st_release_sys(pc.my_ver + 1, my_next_version);  // end
    }
  }
}

// device counter added to the sequence numbers of every collective launched by this process
// (one process drives one GPU); null = sequence numbers are taken as passed
static const uint32_t* g_seq_base = nullptr;

__global__ void seq_advance_kernel(uint32_t* counter, uint32_t k) { *counter += k; }

static int fill_peer_ctx(PeerCtx* c, int rank, int nranks, void* const* src, void* const* reduced,
                         void* const* flags) {
  KFB_REQUIRE(nranks >= 1 && nranks <= kMaxRanks && rank >= 0 && rank < nranks,
              "peer ctx: bad rank %d / %d", rank, nranks);
  c->rank = rank;
  c->nranks = nranks;
  c->seq_base = g_seq_base;
  for (int r = 0; r < kMaxRanks; ++r) {
    c->src[r] = r < nranks ? (float*)src[r] : nullptr;
    c->reduced[r] = (r < nranks && reduced) ? (float*)reduced[r] : nullptr;
    c->flags[r] = r < nranks ? (uint32_t*)flags[r] : nullptr;
  }
  return 0;
}

// The collectives publish a flag only after every CTA of the grid has arrived and then spin on
// flags of the same launch: the whole grid must be resident at once. Clamps a requested CTA
// count to what the device can hold for this kernel (queried, not assumed).
template <typename K>
static int clamp_resident(K kernel, int threads, int blocks) {
  int dev = 0, sms = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0) != cudaSuccess ||
      sms <= 0 || per_sm <= 0)
    return blocks < 1 ? 1 : (blocks > 64 ? 64 : blocks);
  const int cap = sms * per_sm;
  return blocks < 1 ? 1 : (blocks > cap ? cap : blocks);
}

static HyperParams make_hyper(const kfb_hyper* hp) {
  HyperParams h;
  h.opt = hp->opt;
  h.lr = hp->lr;
  h.momentum = hp->momentum;
  h.decay = hp->decay;
  h.epsilon = hp->epsilon;
  h.weight_decay = hp->weight_decay;
  h.grad_scale = hp->grad_scale;
  h.clip = hp->clip;
  h.adam_t = hp->adam_t;
  h.decay_mask = hp->decay_mask;
  return h;
}

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_flag_page_bytes(void) { return 4096; }

// sizeof of the structures that cross the C ABI (0 kfb_conv_desc, 1 kfb_hyper, 2 kfb_push_target);
// bindings compare their mirrors against these. -1 for an unknown index.
extern "C" int kfb_sizeof(int which) {
  switch (which) {
    case 0: return (int)sizeof(kfb_conv_desc);
    case 1: return (int)sizeof(kfb_hyper);
    case 2: return (int)sizeof(kfb_push_target);
    default: return -1;
  }
}

// Sequence numbers from device memory (multi-rank CUDA graphs): after kfb_set_seq_base(p) every
// collective of this process adds *p (a uint32 in device memory, 0 at first) to the `seq` it is
// given. A captured step ends with kfb_seq_advance(p, k), k = collectives per step, so each
// replay of the same captured constants runs with fresh sequence numbers; the host mirrors the
// counter and passes `absolute - mirror` for calls outside the graph. Null switches it off.
extern "C" int kfb_set_seq_base(const void* counter) {
  g_seq_base = static_cast<const uint32_t*>(counter);
  return 0;
}

extern "C" int kfb_seq_advance(void* counter, unsigned k, void* stream) {
  KFB_REQUIRE(counter != nullptr, "seq_advance: null counter");
  seq_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(static_cast<uint32_t*>(counter), k);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_optimizer_step(float* w, const float* g, float* m, float* v, void* w16,
                                  long long n, const kfb_hyper* hp, const float* dyn,
                                  const int* skip_flag, void* stream) {
  KFB_REQUIRE(hp != nullptr && hp->opt >= 0 && hp->opt <= 3, "optimizer_step: bad optimizer");
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  optimizer_step_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      w, g, m, v, (__half*)w16, n, make_hyper(hp), dyn, skip_flag);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_optimizer_step_blend(float* w, const float* avg, const float* g, float* m,
                                        float* v, void* w16, long long n, const kfb_hyper* hp,
                                        const float* dyn, float alpha, void* stream) {
  KFB_REQUIRE(hp != nullptr && avg != nullptr, "optimizer_step_blend: null argument");
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  optimizer_step_blend_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
      w, avg, g, m, v, (__half*)w16, n, make_hyper(hp), dyn, alpha);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_check_finite(const float* g, long long n, int* flag, void* stream) {
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  check_finite_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(g, n, flag);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_set_hyper_dyn(float* dyn, float lr, float grad_scale, float adam_t, int mask,
                                 void* stream) {
  KFB_REQUIRE(dyn != nullptr, "set_hyper_dyn: null dyn");
  set_hyper_dyn_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(dyn, lr, grad_scale, adam_t, mask);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_loss_scale_update(float* state, int* flag, int inc_every, float* dyn,
                                     float base_grad_scale, void* stream) {
  loss_scale_update_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(state, flag, inc_every, dyn,
                                                               base_grad_scale);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// mode 0: gradient all-reduce (+optimizer); mode 1: SMA (src = weight arenas, g_local = my grads)
extern "C" int kfb_allreduce_update(int mode, int rank, int nranks, void* const* src,
                                    void* const* reduced, void* const* flags, long long offset,
                                    long long count, unsigned seq, float* w, const float* g_local,
                                    float* m, float* v, void* w16, const kfb_hyper* hp,
                                    const float* dyn, const int* skip_flag, float sma_alpha,
                                    int blocks, void* stream) {
  KFB_REQUIRE(hp != nullptr, "allreduce_update: null hyper");
  KFB_REQUIRE(offset % 4 == 0, "allreduce_update: offset must be a multiple of 4 floats");
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, src, reduced, flags)) return 2;
  if (blocks <= 0) {
    // enough CTAs to keep the links busy, few enough to leave SMs to the backward kernels;
    // KFB_AR_BLOCKS overrides the choice for sweeps on the device (read once)
    static const int forced = [] {
      const char* e = getenv("KFB_AR_BLOCKS");
      return (e != nullptr && *e != 0) ? atoi(e) : 0;
    }();
    blocks = forced > 0 ? forced
                        : (count >= (4ll << 20) ? 96 : (count >= (1ll << 20) ? 64 : 32));
  }
  blocks = clamp_resident(allreduce_update_kernel<0>, 512, blocks);
  cudaStream_t st = (cudaStream_t)stream;
  // mode 2: reduce only - the sum lands in every rank's `reduced` buffer and is applied later
  // (variable_consistency=relaxed: one-step-stale gradients, T/batch_allreduce.py:353-388)
  if (mode == 1)
    allreduce_update_kernel<1><<<blocks, 512, 0, st>>>(c, offset, count, seq, w, g_local, m, v,
                                                       (__half*)w16, make_hyper(hp), dyn,
                                                       skip_flag, sma_alpha, 0, nullptr, 0);
  else
    allreduce_update_kernel<0><<<blocks, 512, 0, st>>>(c, offset, count, seq, w, g_local, m, v,
                                                       (__half*)w16, make_hyper(hp), dyn,
                                                       skip_flag, sma_alpha, mode == 2 ? 1 : 0,
                                                       nullptr, 0);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// NVLS form of kfb_allreduce_update: `src_mc` / `reduced_mc` are the multicast addresses of
// the source arena and of the landing buffer (element 0 of the arena, like the entries of
// src / reduced), `reduced_local` this rank's own landing buffer.
extern "C" int kfb_allreduce_update_nvls(int mode, int rank, int nranks, const void* src_mc,
                                         void* reduced_mc, void* reduced_local,
                                         void* const* flags, long long offset, long long count,
                                         unsigned seq, float* w, const float* g_local, float* m,
                                         float* v, void* w16, const kfb_hyper* hp,
                                         const float* dyn, const int* skip_flag,
                                         float sma_alpha, int blocks, void* stream) {
  KFB_REQUIRE(hp != nullptr, "allreduce_update_nvls: null hyper");
  KFB_REQUIRE(src_mc != nullptr && reduced_mc != nullptr && reduced_local != nullptr,
              "allreduce_update_nvls: multicast addresses are required (is the allocation bound "
              "to a multicast object?)");
  KFB_REQUIRE(offset % 4 == 0, "allreduce_update_nvls: offset must be a multiple of 4 floats");
  KFB_REQUIRE(((uintptr_t)src_mc | (uintptr_t)reduced_mc | (uintptr_t)reduced_local) % 16 == 0,
              "allreduce_update_nvls: arenas must be 16-byte aligned");
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, flags, nullptr, flags)) return 2;
  for (int r = 0; r < kMaxRanks; ++r) c.src[r] = nullptr;
  c.reduced[rank] = (float*)reduced_local;
  if (blocks <= 0) {
    // the switch does the adds: fewer CTAs than the pull kernel saturate the links
    blocks = count >= (4ll << 20) ? 64 : (count >= (1ll << 20) ? 48 : 24);
  }
  blocks = clamp_resident(allreduce_update_nvls_kernel<0>, 512, blocks);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == 1)
    allreduce_update_nvls_kernel<1><<<blocks, 512, 0, st>>>(
        c, (const float*)src_mc, (float*)reduced_mc, offset, count, seq, w, g_local, m, v,
        (__half*)w16, make_hyper(hp), dyn, skip_flag, sma_alpha, 0);
  else
    allreduce_update_nvls_kernel<0><<<blocks, 512, 0, st>>>(
        c, (const float*)src_mc, (float*)reduced_mc, offset, count, seq, w, g_local, m, v,
        (__half*)w16, make_hyper(hp), dyn, skip_flag, sma_alpha, mode == 2 ? 1 : 0);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Second generation (see allreduce_update2_kernel). Pull form: `src` / `reduced` hold every
// rank's arena / landing buffer addresses and src_mc = reduced_mc = null. NVLS form: src_mc /
// reduced_mc are the multicast aliases and src / reduced may be null. `half` = 1: the source
// arenas and the landing buffers hold fp16 (compact gradient transfer; mode 0 or 2 only).
// `cast_from` (optional, with half): this rank's fp32 gradients of the range are first copied as
// fp16 into `cast_to` (this rank's source arena) on the same stream.
extern "C" int kfb_allreduce_update2(int mode, int rank, int nranks, void* const* src,
                                     void* const* reduced, const void* src_mc, void* reduced_mc,
                                     const void* reduced_local, void* const* flags, int half,
                                     const float* cast_from, void* cast_to, long long offset,
                                     long long count, unsigned seq, float* w,
                                     const float* g_local, float* m, float* v, void* w16,
                                     const kfb_hyper* hp, const float* dyn, const int* skip_flag,
                                     float sma_alpha, int blocks, void* stream) {
  KFB_REQUIRE(hp != nullptr && reduced_local != nullptr, "allreduce_update2: null argument");
  KFB_REQUIRE(offset % 8 == 0 && count % 8 == 0,
              "allreduce_update2: offset (%lld) and count (%lld) must be multiples of 8", offset,
              count);
  KFB_REQUIRE(!(half && (mode == 1 || mode == 3)),
              "allreduce_update2: SMA averages the fp32 model");
  const bool nvls = src_mc != nullptr;
  KFB_REQUIRE(nvls ? reduced_mc != nullptr : (src != nullptr && reduced != nullptr),
              "allreduce_update2: incomplete %s addresses", nvls ? "multicast" : "peer");
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, flags, nullptr, flags)) return 2;
  Ar2Args a;
  for (int r = 0; r < kMaxRanks; ++r) {
    a.src[r] = (!nvls && r < nranks) ? src[r] : nullptr;
    a.red[r] = (!nvls && r < nranks) ? reduced[r] : nullptr;
  }
  a.src_mc = src_mc;
  a.red_mc = reduced_mc;
  a.red_local = reduced_local;
  a.offset = offset;
  a.count = count;
  a.seq = seq;
  a.w = w;
  a.g_local = g_local;
  a.m = m;
  a.v = v;
  a.w16 = (__half*)w16;
  a.dyn = dyn;
  a.skip_flag = skip_flag;
  a.sma_alpha = sma_alpha;
  a.reduce_only = (mode == 2 || mode == 3) ? 1 : 0;   // 3: average of the sources only (SMA)
  if (blocks <= 0) {
    static const int forced = [] {
      const char* e = getenv("KFB_AR_BLOCKS");
      return (e != nullptr && *e != 0) ? atoi(e) : 0;
    }();
    // measured on 2 B200s, 102 MB: 16 / 32 / 64 / 96 / 148 CTAs -> 710 / 376 / 298 / 259 / 263 us
    blocks = forced > 0 ? forced
                        : (count >= (4ll << 20) ? 96 : (count >= (1ll << 20) ? 64 : 32));
  }
  blocks = clamp_resident(allreduce_update2_kernel<0, false, true>, 512, blocks);
  cudaStream_t st = (cudaStream_t)stream;
  if (half && cast_from != nullptr) {
    KFB_REQUIRE(cast_to != nullptr, "allreduce_update2: cast_to missing");
    long long cb = (count / 8 + 255) / 256;
    if (cb > 2 * blocks) cb = 2 * blocks;
    if (cb < 1) cb = 1;
    cast_range_f16_kernel<<<(unsigned)cb, 256, 0, st>>>(cast_from + offset,
                                                       (__half*)cast_to + offset, count);
  }
  const HyperParams h = make_hyper(hp);
#define KFB_AR2(MODE, NVLS, HALF) \
  allreduce_update2_kernel<MODE, NVLS, HALF><<<blocks, 512, 0, st>>>(c, a, h)
  if (mode == 1 || mode == 3) {
    if (nvls) KFB_AR2(1, true, false); else KFB_AR2(1, false, false);
  } else if (half) {
    if (nvls) KFB_AR2(0, true, true); else KFB_AR2(0, false, true);
  } else {
    if (nvls) KFB_AR2(0, true, false); else KFB_AR2(0, false, false);
  }
#undef KFB_AR2
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Second half of the fused weight-gradient reduce-scatter: the gradients of [offset,
// offset+count) were pushed into this rank's `stage` slots (slot r = contribution of rank r,
// `stage_stride` floats apart, indexed like the arena) by kfb_conv_wgrad_rs / kfb_gemm_nt_rs /
// kfb_push_ranges on every rank. READY now means "my pushes are complete"; the shard is summed
// from local memory in rank order (bit-identical to the pull path), pushed to every rank's
// `reduced` buffer and applied as usual.
extern "C" int kfb_allreduce_update_staged(int rank, int nranks, void* const* reduced,
                                           void* const* flags, long long offset, long long count,
                                           unsigned seq, float* w, float* m, float* v, void* w16,
                                           const kfb_hyper* hp, const float* dyn,
                                           const int* skip_flag, const float* stage,
                                           long long stage_stride, int blocks, void* stream) {
  KFB_REQUIRE(hp != nullptr && stage != nullptr, "allreduce_update_staged: null argument");
  KFB_REQUIRE(offset % 4 == 0, "allreduce_update_staged: offset must be a multiple of 4");
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, reduced, reduced, flags)) return 2;
  if (blocks <= 0) {
    blocks = static_cast<int>((count / 4 + 511) / 512);
    if (blocks < 8) blocks = 8;
  }
  blocks = clamp_resident(allreduce_update_kernel<0>, 512, blocks);
  allreduce_update_kernel<0><<<blocks, 512, 0, (cudaStream_t)stream>>>(
      c, offset, count, seq, w, nullptr, m, v, (__half*)w16, make_hyper(hp), dyn, skip_flag, 0.f,
      0, stage, stage_stride);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Pushes gradients that are not produced by a GEMM epilogue (batch-norm scales, biases,
// depthwise filters ...) into the owners' staging slots, same ownership rule as kfb_push_target.
__global__ void __launch_bounds__(256)
    push_ranges_kernel(const float* __restrict__ grad, const long long* __restrict__ ranges,
                       int nranges, long long bucket_lo, long long per, int n, PushDst d) {
  for (int k = blockIdx.y; k < nranges; k += gridDim.y) {
    const long long lo = ranges[2 * k], hi = ranges[2 * k + 1];
    for (long long i = lo + (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
         i < hi; i += static_cast<long long>(gridDim.x) * blockDim.x * 4) {
      long long o = (i - bucket_lo) / per;
      o = o < 0 ? 0 : (o >= n ? n - 1 : o);
      if (i + 4 <= hi) {
        *reinterpret_cast<float4*>(d.base[o] + i) = *reinterpret_cast<const float4*>(grad + i);
      } else {
        for (long long j = i; j < hi; ++j) d.base[o][j] = grad[j];
      }
    }
  }
  __threadfence_system();
}

extern "C" int kfb_push_ranges(const float* grad, const long long* ranges_dev, int nranges,
                               long long max_range_elems, long long bucket_lo, long long per,
                               int nranks, void* const* slot_base, void* stream) {
  KFB_REQUIRE(nranks > 1 && nranks <= kMaxRanks && per > 0 && per % 4 == 0,
              "push_ranges: bad ranks / shard length");
  if (nranges <= 0) return 0;
  PushDst d;
  for (int r = 0; r < kMaxRanks; ++r) d.base[r] = r < nranks ? (float*)slot_base[r] : nullptr;
  long long bx = (max_range_elems / 4 + 255) / 256;
  if (bx < 1) bx = 1;
  if (bx > 64) bx = 64;
  dim3 grid((unsigned)bx, (unsigned)(nranges < 1024 ? nranges : 1024));
  push_ranges_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(grad, ranges_dev, nranges, bucket_lo,
                                                            per, nranks, d);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_broadcast_pull(int rank, int nranks, void* const* src, void* const* flags,
                                  int root, long long offset, long long count, unsigned seq,
                                  float* dst, void* dst16, int blocks, void* stream) {
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, src, nullptr, flags)) return 2;
  if (blocks <= 0) blocks = 64;
  blocks = clamp_resident(broadcast_pull_kernel, 512, blocks);
  broadcast_pull_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(c, root, offset, count, seq,
                                                                  dst, (__half*)dst16);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_device_barrier(int rank, int nranks, void* const* flags, unsigned seq,
                                  void* stream) {
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, flags, nullptr, flags)) return 2;
  device_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(c, seq);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_flag_or(int rank, int nranks, void* const* flags, unsigned seq, int* flag,
                           void* stream) {
  PeerCtx c;
  if (fill_peer_ctx(&c, rank, nranks, flags, nullptr, flags)) return 2;
  flag_or_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(c, seq, flag);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_pair_average_update(const float* peer_pub, const unsigned* peer_ver,
                                       float* my_pub, unsigned* my_ver, long long stride,
                                       long long count, float* w, const float* g, float* m,
                                       float* v, void* w16, const kfb_hyper* hp, const float* dyn,
                                       unsigned my_next_version, int has_peer, int* torn_counter,
                                       float* pull_buf, unsigned* sync, int blocks,
                                       void* stream) {
  KFB_REQUIRE(hp != nullptr, "pair_average_update: null hyper");
  KFB_REQUIRE(!has_peer || (pull_buf != nullptr && sync != nullptr),
              "pair_average_update: needs the pull buffer and four zeroed sync words");
  PairCtx pc;
  pc.peer_pub = peer_pub;
  pc.peer_ver = peer_ver;
  pc.my_pub = my_pub;
  pc.my_ver = my_ver;
  pc.stride = stride;
  if (blocks <= 0) blocks = 64;
  blocks = clamp_resident(pair_average_kernel, 512, blocks);
  pair_average_kernel<<<blocks, 512, 0, (cudaStream_t)stream>>>(
      pc, count, w, g, m, v, (__half*)w16, make_hyper(hp), dyn, my_next_version, has_peer,
      torn_counter, pull_buf, sync);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Symmetric-memory runtime: device allocations exported over CUDA IPC.
// ---------------------------------------------------------------------------
extern "C" int kfb_symm_alloc(long long bytes, void** out) {
  KFB_REQUIRE(bytes > 0 && out != nullptr, "symm_alloc: bad arguments");
  void* p = nullptr;
  KFB_CHECK_CUDA(cudaMalloc(&p, (size_t)bytes));
  KFB_CHECK_CUDA(cudaMemset(p, 0, (size_t)bytes));
  KFB_CHECK_CUDA(cudaDeviceSynchronize());
  *out = p;
  return 0;
}

extern "C" int kfb_symm_free(void* p) {
  KFB_CHECK_CUDA(cudaFree(p));
  return 0;
}

extern "C" int kfb_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int kfb_ipc_get_handle(void* p, void* handle_out) {
  cudaIpcMemHandle_t h;
  KFB_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle_out, &h, sizeof(h));
  return 0;
}

extern "C" int kfb_ipc_open(const void* handle, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  void* p = nullptr;
  KFB_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out = p;
  return 0;
}

extern "C" int kfb_ipc_close(void* p) {
  KFB_CHECK_CUDA(cudaIpcCloseMemHandle(p));
  return 0;
}

extern "C" int kfb_enable_peer_access(int peer_device) {
  int dev = 0;
  KFB_CHECK_CUDA(cudaGetDevice(&dev));
  if (peer_device == dev) return 0;
  int can = 0;
  KFB_CHECK_CUDA(cudaDeviceCanAccessPeer(&can, dev, peer_device));
  KFB_REQUIRE(can, "device %d cannot access peer %d", dev, peer_device);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_device, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  KFB_CHECK_CUDA(e);
  return 0;
}
