// Single-launch batch normalisation for sm_100a: the whole forward (fold of the conv epilogue's
// partial statistics -> mean / variance / moving averages -> apply + residual + ReLU) and the
// whole backward (per-channel reductions -> dgamma / dbeta / coefficients -> dz and the
// residual gradient) each run as ONE persistent kernel whose phases are separated by grid-wide
// barriers instead of kernel boundaries.
//
// Why: the reference gets batch norm as one fused op per direction
// (tf.nn.fused_batch_norm, T/convnet_builder_legacy.py:418-431). The three-kernel forms of
// nn_kernels.cu (reduce, finalize, apply) cost 25-40 us of launch / drain / ramp latency per
// layer on top of their memory time (profiles/node_timing_r2b.md: a 25 MB layer took 60 us
// against a 20 us roof); ResNet-50 has 53 such layers in each direction.
//
// Layout and thread mapping are those of nn_kernels.cu: NHWC fp16 viewed as [rows][C] with a
// row stride, one thread = one 16-byte channel vector of a row lane, each CTA owns a
// contiguous slab of rows. The grid is sized by the host so that every CTA is resident at the
// same time (a grid barrier needs that); `sync` is two zero-initialised words per caller.
#include "kfb_common.cuh"
#include "../include/kfb200.h"

#include <stdlib.h>

namespace kfb {

constexpr int kBnThreads = 256;
constexpr int kBnWarps = kBnThreads / 32;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Grid-wide barrier for a grid whose CTAs are all resident. sync[0] counts arrivals, sync[1] is
// the generation; the last CTA to arrive resets the count and bumps the generation. Bounded:
// a missing CTA becomes a trap (CUDA error), never a hang.
__device__ __forceinline__ void grid_barrier(unsigned* sync, unsigned nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned gen = ld_acquire_u32(sync + 1);
    __threadfence();
    const unsigned prev = atomicAdd(sync, 1u);
    if (prev == nblocks - 1) {
      atomicExch(sync, 0u);
      __threadfence();
      atomicAdd(sync + 1, 1u);
    } else {
      const uint64_t t0 = globaltimer_ns();
      unsigned spins = 0;
      while (ld_acquire_u32(sync + 1) == gen) {
        if ((++spins & 0xff) == 0 && globaltimer_ns() - t0 > 4000000000ull) {
          printf("kfb: grid barrier timeout (block %d of %u)\n", blockIdx.x, nblocks);
          __trap();
        }
      }
    }
    __threadfence();
  }
  __syncthreads();
}

// butterfly sum with a fixed order (deterministic); every lane ends with the total
__device__ __forceinline__ float warp_sum_det(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct SlabMap {
  int vpr, cv_per_pass, row_lanes, tv, tr;
  long long r_begin, r_end;
};

__device__ __forceinline__ SlabMap make_slab_map(long long rows, int C) {
  SlabMap m;
  m.vpr = C >> 3;
  m.cv_per_pass = m.vpr < kBnThreads ? m.vpr : kBnThreads;
  m.row_lanes = kBnThreads / m.cv_per_pass;
  m.tv = threadIdx.x % m.cv_per_pass;
  m.tr = threadIdx.x / m.cv_per_pass;
  const long long per = (rows + gridDim.x - 1) / gridDim.x;
  m.r_begin = static_cast<long long>(blockIdx.x) * per;
  m.r_end = m.r_begin + per;
  if (m.r_end > rows) m.r_end = rows;
  if (m.r_begin > rows) m.r_begin = rows;
  return m;
}

// ReLU mask of y = relu(x*sc + sh) recomputed from x: (x > thr) xor (sc < 0), thr = -sh/sc.
__device__ __forceinline__ void relu_thresholds(const float* __restrict__ relu_affine, int C,
                                                int c0, float (&thr)[8], unsigned& neg_bits) {
  neg_bits = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float sc = __ldcg(relu_affine + c0 + i), sh = __ldcg(relu_affine + C + c0 + i);
    if (sc != 0.f) {
      thr[i] = -sh / sc;
      if (sc < 0.f) neg_bits |= 1u << i;
    } else {
      thr[i] = sh > 0.f ? -INFINITY : INFINITY;
    }
  }
}

// folds the row lanes of one pass through shared memory; lanes with tr == 0 return the sums
__device__ __forceinline__ void fold_row_lanes(float* red, const SlabMap& m, float (&a)[8],
                                               float (&b)[8]) {
  __syncthreads();
  float* mine = red + threadIdx.x * 16;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mine[i] = a[i];
    mine[8 + i] = b[i];
  }
  __syncthreads();
  if (m.tr == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = b[i] = 0.f;
    for (int l = 0; l < m.row_lanes; ++l) {
      const float* o = red + (l * m.cv_per_pass + m.tv) * 16;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        a[i] += o[i];
        b[i] += o[8 + i];
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Forward (training)
// ---------------------------------------------------------------------------
struct BnFwdArgs {
  const __half* x; long long ldx;
  const __half* res; long long ldr;
  __half* y; long long ldy;
  unsigned char* bitmask;     // optional [rows][C/8]: bit i of byte (r, cv) = (y[r][8cv+i] > 0)
  long long rows; int C;
  const float* gamma; const float* beta;
  float* moving_mean; float* moving_var;
  float eps, decay, inv_count, unbias;
  int update_moving, relu;
  float* mean_rstd; float* scale_shift;
  // statistics: either partial rows produced by the conv epilogue (stats_rows > 0): for channel
  // c = nb*stats_bn + j the rows are stats[((nb*stats_rows + r)*2 + s)*stats_bn + j]; or
  // (stats_rows == 0) this kernel reduces x itself into partial[grid][2][C] first
  const float* stats; int stats_rows, stats_bn;
  float* partial;
  unsigned* sync;
  unsigned long long* dbg;   // optional: globaltimer stamps of CTA 0 at the phase boundaries
};

#define KFB_BN_STAMP(i)                                                     \
  do {                                                                      \
    if (p.dbg != nullptr && blockIdx.x == 0 && threadIdx.x == 0) p.dbg[i] = globaltimer_ns(); \
  } while (0)

template <int KU, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
    bn_fwd_fused_kernel(const BnFwdArgs p) {
  extern __shared__ float red[];
  const SlabMap m = make_slab_map(p.rows, p.C);
  const int C = p.C;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool own_stats = p.stats_rows == 0;

  // ---- phase 0 (only without conv-epilogue statistics): per-CTA partial sums of x, x^2 ----
  if (own_stats) {
    for (int cv0 = 0; cv0 < m.vpr; cv0 += m.cv_per_pass) {
      const int cv = cv0 + m.tv;
      const bool active = cv < m.vpr && m.tr < m.row_lanes;
      float a[8], b[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = b[i] = 0.f;
      if (active) {
        constexpr int kU = 4;
        long long r = m.r_begin + m.tr;
        for (; r + (kU - 1) * m.row_lanes < m.r_end; r += kU * m.row_lanes) {
          Half8 q[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u)
            q[u] = *reinterpret_cast<const Half8*>(p.x + (r + u * m.row_lanes) * p.ldx + cv * 8);
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            float v[8];
            unpack8(q[u], v);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              a[i] += v[i];
              b[i] = fmaf(v[i], v[i], b[i]);
            }
          }
        }
        for (; r < m.r_end; r += m.row_lanes) {
          float v[8];
          load8(p.x + r * p.ldx + cv * 8, v);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            a[i] += v[i];
            b[i] = fmaf(v[i], v[i], b[i]);
          }
        }
      }
      fold_row_lanes(red, m, a, b);
      if (m.tr == 0 && cv < m.vpr) {
        float* pa = p.partial + static_cast<long long>(blockIdx.x) * 2 * C + cv * 8;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          pa[i] = a[i];
          pa[C + i] = b[i];
        }
      }
    }
    grid_barrier(p.sync, gridDim.x);
  }

  // ---- phase 1: one warp per channel folds the partial rows and finalises the channel ----
  for (int c = blockIdx.x * kBnWarps + warp; c < C; c += gridDim.x * kBnWarps) {
    float s = 0.f, ss = 0.f;
    {
      // rows of partial sums for channel c: `nrows` rows, `rstride` floats apart, the sum of
      // squares `sq_off` floats after the sum. Eight loads per lane are issued before the first
      // add: the fold is a chain of L2 / DRAM round trips otherwise (it was 15-20 us of the
      // ~40 us a 25 MB layer took; profiles/resnet50_bs256_launches_by_layer_r2.txt)
      const float* base;
      int nrows;
      long long rstride, sq_off;
      if (own_stats) {
        base = p.partial + c;
        nrows = static_cast<int>(gridDim.x);
        rstride = 2ll * C;
        sq_off = C;
      } else {
        const int nb = c / p.stats_bn, j = c - nb * p.stats_bn;
        base = p.stats + static_cast<long long>(nb) * p.stats_rows * 2 * p.stats_bn + j;
        nrows = p.stats_rows;
        rstride = 2ll * p.stats_bn;
        sq_off = p.stats_bn;
      }
      constexpr int kF = 8;
      for (int r0 = lane; r0 < nrows; r0 += 32 * kF) {
        float va[kF], vb[kF];
#pragma unroll
        for (int u = 0; u < kF; ++u) {
          const int r = r0 + 32 * u;
          const float* q = base + static_cast<long long>(r < nrows ? r : lane) * rstride;
          va[u] = __ldcg(q);
          vb[u] = __ldcg(q + sq_off);
          if (r >= nrows) va[u] = vb[u] = 0.f;
        }
#pragma unroll
        for (int u = 0; u < kF; ++u) {
          s += va[u];
          ss += vb[u];
        }
      }
    }
    s = warp_sum_det(s);
    ss = warp_sum_det(ss);
    if (lane == 0) {
      const float mean = s * p.inv_count;
      float var = ss * p.inv_count - mean * mean;
      var = var < 0.f ? 0.f : var;
      const float rstd = rsqrtf(var + p.eps);
      p.mean_rstd[c] = mean;
      p.mean_rstd[C + c] = rstd;
      const float g = p.gamma ? p.gamma[c] : 1.f;
      const float sc = g * rstd;
      p.scale_shift[c] = sc;
      p.scale_shift[C + c] = p.beta[c] - mean * sc;
      if (p.update_moving) {
        p.moving_mean[c] = p.moving_mean[c] * p.decay + mean * (1.f - p.decay);
        p.moving_var[c] = p.moving_var[c] * p.decay + var * p.unbias * (1.f - p.decay);
      }
    }
  }
  grid_barrier(p.sync, gridDim.x);

  // ---- phase 2: y = act(x*scale + shift (+ residual)); rows last-to-first (the end of x was
  //      written most recently by the producing conv and is still in L2) ----
  for (int cv0 = 0; cv0 < m.vpr; cv0 += m.cv_per_pass) {
    const int cv = cv0 + m.tv;
    if (!(cv < m.vpr && m.tr < m.row_lanes)) continue;
    float sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sc[i] = __ldcg(p.scale_shift + cv * 8 + i);
      sh[i] = __ldcg(p.scale_shift + C + cv * 8 + i);
    }
    constexpr int kU = KU;
    const long long n_mine = m.r_end > m.r_begin + m.tr
                                 ? (m.r_end - m.r_begin - m.tr + m.row_lanes - 1) / m.row_lanes
                                 : 0;
    for (long long k0 = 0; k0 < n_mine; k0 += kU) {
      long long ru[kU];
      bool ok[kU];
      Half8 xq[kU], rq[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long k = n_mine - 1 - (k0 + u);
        ok[u] = k >= 0;
        ru[u] = m.r_begin + m.tr + (ok[u] ? k : 0) * m.row_lanes;
        xq[u] = *reinterpret_cast<const Half8*>(p.x + ru[u] * p.ldx + cv * 8);
      }
      if (p.res != nullptr) {
#pragma unroll
        for (int u = 0; u < kU; ++u)
          rq[u] = *reinterpret_cast<const Half8*>(p.res + ru[u] * p.ldr + cv * 8);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        float v[8];
        unpack8(xq[u], v);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = fmaf(v[i], sc[i], sh[i]);
        if (p.res != nullptr) {
          float w[8];
          unpack8(rq[u], w);
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] += w[i];
        }
        unsigned bits = 0;
        if (p.relu) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            bits |= (v[i] > 0.f ? 1u : 0u) << i;
            v[i] = fmaxf(v[i], 0.f);
          }
        }
        if (ok[u]) {
          store8(p.y + ru[u] * p.ldy + cv * 8, v);
          if (p.bitmask != nullptr)
            p.bitmask[ru[u] * m.vpr + cv] = static_cast<unsigned char>(bits);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Backward
// ---------------------------------------------------------------------------
struct BnBwdArgs {
  const __half* dy; long long lddy;
  const __half* mask; long long ldm;        // forward output (ReLU mask source), optional
  const unsigned char* bitmask;            // or the packed mask written by the forward, optional
  const float* relu_affine;                // or recompute the mask from x (no residual), optional
  const __half* x; long long ldx;
  long long rows; int C;
  const float* gamma; const float* mean_rstd;
  float* dgamma; float* dbeta; int accumulate; float grad_scale; float inv_count;
  float* coef;                             // [3][C]
  float* partial;                          // [grid][2][C]
  __half* dz; long long lddz;
  __half* dres; long long lddr; int dres_accumulate;
  unsigned* sync;
  unsigned long long* dbg;
};

// g = dy masked by the ReLU of the forward output; MASK: 0 none, 1 from y, 2 from x, 3 bitmask
template <int MASK>
__device__ __forceinline__ void masked_grad(const Half8& gq, const Half8& mq, unsigned bits,
                                            const float (&v)[8], const float (&thr)[8],
                                            unsigned neg_bits, float (&g)[8]) {
  unpack8(gq, g);
  if (MASK == 1) {
    float mm[8];
    unpack8(mq, mm);
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = mm[i] > 0.f ? g[i] : 0.f;
  } else if (MASK == 2) {
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = ((v[i] > thr[i]) != (((neg_bits >> i) & 1u) != 0)) ? g[i] : 0.f;
  } else if (MASK == 3) {
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = ((bits >> i) & 1u) ? g[i] : 0.f;
  }
}

template <int MASK, int KU, int MINB>
__global__ void __launch_bounds__(kBnThreads, MINB)
    bn_bwd_fused_kernel(const BnBwdArgs p) {
  extern __shared__ float red[];
  const SlabMap m = make_slab_map(p.rows, p.C);
  const int C = p.C;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int kU = KU;
  KFB_BN_STAMP(0);

  // ---- phase 1: a = sum g, b = sum g * (x - mean) * rstd over this CTA's slab ----
  for (int cv0 = 0; cv0 < m.vpr; cv0 += m.cv_per_pass) {
    const int cv = cv0 + m.tv;
    const bool active = cv < m.vpr && m.tr < m.row_lanes;
    float a[8], b[8], mu[8], thr[8];
    unsigned neg_bits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = b[i] = mu[i] = thr[i] = 0.f;
    if (active) {
#pragma unroll
      for (int i = 0; i < 8; ++i) mu[i] = p.mean_rstd[cv * 8 + i];
      if (MASK == 2) relu_thresholds(p.relu_affine, C, cv * 8, thr, neg_bits);
      for (long long r = m.r_begin + m.tr; r < m.r_end; r += kU * m.row_lanes) {
        long long ru[kU];
        bool ok[kU];
        Half8 gq[kU], xq[kU], mq[kU];
        unsigned bits[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          ru[u] = r + static_cast<long long>(u) * m.row_lanes;
          ok[u] = ru[u] < m.r_end;
          if (!ok[u]) ru[u] = r;
          gq[u] = *reinterpret_cast<const Half8*>(p.dy + ru[u] * p.lddy + cv * 8);
          bits[u] = 0;
        }
#pragma unroll
        for (int u = 0; u < kU; ++u)
          xq[u] = *reinterpret_cast<const Half8*>(p.x + ru[u] * p.ldx + cv * 8);
        if (MASK == 1) {
#pragma unroll
          for (int u = 0; u < kU; ++u)
            mq[u] = *reinterpret_cast<const Half8*>(p.mask + ru[u] * p.ldm + cv * 8);
        }
        if (MASK == 3) {
#pragma unroll
          for (int u = 0; u < kU; ++u) bits[u] = p.bitmask[ru[u] * m.vpr + cv];
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          float g[8], v[8];
          unpack8(xq[u], v);
          masked_grad<MASK>(gq[u], mq[u], bits[u], v, thr, neg_bits, g);
          if (ok[u]) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              a[i] += g[i];
              b[i] = fmaf(g[i], v[i] - mu[i], b[i]);
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) b[i] *= p.mean_rstd[C + cv * 8 + i];
    }
    fold_row_lanes(red, m, a, b);
    if (m.tr == 0 && cv < m.vpr) {
      float* pa = p.partial + static_cast<long long>(blockIdx.x) * 2 * C + cv * 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        pa[i] = a[i];
        pa[C + i] = b[i];
      }
    }
  }
  KFB_BN_STAMP(1);
  grid_barrier(p.sync, gridDim.x);
  KFB_BN_STAMP(2);

  // ---- phase 2: one warp per channel: dbeta / dgamma and the coefficients of
  //      dz = A*g + B*x + D with A = gamma*rstd, B = -A*rstd*dgamma/M,
  //      D = A*(rstd*mean*dgamma/M - dbeta/M) ----
  for (int c = blockIdx.x * kBnWarps + warp; c < C; c += gridDim.x * kBnWarps) {
    float db = 0.f, dg = 0.f;
    {
      // all loads of a lane are issued before the first add (one L2 round trip, fixed order)
      constexpr int kF = 8;
      const int nb = static_cast<int>(gridDim.x);
      for (int b0 = lane; b0 < nb; b0 += 32 * kF) {
        float va[kF], vb[kF];
#pragma unroll
        for (int u = 0; u < kF; ++u) {
          const int b = b0 + 32 * u;
          const float* q = p.partial + static_cast<long long>(b < nb ? b : lane) * 2 * C + c;
          va[u] = __ldcg(q);
          vb[u] = __ldcg(q + C);
          if (b >= nb) va[u] = vb[u] = 0.f;
        }
#pragma unroll
        for (int u = 0; u < kF; ++u) {
          db += va[u];
          dg += vb[u];
        }
      }
    }
    db = warp_sum_det(db);
    dg = warp_sum_det(dg);
    if (lane == 0) {
      if (p.dgamma) p.dgamma[c] = (p.accumulate ? p.dgamma[c] : 0.f) + dg * p.grad_scale;
      if (p.dbeta) p.dbeta[c] = (p.accumulate ? p.dbeta[c] : 0.f) + db * p.grad_scale;
      const float mean = p.mean_rstd[c], rstd = p.mean_rstd[C + c];
      const float A = (p.gamma ? p.gamma[c] : 1.f) * rstd;
      const float c1 = db * p.inv_count, c2 = dg * p.inv_count;
      p.coef[c] = A;
      p.coef[C + c] = -A * rstd * c2;
      p.coef[2 * C + c] = A * (rstd * mean * c2 - c1);
    }
  }
  if (p.dz == nullptr) return;  // only the parameter gradients were wanted
  KFB_BN_STAMP(3);
  grid_barrier(p.sync, gridDim.x);
  KFB_BN_STAMP(4);

  // ---- phase 3: dz = A*g + B*x + D (and d_res (+)= g); the slab is walked last-to-first so
  //      its tail, read last in phase 1, is served from L2 ----
  for (int cv0 = 0; cv0 < m.vpr; cv0 += m.cv_per_pass) {
    const int cv = cv0 + m.tv;
    if (!(cv < m.vpr && m.tr < m.row_lanes)) continue;
    float A[8], B[8], D[8], thr[8];
    unsigned neg_bits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      A[i] = __ldcg(p.coef + cv * 8 + i);
      B[i] = __ldcg(p.coef + C + cv * 8 + i);
      D[i] = __ldcg(p.coef + 2 * C + cv * 8 + i);
      thr[i] = 0.f;
    }
    if (MASK == 2) relu_thresholds(p.relu_affine, C, cv * 8, thr, neg_bits);
    const long long n_mine = m.r_end > m.r_begin + m.tr
                                 ? (m.r_end - m.r_begin - m.tr + m.row_lanes - 1) / m.row_lanes
                                 : 0;
    for (long long k0 = 0; k0 < n_mine; k0 += kU) {
      long long ru[kU];
      bool ok[kU];
      Half8 gq[kU], xq[kU], mq[kU], oq[kU];
      unsigned bits[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long k = n_mine - 1 - (k0 + u);
        ok[u] = k >= 0;
        ru[u] = m.r_begin + m.tr + (ok[u] ? k : 0) * m.row_lanes;
        gq[u] = *reinterpret_cast<const Half8*>(p.dy + ru[u] * p.lddy + cv * 8);
        bits[u] = 0;
      }
#pragma unroll
      for (int u = 0; u < kU; ++u)
        xq[u] = *reinterpret_cast<const Half8*>(p.x + ru[u] * p.ldx + cv * 8);
      if (MASK == 1) {
#pragma unroll
        for (int u = 0; u < kU; ++u)
          mq[u] = *reinterpret_cast<const Half8*>(p.mask + ru[u] * p.ldm + cv * 8);
      }
      if (MASK == 3) {
#pragma unroll
        for (int u = 0; u < kU; ++u) bits[u] = p.bitmask[ru[u] * m.vpr + cv];
      }
      if (p.dres != nullptr && p.dres_accumulate) {
#pragma unroll
        for (int u = 0; u < kU; ++u)
          oq[u] = *reinterpret_cast<const Half8*>(p.dres + ru[u] * p.lddr + cv * 8);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        float g[8], v[8], out[8];
        unpack8(xq[u], v);
        masked_grad<MASK>(gq[u], mq[u], bits[u], v, thr, neg_bits, g);
#pragma unroll
        for (int i = 0; i < 8; ++i) out[i] = fmaf(A[i], g[i], fmaf(B[i], v[i], D[i]));
        if (!ok[u]) continue;
        store8(p.dz + ru[u] * p.lddz + cv * 8, out);
        if (p.dres != nullptr) {
          if (p.dres_accumulate) {
            float o[8];
            unpack8(oq[u], o);
#pragma unroll
            for (int i = 0; i < 8; ++i) g[i] += o[i];
          }
          store8(p.dres + ru[u] * p.lddr + cv * 8, g);
        }
      }
    }
  }
  KFB_BN_STAMP(5);
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
constexpr int kBnSmemBytes = kBnThreads * 16 * sizeof(float);

// Largest grid whose CTAs are all resident, capped at KFB_BN_CTAS_PER_SM (default 3) per SM so
// that CTAs of a concurrently running exchange kernel still find room.
template <typename K>
static int resident_grid(K kernel, long long rows) {
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kBnThreads, kBnSmemBytes) !=
          cudaSuccess ||
      per_sm < 1)
    return 0;
  static int cap = -1;
  if (cap < 0) {
    const char* e = getenv("KFB_BN_CTAS_PER_SM");
    cap = (e != nullptr && atoi(e) > 0) ? atoi(e) : 8;
  }
  if (per_sm > cap) per_sm = cap;
  long long g = static_cast<long long>(sms) * per_sm;
  const long long by_rows = (rows + 31) / 32;  // at least 32 rows per CTA
  if (g > by_rows) g = by_rows;
  return g < 1 ? 1 : static_cast<int>(g);
}

static unsigned long long* g_bn_dbg = nullptr;

}  // namespace kfb

using namespace kfb;

// debugging aid (tools/bn_microbench.py --phases): six device words that receive the
// globaltimer of CTA 0 at the phase boundaries of the next backward kernels; null disables
extern "C" int kfb_bn_fused_debug_stamps(unsigned long long* dev_words) {
  g_bn_dbg = dev_words;
  return 0;
}

extern "C" int kfb_bn_fused_max_grid(void) {
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return (sms > 0 ? sms : 148) * 8;
}

extern "C" int kfb_bn_forward_train_fused(
    const void* x, long long ldx, long long rows, int C, const float* gamma, const float* beta,
    float* moving_mean, float* moving_var, float eps, float decay, int update_moving,
    float* mean_rstd, float* scale_shift, const float* stats, int stats_rows, int stats_bn,
    float* partial, const void* residual, long long ldr, int relu, void* y, long long ldy,
    unsigned char* bitmask, unsigned* sync, void* stream) {
  KFB_REQUIRE(C > 0 && C % 8 == 0, "bn_forward_train_fused: C (%d) must be a multiple of 8", C);
  KFB_REQUIRE(sync != nullptr, "bn_forward_train_fused: needs the two barrier words");
  KFB_REQUIRE(stats_rows == 0 || (stats != nullptr && stats_bn > 0),
              "bn_forward_train_fused: bad statistics layout");
  // 4 rows in flight per thread at 2 CTAs per SM (tools/bn_microbench.py: 2 and 8 rows and
  // 1 / 3 / 4 CTAs per SM were equal or slower)
  const int grid = resident_grid(bn_fwd_fused_kernel<4, 2>, rows);
  KFB_REQUIRE(grid > 0, "bn_forward_train_fused: occupancy query failed");
  BnFwdArgs a;
  a.x = (const __half*)x; a.ldx = ldx;
  a.res = (const __half*)residual; a.ldr = ldr;
  a.y = (__half*)y; a.ldy = ldy;
  a.bitmask = bitmask;
  a.rows = rows; a.C = C;
  a.gamma = gamma; a.beta = beta;
  a.moving_mean = moving_mean; a.moving_var = moving_var;
  a.eps = eps; a.decay = decay;
  a.inv_count = (float)(1.0 / (double)rows);
  a.unbias = rows > 1 ? (float)((double)rows / (double)(rows - 1)) : 1.f;
  a.update_moving = update_moving; a.relu = relu;
  a.mean_rstd = mean_rstd; a.scale_shift = scale_shift;
  a.stats = stats; a.stats_rows = stats_rows; a.stats_bn = stats_bn;
  a.partial = partial;
  a.sync = sync;
  a.dbg = nullptr;
  KFB_REQUIRE(stats_rows > 0 || partial != nullptr,
              "bn_forward_train_fused: needs scratch for its own statistics pass");
  bn_fwd_fused_kernel<4, 2><<<grid, kBnThreads, kBnSmemBytes, (cudaStream_t)stream>>>(a);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_bn_backward_fused(
    const void* dy, long long lddy, const void* mask, long long ldm, const unsigned char* bitmask,
    const float* relu_affine, const void* x, long long ldx, long long rows, int C,
    const float* gamma, const float* mean_rstd, float* dgamma, float* dbeta,
    int accumulate_param_grads, float grad_scale, float* coef, float* partial, void* dz,
    long long lddz, void* dres, long long lddr, int dres_accumulate, unsigned* sync,
    void* stream) {
  KFB_REQUIRE(C > 0 && C % 8 == 0, "bn_backward_fused: C (%d) must be a multiple of 8", C);
  KFB_REQUIRE(sync != nullptr && partial != nullptr && coef != nullptr,
              "bn_backward_fused: needs scratch and the two barrier words");
  const int mode = mask != nullptr ? 1 : (bitmask != nullptr ? 3 : (relu_affine != nullptr ? 2 : 0));
  BnBwdArgs a;
  a.dy = (const __half*)dy; a.lddy = lddy;
  a.mask = (const __half*)mask; a.ldm = ldm;
  a.bitmask = bitmask;
  a.relu_affine = relu_affine;
  a.x = (const __half*)x; a.ldx = ldx;
  a.rows = rows; a.C = C;
  a.gamma = gamma; a.mean_rstd = mean_rstd;
  a.dgamma = dgamma; a.dbeta = dbeta;
  a.accumulate = accumulate_param_grads; a.grad_scale = grad_scale;
  a.inv_count = (float)(1.0 / (double)rows);
  a.coef = coef; a.partial = partial;
  a.dz = (__half*)dz; a.lddz = lddz;
  a.dres = (__half*)dres; a.lddr = lddr; a.dres_accumulate = dres_accumulate;
  a.sync = sync;
  a.dbg = g_bn_dbg;
  cudaStream_t st = (cudaStream_t)stream;
  int grid = 0;
  // 4 rows x 16 bytes per tensor in flight per thread at 2 CTAs per SM measured best (tools/
  // bn_microbench.py: 1, 2, 6 and 8 rows and 1 / 3 / 4 / 6 CTAs per SM were all slower)
#define KFB_BN_LAUNCH_V(MASK)                                                                  \
  do {                                                                                         \
    grid = resident_grid(bn_bwd_fused_kernel<MASK, 4, 2>, rows);                               \
    if (grid > 0) bn_bwd_fused_kernel<MASK, 4, 2><<<grid, kBnThreads, kBnSmemBytes, st>>>(a);  \
  } while (0)
  switch (mode) {
    case 0: KFB_BN_LAUNCH_V(0); break;
    case 1: KFB_BN_LAUNCH_V(1); break;
    case 2: KFB_BN_LAUNCH_V(2); break;
    default: KFB_BN_LAUNCH_V(3); break;
  }
  KFB_REQUIRE(grid > 0, "bn_backward_fused: occupancy query failed");
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
