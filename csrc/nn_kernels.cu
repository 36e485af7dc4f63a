// Memory-bound layer kernels for sm_100a: batch-norm (stats / finalize / fused
// apply+ReLU+residual, backward reduce + apply), pooling, spatial mean,
// bias/ReLU backward, strided copies (concat / split), explicit im2col for
// tiny-C stems, dropout, LRN and the fused softmax-cross-entropy(+top-k).
//
// All tensors are NHWC fp16 viewed as [rows = N*H*W][C] with an explicit row
// stride (ld) so producers can write straight into channel slices of a concat
// buffer. Channel counts must be multiples of 8 (16-byte vectors).
// Reference behaviour: ConvNetBuilder.batch_norm / mpool / apool / spatial_mean /
// dropout / lrn (T/convnet_builder_legacy.py:228-468) and CNNModel.loss_function /
// accuracy_function (T/models/model.py:285-312).
#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

constexpr int kRedThreads = 256;

// Generic per-channel two-value reduction over rows.
//   MODE 0 (bn stats)      : a = x,            b = x*x
//   MODE 1 (bn backward)   : g = dy*(mask>0);  a = g, b = g * (x - mean) * rstd
//   MODE 2 (bias backward) : g = dy*(mask>0);  a = g (b unused); also writes g to dz
// Stage 1 writes partial[block][2][C]; stage 2 (finalize kernels) folds the blocks in a
// fixed order, so results are run-to-run deterministic.
template <int MODE>
__global__ void __launch_bounds__(kRedThreads)
    channel_reduce_kernel(const __half* __restrict__ x, long long ldx,
                          const __half* __restrict__ dy, long long lddy,
                          const __half* __restrict__ mask, long long ldm,
                          const float* __restrict__ mean_rstd,  // [2][C] (MODE 1)
                          const float* __restrict__ relu_affine,  // [2][C] or null (MODE 1)
                          __half* __restrict__ dz, long long lddz, long long rows, int C,
                          float* __restrict__ partial) {
  extern __shared__ float red[];  // [kRedThreads][16] worst case handled by looping
  const int vpr = C >> 3;         // vectors per row
  // thread -> (row lane, channel vector); channel vectors beyond 256 are looped
  const int cv_per_pass = vpr < kRedThreads ? vpr : kRedThreads;
  const int row_lanes = kRedThreads / cv_per_pass;
  const int tv = threadIdx.x % cv_per_pass;
  const int tr = threadIdx.x / cv_per_pass;
  const long long rows_per_block = (rows + gridDim.x - 1) / gridDim.x;
  // MODE 1 (BN backward): slabs are handed out last-to-first so the blocks scheduled first read
  // the end of dy, which the producing dgrad wrote last (L2 reuse); the apply pass that follows
  // runs first-to-last and finds the beginning fresh.
  const long long slab = (MODE == 1) ? (gridDim.x - 1 - blockIdx.x) : blockIdx.x;
  const long long r_begin = slab * rows_per_block;
  long long r_end = r_begin + rows_per_block;
  if (r_end > rows) r_end = rows;

  for (int cv0 = 0; cv0 < vpr; cv0 += cv_per_pass) {
    const int cv = cv0 + tv;
    float a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = b[i] = 0.f;
    // MODE 1 per-channel constants. The ReLU mask of y = relu(x*sc + sh) is recomputed from x
    // as (x > thr) xor (sc < 0) with thr = -sh/sc, so y does not have to be re-read.
    float mu[8], thr[8];
    unsigned neg_bits = 0;
    const bool recompute_mask = MODE == 1 && relu_affine != nullptr;
    const bool active = cv < vpr && tr < row_lanes;
    if (MODE == 1 && active) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        mu[i] = mean_rstd[cv * 8 + i];
        thr[i] = 0.f;
        if (recompute_mask) {
          const float sc = relu_affine[cv * 8 + i], sh = relu_affine[C + cv * 8 + i];
          if (sc > 0.f) {
            thr[i] = -sh / sc;
          } else if (sc < 0.f) {
            thr[i] = -sh / sc;
            neg_bits |= 1u << i;
          } else {
            thr[i] = sh > 0.f ? -INFINITY : INFINITY;  // constant output
          }
        }
      }
    }
    if (active) {
      // kRowUnroll rows per iteration, all their 16-byte loads issued before the first use:
      // the kernel is bound by bytes in flight per SM (512-1024 resident threads), not by math.
      constexpr int kRowUnroll = 4;
      if (MODE == 0) {
        long long r = r_begin + tr;
        for (; r + (kRowUnroll - 1) * row_lanes < r_end; r += kRowUnroll * row_lanes) {
          Half8 q[kRowUnroll];
#pragma unroll
          for (int u = 0; u < kRowUnroll; ++u)
            q[u] = *reinterpret_cast<const Half8*>(x + (r + u * row_lanes) * ldx + cv * 8);
#pragma unroll
          for (int u = 0; u < kRowUnroll; ++u) {
            float v[8];
            unpack8(q[u], v);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              a[i] += v[i];
              b[i] += v[i] * v[i];
            }
          }
        }
        for (; r < r_end; r += row_lanes) {
          float v[8];
          load8(x + r * ldx + cv * 8, v);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            a[i] += v[i];
            b[i] += v[i] * v[i];
          }
        }
      } else {
        for (long long r = r_begin + tr; r < r_end; r += kRowUnroll * row_lanes) {
          long long ru[kRowUnroll];
          bool ok[kRowUnroll];
          Half8 gq[kRowUnroll], xq[kRowUnroll], mq[kRowUnroll];
#pragma unroll
          for (int u = 0; u < kRowUnroll; ++u) {
            ru[u] = r + static_cast<long long>(u) * row_lanes;
            ok[u] = ru[u] < r_end;
            if (!ok[u]) ru[u] = r;  // any valid row; its contribution is zeroed below
            gq[u] = *reinterpret_cast<const Half8*>(dy + ru[u] * lddy + cv * 8);
          }
          if (MODE == 1) {
#pragma unroll
            for (int u = 0; u < kRowUnroll; ++u)
              xq[u] = *reinterpret_cast<const Half8*>(x + ru[u] * ldx + cv * 8);
          }
          if (mask != nullptr) {
#pragma unroll
            for (int u = 0; u < kRowUnroll; ++u)
              mq[u] = *reinterpret_cast<const Half8*>(mask + ru[u] * ldm + cv * 8);
          }
#pragma unroll
          for (int u = 0; u < kRowUnroll; ++u) {
            float g[8], v[8];
            unpack8(gq[u], g);
            if (MODE == 1) unpack8(xq[u], v);
            if (mask != nullptr) {
              float m[8];
              unpack8(mq[u], m);
#pragma unroll
              for (int i = 0; i < 8; ++i) g[i] = m[i] > 0.f ? g[i] : 0.f;
            } else if (recompute_mask) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const bool ng = (neg_bits >> i) & 1u;
                g[i] = ((v[i] > thr[i]) != ng) ? g[i] : 0.f;
              }
            }
            if (!ok[u]) {
#pragma unroll
              for (int i = 0; i < 8; ++i) g[i] = 0.f;
            }
            if (MODE == 1) {
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                a[i] += g[i];
                b[i] += g[i] * (v[i] - mu[i]);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 8; ++i) a[i] += g[i];
              if (dz != nullptr && ok[u]) store8(dz + ru[u] * lddz + cv * 8, g);
            }
          }
        }
        if (MODE == 1) {
#pragma unroll
          for (int i = 0; i < 8; ++i) b[i] *= mean_rstd[C + cv * 8 + i];
        }
      }
    }
    // fold the row lanes through shared memory (fixed order)
    __syncthreads();
    float* mine = red + threadIdx.x * 16;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      mine[i] = a[i];
      mine[8 + i] = b[i];
    }
    __syncthreads();
    if (tr == 0 && cv < vpr) {
      float sa[8], sb[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) sa[i] = sb[i] = 0.f;
      for (int l = 0; l < row_lanes; ++l) {
        const float* o = red + (l * cv_per_pass + tv) * 16;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          sa[i] += o[i];
          sb[i] += o[8 + i];
        }
      }
      float* pa = partial + static_cast<long long>(slab) * 2 * C + cv * 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        pa[i] = sa[i];
        pa[C + i] = sb[i];
      }
    }
  }
}

// Folds the stage-1 partials of channel c in a fixed order. Block = (32 channels, 32 lanes):
// lane ty sums blocks ty, ty+32, ... (coalesced 128-byte rows) through four independent
// accumulators so the L2 round trips overlap, then lane 0 adds the lane sums.
constexpr int kFoldLanes = 32;
__device__ __forceinline__ bool fold_partials(const float* __restrict__ partial, int nblocks,
                                              int C, int c, float& s0, float& s1) {
  __shared__ float sh[2][kFoldLanes][32];
  const int tx = threadIdx.x, ty = threadIdx.y;
  float a = 0.f, b = 0.f;
  if (c < C) {
    float a4[4] = {0.f, 0.f, 0.f, 0.f}, b4[4] = {0.f, 0.f, 0.f, 0.f};
    const long long row = 2ll * C;
    int blk = ty;
    for (; blk + 3 * kFoldLanes < nblocks; blk += 4 * kFoldLanes) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* q = partial + static_cast<long long>(blk + u * kFoldLanes) * row + c;
        a4[u] += q[0];
        b4[u] += q[C];
      }
    }
    for (int u = 0; blk < nblocks; blk += kFoldLanes, ++u) {
      const float* q = partial + static_cast<long long>(blk) * row + c;
      a4[u] += q[0];
      b4[u] += q[C];
    }
    a = (a4[0] + a4[1]) + (a4[2] + a4[3]);
    b = (b4[0] + b4[1]) + (b4[2] + b4[3]);
  }
  sh[0][ty][tx] = a;
  sh[1][ty][tx] = b;
  __syncthreads();
  if (ty != 0 || c >= C) return false;
  s0 = 0.f;
  s1 = 0.f;
#pragma unroll
  for (int l = 0; l < kFoldLanes; ++l) {
    s0 += sh[0][l][tx];
    s1 += sh[1][l][tx];
  }
  return true;
}

// BN forward finalize: mean/var from partial sums, moving-average update, and the
// per-channel affine (scale, shift) the apply kernel uses.
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int nblocks, int C,
                                   float inv_count, float unbias, float eps, float decay,
                                   const float* __restrict__ gamma,  // may be null (no scale)
                                   const float* __restrict__ beta, float* __restrict__ moving_mean,
                                   float* __restrict__ moving_var,
                                   float* __restrict__ mean_rstd,     // [2][C] saved for bwd
                                   float* __restrict__ scale_shift,   // [2][C]
                                   int update_moving) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s, ss;
  if (!fold_partials(partial, nblocks, C, c, s, ss)) return;
  const float mean = s * inv_count;
  float var = ss * inv_count - mean * mean;
  var = var < 0.f ? 0.f : var;
  const float rstd = rsqrtf(var + eps);
  mean_rstd[c] = mean;
  mean_rstd[C + c] = rstd;
  const float g = gamma ? gamma[c] : 1.f;
  const float sc = g * rstd;
  scale_shift[c] = sc;
  scale_shift[C + c] = beta[c] - mean * sc;
  if (update_moving) {
    moving_mean[c] = moving_mean[c] * decay + mean * (1.f - decay);
    moving_var[c] = moving_var[c] * decay + var * unbias * (1.f - decay);
  }
}

// Inference-mode affine from the moving statistics.
__global__ void bn_inference_affine_kernel(int C, float eps, const float* __restrict__ gamma,
                                           const float* __restrict__ beta,
                                           const float* __restrict__ moving_mean,
                                           const float* __restrict__ moving_var,
                                           float* __restrict__ scale_shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = (gamma ? gamma[c] : 1.f) * rsqrtf(moving_var[c] + eps);
  scale_shift[c] = sc;
  scale_shift[C + c] = beta[c] - moving_mean[c] * sc;
}

// y = act(x * scale + shift (+ residual)); act = ReLU if relu.
template <int kU>
__global__ void __launch_bounds__(256)
    bn_apply_kernel(const __half* __restrict__ x, long long ldx,
                    const float* __restrict__ scale_shift, const __half* __restrict__ res,
                    long long ldr, __half* __restrict__ y, long long ldy, long long rows, int C,
                    int relu, unsigned char* __restrict__ bitmask) {
  const int vpr = C >> 3;
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  if (cv >= vpr) return;
  float sc[8], sh[8];
  {
    const float4 s0 = *reinterpret_cast<const float4*>(scale_shift + cv * 8);
    const float4 s1 = *reinterpret_cast<const float4*>(scale_shift + cv * 8 + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(scale_shift + C + cv * 8);
    const float4 b1 = *reinterpret_cast<const float4*>(scale_shift + C + cv * 8 + 4);
    sc[0] = s0.x; sc[1] = s0.y; sc[2] = s0.z; sc[3] = s0.w;
    sc[4] = s1.x; sc[5] = s1.y; sc[6] = s1.z; sc[7] = s1.w;
    sh[0] = b0.x; sh[1] = b0.y; sh[2] = b0.z; sh[3] = b0.w;
    sh[4] = b1.x; sh[5] = b1.y; sh[6] = b1.z; sh[7] = b1.w;
  }
  // rows are visited last-to-first: the end of x was written most recently by the producing
  // conv and is still in L2; the next conv then finds the beginning of y fresh
  auto finish = [&](long long r, const Half8& xv, const Half8& rv) {
    float v[8];
    unpack8(xv, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = fmaf(v[k], sc[k], sh[k]);
    if (res != nullptr) {
      float rr[8];
      unpack8(rv, rr);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += rr[k];
    }
    if (relu) {
      unsigned bits = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        bits |= (v[k] > 0.f ? 1u : 0u) << k;
        v[k] = fmaxf(v[k], 0.f);
      }
      if (bitmask != nullptr) bitmask[r * vpr + cv] = static_cast<unsigned char>(bits);
    }
    store8(y + r * ldy + cv * 8, v);
  };
  const long long stride = static_cast<long long>(gridDim.y) * blockDim.y;
  long long rr = static_cast<long long>(blockIdx.y) * blockDim.y + threadIdx.y;
  // kU rows per iteration: all loads are issued before the first use (the bytes in flight
  // per SM, not the thread count, set the achieved HBM bandwidth of a streaming kernel)
  for (; kU > 1 && rr + (kU - 1) * stride < rows; rr += kU * stride) {
    Half8 xv[kU], rv[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long r = rows - 1 - (rr + u * stride);
      xv[u] = *reinterpret_cast<const Half8*>(x + r * ldx + cv * 8);
      if (res != nullptr) rv[u] = *reinterpret_cast<const Half8*>(res + r * ldr + cv * 8);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) finish(rows - 1 - (rr + u * stride), xv[u], rv[u]);
  }
  for (; rr < rows; rr += stride) {
    const long long r = rows - 1 - rr;
    Half8 xv, rv;
    xv = *reinterpret_cast<const Half8*>(x + r * ldx + cv * 8);
    if (res != nullptr) rv = *reinterpret_cast<const Half8*>(res + r * ldr + cv * 8);
    finish(r, xv, rv);
  }
}

// BN forward finalize from the conv-epilogue statistics (layout of kfb_gemm_stats_layout: for
// channel c = nb*bn + j the partial rows are stats[((nb*nrows + r)*2 + s)*bn + j]). Block =
// (32 channels, 32 row lanes), like fold_partials.
__global__ void bn_finalize_stats_kernel(const float* __restrict__ stats, int nrows, int bn, int C,
                                         float inv_count, float unbias, float eps, float decay,
                                         const float* __restrict__ gamma,
                                         const float* __restrict__ beta,
                                         float* __restrict__ moving_mean,
                                         float* __restrict__ moving_var,
                                         float* __restrict__ mean_rstd,
                                         float* __restrict__ scale_shift, int update_moving) {
  __shared__ float sh[2][kFoldLanes][32];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int c = blockIdx.x * 32 + tx;
  float a = 0.f, b = 0.f;
  if (c < C) {
    const int nb = c / bn, j = c - nb * bn;
    const float* base = stats + static_cast<long long>(nb) * nrows * 2 * bn + j;
    float a4[4] = {0.f, 0.f, 0.f, 0.f}, b4[4] = {0.f, 0.f, 0.f, 0.f};
    int r = ty;
    for (; r + 3 * kFoldLanes < nrows; r += 4 * kFoldLanes) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float* q = base + static_cast<long long>(r + u * kFoldLanes) * 2 * bn;
        a4[u] += q[0];
        b4[u] += q[bn];
      }
    }
    for (int u = 0; r < nrows; r += kFoldLanes, ++u) {
      const float* q = base + static_cast<long long>(r) * 2 * bn;
      a4[u] += q[0];
      b4[u] += q[bn];
    }
    a = (a4[0] + a4[1]) + (a4[2] + a4[3]);
    b = (b4[0] + b4[1]) + (b4[2] + b4[3]);
  }
  sh[0][ty][tx] = a;
  sh[1][ty][tx] = b;
  __syncthreads();
  if (ty != 0 || c >= C) return;
  float s = 0.f, ss = 0.f;
#pragma unroll
  for (int l = 0; l < kFoldLanes; ++l) {
    s += sh[0][l][tx];
    ss += sh[1][l][tx];
  }
  const float mean = s * inv_count;
  float var = ss * inv_count - mean * mean;
  var = var < 0.f ? 0.f : var;
  const float rstd = rsqrtf(var + eps);
  mean_rstd[c] = mean;
  mean_rstd[C + c] = rstd;
  const float g = gamma ? gamma[c] : 1.f;
  const float sc = g * rstd;
  scale_shift[c] = sc;
  scale_shift[C + c] = beta[c] - mean * sc;
  if (update_moving) {
    moving_mean[c] = moving_mean[c] * decay + mean * (1.f - decay);
    moving_var[c] = moving_var[c] * decay + var * unbias * (1.f - decay);
  }
}

// BN backward finalize: dgamma/dbeta (optionally accumulated into the gradient arena) and
// the per-channel coefficients of the apply pass, dz = A*g + B*x + D with
//   A = gamma*rstd, B = -A*rstd*dgamma/M, D = A*(rstd*mean*dgamma/M - dbeta/M).
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int nblocks, int C,
                                       float inv_count, const float* __restrict__ gamma,
                                       const float* __restrict__ mean_rstd,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       int accumulate, float grad_scale,
                                       float* __restrict__ coef /* [3][C]: A, B, D */) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  float db, dg;
  if (!fold_partials(partial, nblocks, C, c, db, dg)) return;
  if (dgamma) dgamma[c] = (accumulate ? dgamma[c] : 0.f) + dg * grad_scale;
  if (dbeta) dbeta[c] = (accumulate ? dbeta[c] : 0.f) + db * grad_scale;
  const float mean = mean_rstd[c], rstd = mean_rstd[C + c];
  const float A = (gamma ? gamma[c] : 1.f) * rstd;
  const float c1 = db * inv_count, c2 = dg * inv_count;
  coef[c] = A;
  coef[C + c] = -A * rstd * c2;
  coef[2 * C + c] = A * (rstd * mean * c2 - c1);
}

// dz = A*g + B*x + D, g = dy * (mask > 0); optionally d_res (+)= g.
__device__ __forceinline__ void load8f(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w;
  f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__global__ void __launch_bounds__(256)
    bn_bwd_apply_kernel(const __half* __restrict__ dy, long long lddy,
                        const __half* __restrict__ mask, long long ldm,
                        const __half* __restrict__ x, long long ldx,
                        const float* __restrict__ coef, const float* __restrict__ relu_affine,
                        __half* __restrict__ dz, long long lddz, __half* __restrict__ dres,
                        long long lddr, int dres_accumulate, long long rows, int C) {
  const int vpr = C >> 3;
  // blockDim = (vx, vy): vx channel vectors x vy rows; channel coefficients live in registers
  const int cv = blockIdx.x * blockDim.x + threadIdx.x;
  if (cv >= vpr) return;
  float A[8], B[8], D[8], thr[8];
  unsigned neg_bits = 0;
  load8f(coef + cv * 8, A);
  load8f(coef + C + cv * 8, B);
  load8f(coef + 2 * C + cv * 8, D);
  if (relu_affine != nullptr) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float sc = relu_affine[cv * 8 + k], sh = relu_affine[C + cv * 8 + k];
      thr[k] = sc != 0.f ? -sh / sc : (sh > 0.f ? -INFINITY : INFINITY);
      if (sc < 0.f) neg_bits |= 1u << k;
    }
  }
  for (long long r = static_cast<long long>(blockIdx.y) * blockDim.y + threadIdx.y; r < rows;
       r += static_cast<long long>(gridDim.y) * blockDim.y) {
    float g[8], v[8];
    load8(dy + r * lddy + cv * 8, g);
    load8(x + r * ldx + cv * 8, v);
    if (mask != nullptr) {
      float m[8];
      load8(mask + r * ldm + cv * 8, m);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = m[k] > 0.f ? g[k] : 0.f;
    } else if (relu_affine != nullptr) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        g[k] = ((v[k] > thr[k]) != (((neg_bits >> k) & 1u) != 0)) ? g[k] : 0.f;
    }
    float out[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) out[k] = fmaf(A[k], g[k], fmaf(B[k], v[k], D[k]));
    store8(dz + r * lddz + cv * 8, out);
    if (dres != nullptr) {
      if (dres_accumulate) {
        float o[8];
        load8(dres + r * lddr + cv * 8, o);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] += o[k];
      }
      store8(dres + r * lddr + cv * 8, g);
    }
  }
}

// Fold bias-gradient partials: db[c] (+)= scale * sum_b partial[b][0][c]
__global__ void colsum_finalize_kernel(const float* __restrict__ partial, int nblocks, int C,
                                       float* __restrict__ out, int accumulate, float scale) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s, unused;
  if (!fold_partials(partial, nblocks, C, c, s, unused)) return;
  out[c] = (accumulate ? out[c] : 0.f) + s * scale;
}

// ---------------------------------------------------------------------------
// Elementwise: out = act(a (+ b)); act: 0 none, 1 relu. Also used for grad accumulation.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    add_act_kernel(const __half* __restrict__ a, long long lda, const __half* __restrict__ b,
                   long long ldb, __half* __restrict__ y, long long ldy, long long rows, int C,
                   int relu) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float v[8];
    load8(a + r * lda + cv * 8, v);
    if (b != nullptr) {
      float w[8];
      load8(b + r * ldb + cv * 8, w);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += w[k];
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
    }
    store8(y + r * ldy + cv * 8, v);
  }
}

// dx (+)= dy * (mask > 0)
__global__ void __launch_bounds__(256)
    relu_bwd_kernel(const __half* __restrict__ dy, long long lddy, const __half* __restrict__ mask,
                    long long ldm, __half* __restrict__ dx, long long lddx, int accumulate,
                    long long rows, int C) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float g[8];
    load8(dy + r * lddy + cv * 8, g);
    if (mask != nullptr) {
      float m[8];
      load8(mask + r * ldm + cv * 8, m);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = m[k] > 0.f ? g[k] : 0.f;
    }
    if (accumulate) {
      float o[8];
      load8(dx + r * lddx + cv * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += o[k];
    }
    store8(dx + r * lddx + cv * 8, g);
  }
}

// ---------------------------------------------------------------------------
// Pooling (NHWC). Max pooling stores the arg-max tap (uint8) for the backward pass.
// Average pooling divides by the number of in-bounds taps (TF 'SAME' semantics).
// ---------------------------------------------------------------------------
struct PoolGeom {
  int n, h, w, c, p, q, kh, kw, sh, sw, pad_h, pad_w;
};

// TK / TS > 0 fix the (square) window and stride at compile time: the 3x3/2 and 2x2/2 pools of
// the zoo then run without integer divisions and with fully unrolled window loops.
template <int TK, int TS>
__global__ void __launch_bounds__(256)
    pool_fwd_kernel(const __half* __restrict__ x, long long ldx, __half* __restrict__ y,
                    long long ldy, unsigned char* __restrict__ argmax, PoolGeom gg, int is_max) {
  PoolGeom g = gg;
  if (TK > 0) {
    g.kh = g.kw = TK;
    g.sh = g.sw = TS;
  }
  const int vpr = g.c >> 3;
  // grid = (column x channel-vector chunks, output rows, images): no division of a flat index
  for (int n = blockIdx.z; n < g.n; n += gridDim.z)
  for (int p = blockIdx.y; p < g.p; p += gridDim.y)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < g.q * vpr;
       t += gridDim.x * blockDim.x) {
    const int q = t / vpr;
    const int cv = t - q * vpr;
    const int h0 = p * g.sh - g.pad_h, w0 = q * g.sw - g.pad_w;
    float acc[8];
    int arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      acc[k] = is_max ? -65504.f : 0.f;
      arg[k] = 0;
    }
    int count = 0;
#pragma unroll
    for (int r = 0; r < (TK > 0 ? TK : g.kh); ++r) {
      const int hh = h0 + r;
      if (hh < 0 || hh >= g.h) continue;
#pragma unroll
      for (int s = 0; s < (TK > 0 ? TK : g.kw); ++s) {
        const int ww = w0 + s;
        if (ww < 0 || ww >= g.w) continue;
        float v[8];
        load8(x + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * ldx + cv * 8, v);
        if (is_max) {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (v[k] > acc[k] || count == 0) {
              acc[k] = v[k];
              arg[k] = r * g.kw + s;
            }
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += v[k];
        }
        ++count;
      }
    }
    if (!is_max) {
      const float inv = count > 0 ? 1.f / count : 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] *= inv;
    }
    const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
    store8(y + orow * ldy + cv * 8, acc);
    if (is_max && argmax != nullptr) {
      uint2 packed;
      packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
      packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
      *reinterpret_cast<uint2*>(argmax + orow * g.c + cv * 8) = packed;
    }
  }
}

// Max pooling forward with a compile-time window, packed: the eight channels of a vector stay
// four half2 words - HSETP2 mask, HMNMX2 and one LOP3 for the arg-max lanes per tap and word,
// no fp32 conversions. The scalar kernel above spends ~300 instructions per output vector of a
// 3x3 window and is issue-bound at a third of the HBM roof on ResNet's stem pool.
template <int TK, int TS>
__global__ void __launch_bounds__(256)
    pool_fwd_max_packed_kernel(const __half* __restrict__ x, long long ldx,
                               __half* __restrict__ y, long long ldy,
                               unsigned char* __restrict__ argmax, PoolGeom g) {
  const int vpr = g.c >> 3;
  for (int n = blockIdx.z; n < g.n; n += gridDim.z)
  for (int p = blockIdx.y; p < g.p; p += gridDim.y)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < g.q * vpr;
       t += gridDim.x * blockDim.x) {
    const int q = t / vpr;
    const int cv = t - q * vpr;
    const int h0 = p * TS - g.pad_h, w0 = q * TS - g.pad_w;
    uint4 v[TK * TK];
    bool ok[TK * TK];
#pragma unroll
    for (int r = 0; r < TK; ++r) {
#pragma unroll
      for (int s = 0; s < TK; ++s) {
        const int hh = h0 + r, ww = w0 + s;
        ok[r * TK + s] = hh >= 0 && hh < g.h && ww >= 0 && ww < g.w;
        if (ok[r * TK + s])
          v[r * TK + s] = *reinterpret_cast<const uint4*>(
              x + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * ldx + cv * 8);
      }
    }
    unsigned acc[4], arg[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      acc[k] = 0xFC00FC00u;  // -inf, -inf
      arg[k] = 0u;
    }
    unsigned first = 0xFFFFFFFFu;  // the first in-bounds tap always wins (as the scalar kernel)
#pragma unroll
    for (int tap = 0; tap < TK * TK; ++tap) {
      if (!ok[tap]) continue;
      const unsigned w4[4] = {v[tap].x, v[tap].y, v[tap].z, v[tap].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const __half2 a = *reinterpret_cast<const __half2*>(&acc[k]);
        const __half2 b = *reinterpret_cast<const __half2*>(&w4[k]);
        const unsigned m = __hgt2_mask(b, a) | first;
        acc[k] = (acc[k] & ~m) | (w4[k] & m);
        arg[k] = (arg[k] & ~m) | ((static_cast<unsigned>(tap) * 0x00010001u) & m);
      }
      first = 0u;
    }
    const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
    *reinterpret_cast<uint4*>(y + orow * ldy + cv * 8) = make_uint4(acc[0], acc[1], acc[2], acc[3]);
    if (argmax != nullptr) {
      uint2 packed;
      packed.x = __byte_perm(arg[0], arg[1], 0x6420);
      packed.y = __byte_perm(arg[2], arg[3], 0x6420);
      *reinterpret_cast<uint2*>(argmax + orow * g.c + cv * 8) = packed;
    }
  }
}

// Gather formulation: one thread per input vector visits the output windows that cover it.
template <int TK, int TS>
__global__ void __launch_bounds__(256)
    pool_bwd_kernel(const __half* __restrict__ dy, long long lddy,
                    const unsigned char* __restrict__ argmax, __half* __restrict__ dx,
                    long long lddx, int accumulate, PoolGeom gg, int is_max) {
  PoolGeom g = gg;
  if (TK > 0) {
    g.kh = g.kw = TK;
    g.sh = g.sw = TS;
  }
  const int vpr = g.c >> 3;
  for (int n = blockIdx.z; n < g.n; n += gridDim.z)
  for (int hh = blockIdx.y; hh < g.h; hh += gridDim.y)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < g.w * vpr;
       t += gridDim.x * blockDim.x) {
    const int ww = t / vpr;
    const int cv = t - ww * vpr;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    // output rows p with p*sh - pad <= hh < p*sh - pad + kh
    int p_lo = (hh + g.pad_h - g.kh + g.sh) / g.sh;
    if (hh + g.pad_h - g.kh + 1 <= 0) p_lo = 0;
    int p_hi = (hh + g.pad_h) / g.sh;
    if (p_hi >= g.p) p_hi = g.p - 1;
    int q_lo = (ww + g.pad_w - g.kw + g.sw) / g.sw;
    if (ww + g.pad_w - g.kw + 1 <= 0) q_lo = 0;
    int q_hi = (ww + g.pad_w) / g.sw;
    if (q_hi >= g.q) q_hi = g.q - 1;
    __half* o = dx + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * lddx + cv * 8;
    if (is_max) {
      // Max pooling: the kernel is issue-bound, so stay in packed form. The 8 arg-max bytes
      // of a window are compared with the tap this pixel is in it by a SWAR "byte == 0" test,
      // widened to 16-bit lane masks with PRMT, ANDed onto the raw half2 gradient words and
      // summed with HADD2 (at most kh*kw/(sh*sw) terms, almost always 0 or 1 non-zero).
      __half2 hacc[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) hacc[k] = __floats2half2_rn(0.f, 0.f);
      if (TK > 0 && g.pad_h >= 0 && g.pad_w >= 0) {
        // Compile-time window: the covering windows are p0 - j (tap row r0 + j*TS), j = 0..J-1,
        // straight-line and predicated, so all their loads are in flight together.
        constexpr int TSS = TS > 0 ? TS : 1;
        constexpr int J = TK > 0 ? (TK + TSS - 1) / TSS : 1;
        const int ph = hh + g.pad_h, pw = ww + g.pad_w;
        const int p0 = ph / TSS, r0 = ph - p0 * TSS;
        const int q0 = pw / TSS, s0 = pw - q0 * TSS;
        uint4 dv[J * J];
        uint2 am[J * J];
        bool ok[J * J];
#pragma unroll
        for (int jp = 0; jp < J; ++jp) {
#pragma unroll
          for (int jq = 0; jq < J; ++jq) {
            const int p = p0 - jp, q = q0 - jq;
            const int idx = jp * J + jq;
            ok[idx] = (r0 + jp * TSS < TK) && (s0 + jq * TSS < TK) && p >= 0 && p < g.p &&
                      q >= 0 && q < g.q;
            dv[idx] = make_uint4(0u, 0u, 0u, 0u);
            am[idx] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
            if (ok[idx]) {
              const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
              dv[idx] = *reinterpret_cast<const uint4*>(dy + orow * lddy + cv * 8);
              am[idx] = *reinterpret_cast<const uint2*>(argmax + orow * g.c + cv * 8);
            }
          }
        }
#pragma unroll
        for (int jp = 0; jp < J; ++jp) {
#pragma unroll
          for (int jq = 0; jq < J; ++jq) {
            const int idx = jp * J + jq;
            const unsigned tapv =
                static_cast<unsigned>((r0 + jp * TSS) * TK + s0 + jq * TSS) * 0x01010101u;
            const unsigned e0 = am[idx].x ^ tapv, e1 = am[idx].y ^ tapv;
            const unsigned z0 = ~(((e0 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e0 | 0x7F7F7F7Fu);
            const unsigned z1 = ~(((e1 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e1 | 0x7F7F7F7Fu);
            const unsigned b0 = ok[idx] ? (z0 >> 7) * 0xFFu : 0u;
            const unsigned b1 = ok[idx] ? (z1 >> 7) * 0xFFu : 0u;
            unsigned w[4];
            w[0] = dv[idx].x & __byte_perm(b0, 0, 0x1100);
            w[1] = dv[idx].y & __byte_perm(b0, 0, 0x3322);
            w[2] = dv[idx].z & __byte_perm(b1, 0, 0x1100);
            w[3] = dv[idx].w & __byte_perm(b1, 0, 0x3322);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              hacc[k] = __hadd2(hacc[k], *reinterpret_cast<const __half2*>(&w[k]));
          }
        }
      } else
      for (int p = p_lo; p <= p_hi; ++p) {
        const int r = hh - (p * g.sh - g.pad_h);
        if (r < 0 || r >= g.kh) continue;
        for (int q = q_lo; q <= q_hi; ++q) {
          const int s = ww - (q * g.sw - g.pad_w);
          if (s < 0 || s >= g.kw) continue;
          const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
          const uint4 dv = *reinterpret_cast<const uint4*>(dy + orow * lddy + cv * 8);
          const uint2 am = *reinterpret_cast<const uint2*>(argmax + orow * g.c + cv * 8);
          const unsigned tapv = static_cast<unsigned>(r * g.kw + s) * 0x01010101u;
          const unsigned e0 = am.x ^ tapv, e1 = am.y ^ tapv;
          // 0x80 in every byte of e that is zero, then 0xFF
          const unsigned z0 = ~(((e0 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e0 | 0x7F7F7F7Fu);
          const unsigned z1 = ~(((e1 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e1 | 0x7F7F7F7Fu);
          const unsigned b0 = (z0 >> 7) * 0xFFu, b1 = (z1 >> 7) * 0xFFu;
          unsigned w[4];
          w[0] = dv.x & __byte_perm(b0, 0, 0x1100);
          w[1] = dv.y & __byte_perm(b0, 0, 0x3322);
          w[2] = dv.z & __byte_perm(b1, 0, 0x1100);
          w[3] = dv.w & __byte_perm(b1, 0, 0x3322);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            hacc[k] = __hadd2(hacc[k], *reinterpret_cast<const __half2*>(&w[k]));
        }
      }
      if (accumulate) {
        const Half8 old = *reinterpret_cast<const Half8*>(o);
#pragma unroll
        for (int k = 0; k < 4; ++k) hacc[k] = __hadd2(hacc[k], old.h[k]);
      }
      Half8 outv;
#pragma unroll
      for (int k = 0; k < 4; ++k) outv.h[k] = hacc[k];
      *reinterpret_cast<Half8*>(o) = outv;
      continue;
    }
    for (int p = p_lo; p <= p_hi; ++p) {
      const int r = hh - (p * g.sh - g.pad_h);
      if (r < 0 || r >= g.kh) continue;
      for (int q = q_lo; q <= q_hi; ++q) {
        const int s = ww - (q * g.sw - g.pad_w);
        if (s < 0 || s >= g.kw) continue;
        const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
        float d[8];
        load8(dy + orow * lddy + cv * 8, d);
        if (is_max) {
          const uint2 packed = *reinterpret_cast<const uint2*>(argmax + orow * g.c + cv * 8);
          const int tap = r * g.kw + s;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const unsigned a = ((k < 4 ? packed.x : packed.y) >> (8 * (k & 3))) & 0xffu;
            if (static_cast<int>(a) == tap) acc[k] += d[k];
          }
        } else {
          // number of in-bounds taps of window (p, q)
          const int h0 = p * g.sh - g.pad_h, w0 = q * g.sw - g.pad_w;
          const int hc = min(h0 + g.kh, g.h) - max(h0, 0);
          const int wc = min(w0 + g.kw, g.w) - max(w0, 0);
          const float inv = 1.f / static_cast<float>(hc * wc);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] += d[k] * inv;
        }
      }
    }
    if (accumulate) {
      float old[8];
      load8(o, old);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += old[k];
    }
    store8(o, acc);
  }
}

// 3x3 / 2 max-pool backward, one thread per 2x2 block of input pixels and channel vector. The
// gather kernel above loads up to four windows (gradient + arg-max, 24 B each) for every input
// pixel - 2.25 valid ones on average, 2.4 GB of L2 -> SM traffic for ResNet's stem pool at batch
// 256, which is what bounded it. The four pixels of a block (padded coordinates 2a..2a+1,
// 2b..2b+1) are covered by the same four windows (a-1..a, b-1..b): they are loaded once, and
// the nine (window, pixel) pairs are resolved in registers with the same SWAR byte compare.
__global__ void __launch_bounds__(256)
    pool_bwd_max3s2_block_kernel(const __half* __restrict__ dy, long long lddy,
                                 const unsigned char* __restrict__ argmax,
                                 __half* __restrict__ dx, long long lddx, int accumulate,
                                 PoolGeom g) {
  const int vpr = g.c >> 3;
  const int nb_w = (g.w + g.pad_w + 1) >> 1, nb_h = (g.h + g.pad_h + 1) >> 1;
  for (int n = blockIdx.z; n < g.n; n += gridDim.z)
  for (int a = blockIdx.y; a < nb_h; a += gridDim.y)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < nb_w * vpr;
       t += gridDim.x * blockDim.x) {
    const int b = t / vpr;
    const int cv = t - b * vpr;
    uint4 dv[4];
    uint2 am[4];
#pragma unroll
    for (int jp = 0; jp < 2; ++jp) {
#pragma unroll
      for (int jq = 0; jq < 2; ++jq) {
        const int p = a - jp, q = b - jq;
        const int idx = jp * 2 + jq;
        dv[idx] = make_uint4(0u, 0u, 0u, 0u);
        am[idx] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);  // matches no tap
        if (p >= 0 && p < g.p && q >= 0 && q < g.q) {
          const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
          dv[idx] = *reinterpret_cast<const uint4*>(dy + orow * lddy + cv * 8);
          am[idx] = *reinterpret_cast<const uint2*>(argmax + orow * g.c + cv * 8);
        }
      }
    }
#pragma unroll
    for (int dh = 0; dh < 2; ++dh) {
      const int hh = 2 * a + dh - g.pad_h;
      if (hh < 0 || hh >= g.h) continue;
#pragma unroll
      for (int dw = 0; dw < 2; ++dw) {
        const int ww = 2 * b + dw - g.pad_w;
        if (ww < 0 || ww >= g.w) continue;
        __half2 hacc[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) hacc[k] = __floats2half2_rn(0.f, 0.f);
#pragma unroll
        for (int jp = 0; jp < 2 - dh; ++jp) {
#pragma unroll
          for (int jq = 0; jq < 2 - dw; ++jq) {
            const int idx = jp * 2 + jq;
            const unsigned tapv =
                static_cast<unsigned>((dh + 2 * jp) * 3 + dw + 2 * jq) * 0x01010101u;
            const unsigned e0 = am[idx].x ^ tapv, e1 = am[idx].y ^ tapv;
            const unsigned z0 = ~(((e0 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e0 | 0x7F7F7F7Fu);
            const unsigned z1 = ~(((e1 & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | e1 | 0x7F7F7F7Fu);
            const unsigned b0 = (z0 >> 7) * 0xFFu, b1 = (z1 >> 7) * 0xFFu;
            unsigned w[4];
            w[0] = dv[idx].x & __byte_perm(b0, 0, 0x1100);
            w[1] = dv[idx].y & __byte_perm(b0, 0, 0x3322);
            w[2] = dv[idx].z & __byte_perm(b1, 0, 0x1100);
            w[3] = dv[idx].w & __byte_perm(b1, 0, 0x3322);
#pragma unroll
            for (int k = 0; k < 4; ++k)
              hacc[k] = __hadd2(hacc[k], *reinterpret_cast<const __half2*>(&w[k]));
          }
        }
        __half* o = dx + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * lddx + cv * 8;
        if (accumulate) {
          const Half8 old = *reinterpret_cast<const Half8*>(o);
#pragma unroll
          for (int k = 0; k < 4; ++k) hacc[k] = __hadd2(hacc[k], old.h[k]);
        }
        Half8 outv;
#pragma unroll
        for (int k = 0; k < 4; ++k) outv.h[k] = hacc[k];
        *reinterpret_cast<Half8*>(o) = outv;
      }
    }
  }
}

// y[n][c] = mean over hw of x[n][hw][c]
__global__ void __launch_bounds__(256)
    spatial_mean_fwd_kernel(const __half* __restrict__ x, long long ldx, __half* __restrict__ y,
                            long long ldy, int n, int hw, int C) {
  const int vpr = C >> 3;
  const long long total = static_cast<long long>(n) * vpr;
  const long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int cv = static_cast<int>(i % vpr);
  const long long img = i / vpr;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  for (int j = 0; j < hw; ++j) {
    float v[8];
    load8(x + (img * hw + j) * ldx + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += v[k];
  }
  const float inv = 1.f / hw;
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] *= inv;
  store8(y + img * ldy + cv * 8, acc);
}

__global__ void __launch_bounds__(256)
    spatial_mean_bwd_kernel(const __half* __restrict__ dy, long long lddy, __half* __restrict__ dx,
                            long long lddx, int accumulate, int n, int hw, int C) {
  const int vpr = C >> 3;
  const long long total = static_cast<long long>(n) * hw * vpr;
  const float inv = 1.f / hw;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int cv = static_cast<int>(i % vpr);
    const long long row = i / vpr;
    const long long img = row / hw;
    float d[8];
    load8(dy + img * lddy + cv * 8, d);
#pragma unroll
    for (int k = 0; k < 8; ++k) d[k] *= inv;
    __half* o = dx + row * lddx + cv * 8;
    if (accumulate) {
      float old[8];
      load8(o, old);
#pragma unroll
      for (int k = 0; k < 8; ++k) d[k] += old[k];
    }
    store8(o, d);
  }
}

// Strided 2-D copy of 16-byte vectors (concat / split / pad). accumulate: dst += src.
__global__ void __launch_bounds__(256)
    copy2d_kernel(const __half* __restrict__ src, long long lds, __half* __restrict__ dst,
                  long long ldd, long long rows, int C, int accumulate) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    if (accumulate) {
      float a[8], b[8];
      load8(src + r * lds + cv * 8, a);
      load8(dst + r * ldd + cv * 8, b);
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] += b[k];
      store8(dst + r * ldd + cv * 8, a);
    } else {
      *reinterpret_cast<uint4*>(dst + r * ldd + cv * 8) =
          *reinterpret_cast<const uint4*>(src + r * lds + cv * 8);
    }
  }
}

// Explicit im2col for convolutions whose input channel count is not a multiple of 8
// (image stems, C = 1 or 3): col[row][(r*S+s)*C + c], zero padded to kpad columns.
// TC/TKH/TKW > 0 make the filter geometry compile-time (cheap index arithmetic); 0 = runtime.
template <int TC, int TKH, int TKW>
__global__ void __launch_bounds__(256)
    im2col_small_c_kernel(const __half* __restrict__ x, __half* __restrict__ col, int n, int h,
                          int w, int c_rt, int p, int q, int kh_rt, int kw_rt, int sh, int sw,
                          int pad_h, int pad_w, int kpad_rt) {
  const int c = TC > 0 ? TC : c_rt;
  const int kh = TKH > 0 ? TKH : kh_rt;
  const int kw = TKW > 0 ? TKW : kw_rt;
  const int ktot = kh * kw * c;
  const int kpad = TC > 0 ? ((TC * TKH * TKW + 7) / 8 * 8) : kpad_rt;
  const int vpr = kpad >> 3;
  // grid = (column x K-vector chunks, output rows, images): no division of a flat index
  for (int nn = blockIdx.z; nn < n; nn += gridDim.z)
  for (int pp = blockIdx.y; pp < p; pp += gridDim.y)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < q * vpr; t += gridDim.x * blockDim.x) {
    const int qq = t / vpr;
    const int kv = t - qq * vpr;
    const long long row = (static_cast<long long>(nn) * p + pp) * q + qq;
    const int h0 = pp * sh - pad_h, w0 = qq * sw - pad_w;
    const __half* xin = x + static_cast<long long>(nn) * h * w * c;
    if (kw * c >= 8 && h0 >= 0 && h0 + kh <= h && w0 >= 0 && w0 + kw <= w) {
      // Interior window (~95 % of a 7x7/2 stem): a patch row (kw * c halves) is contiguous in
      // the NHWC input, so the 8 outputs of this thread are 8 consecutive halves of at most two
      // patch rows - raw 16-bit copies, no bounds checks, no conversions.
      const unsigned short* src = reinterpret_cast<const unsigned short*>(xin);
      const int rowlen = kw * c;
      const int k0 = kv * 8;
      const int r0 = k0 / rowlen;
      const int o0 = k0 - r0 * rowlen;
      const int base0 = ((h0 + r0) * w + w0) * c + o0;
      const int base1 = base0 + w * c - rowlen;       // same offset in the next patch row
      unsigned short e[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int addr = (o0 + j < rowlen) ? base0 + j : base1 + j;
        e[j] = (k0 + j < ktot) ? __ldg(src + addr) : static_cast<unsigned short>(0);
      }
      uint4 packed;
      packed.x = e[0] | (static_cast<unsigned>(e[1]) << 16);
      packed.y = e[2] | (static_cast<unsigned>(e[3]) << 16);
      packed.z = e[4] | (static_cast<unsigned>(e[5]) << 16);
      packed.w = e[6] | (static_cast<unsigned>(e[7]) << 16);
      *reinterpret_cast<uint4*>(col + row * kpad + kv * 8) = packed;
      continue;
    }
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = kv * 8 + j;
      float val = 0.f;
      if (k < ktot) {
        const int tap = k / c;
        const int cc = k - tap * c;
        const int r = tap / kw;
        const int s = tap - r * kw;
        const int hh = h0 + r, ww = w0 + s;
        if (hh >= 0 && hh < h && ww >= 0 && ww < w)
          val = __half2float(xin[(hh * w + ww) * c + cc]);  // one image < 2^31 elements
      }
      v[j] = val;
    }
    store8(col + row * kpad + kv * 8, v);
  }
}

// ---------------------------------------------------------------------------
// Philox-4x32-10 (counter-based RNG) for dropout masks and synthetic inputs.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u;
    key.y += 0xBB67AE85u;
  }
  return ctr;
}
__device__ __forceinline__ float u01(uint32_t r) { return (r >> 8) * (1.0f / 16777216.0f); }

// y = x * mask / keep, mask regenerated from (seed, offset) in backward (same kernel).
__global__ void __launch_bounds__(256)
    dropout_kernel(const __half* __restrict__ x, __half* __restrict__ y, long long nvec,
                   float keep_prob, unsigned long long seed, unsigned long long offset,
                   const unsigned long long* __restrict__ step, int step_shift) {
  // device-resident step counter (CUDA-graph replay): the Philox offset moves with the step
  // although the by-value arguments of the captured launch stay the same
  if (step != nullptr) offset += (*step) << step_shift;
  const float inv_keep = 1.f / keep_prob;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    const unsigned long long cnt = offset + static_cast<unsigned long long>(i);
    const uint4 r0 = philox4x32(
        make_uint4(static_cast<uint32_t>(cnt), static_cast<uint32_t>(cnt >> 32), 0u, 0u), key);
    const uint4 r1 = philox4x32(
        make_uint4(static_cast<uint32_t>(cnt), static_cast<uint32_t>(cnt >> 32), 1u, 0u), key);
    const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    float v[8];
    load8(x + i * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = u01(rr[k]) < keep_prob ? v[k] * inv_keep : 0.f;
    store8(y + i * 8, v);
  }
}

// Synthetic images: truncated normal(mean, std) clipped at 2 sigma (resampled), fp16.
// T/models/model.py:220-237 (tf.truncated_normal mean 127 stddev 60).
__global__ void __launch_bounds__(256)
    synth_images_kernel(__half* __restrict__ out, long long n, float mean, float stddev,
                        unsigned long long seed) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
    float z = 0.f;
    for (uint32_t attempt = 0; attempt < 8; ++attempt) {
      const uint4 r = philox4x32(
          make_uint4(static_cast<uint32_t>(i), static_cast<uint32_t>(i >> 32), attempt, 7u), key);
      const float u1 = fmaxf(u01(r.x), 1e-7f), u2 = u01(r.y);
      const float rad = sqrtf(-2.f * __logf(u1));
      float s, c;
      __sincosf(6.28318530718f * u2, &s, &c);
      z = rad * c;
      if (fabsf(z) <= 2.f) break;
      z = rad * s;
      if (fabsf(z) <= 2.f) break;
      z = fminf(fmaxf(z, -2.f), 2.f);
    }
    out[i] = __float2half_rn(mean + stddev * z);
  }
}

// Labels uniform in [lo, hi)
__global__ void synth_labels_kernel(int* __restrict__ out, int n, int lo, int hi,
                                    unsigned long long seed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint2 key = make_uint2(static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
  const uint4 r = philox4x32(make_uint4(static_cast<uint32_t>(i), 0u, 0u, 11u), key);
  out[i] = lo + static_cast<int>(r.x % static_cast<uint32_t>(hi - lo));
}

// ---------------------------------------------------------------------------
// Local response normalisation across channels (T/convnet_builder_legacy.py:462-468).
// y = x / (bias + alpha * sum_{|j-c|<=radius} x_j^2)^beta
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    lrn_fwd_kernel(const __half* __restrict__ x, __half* __restrict__ y, long long rows, int C,
                   int radius, float bias, float alpha, float beta) {
  const long long total = rows * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / C;
    const int c = static_cast<int>(i - r * C);
    float ss = 0.f;
    const int lo = max(c - radius, 0), hi = min(c + radius, C - 1);
    for (int j = lo; j <= hi; ++j) {
      const float v = __half2float(x[r * C + j]);
      ss += v * v;
    }
    const float d = bias + alpha * ss;
    y[i] = __float2half_rn(__half2float(x[i]) * __powf(d, -beta));
  }
}

__global__ void __launch_bounds__(256)
    lrn_bwd_kernel(const __half* __restrict__ x, const __half* __restrict__ dy,
                   __half* __restrict__ dx, int accumulate, long long rows, int C, int radius,
                   float bias, float alpha, float beta) {
  const long long total = rows * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / C;
    const int c = static_cast<int>(i - r * C);
    const float xc = __half2float(x[i]);
    // dx_c = dy_c * d_c^-beta - 2 alpha beta x_c * sum_{j: |j-c|<=radius} dy_j x_j d_j^(-beta-1)
    float acc = 0.f, own = 0.f;
    const int lo = max(c - radius, 0), hi = min(c + radius, C - 1);
    for (int j = lo; j <= hi; ++j) {
      float ss = 0.f;
      const int l2 = max(j - radius, 0), h2 = min(j + radius, C - 1);
      for (int k = l2; k <= h2; ++k) {
        const float v = __half2float(x[r * C + k]);
        ss += v * v;
      }
      const float d = bias + alpha * ss;
      const float dyj = __half2float(dy[r * C + j]);
      if (j == c) own = dyj * __powf(d, -beta);
      acc += dyj * __half2float(x[r * C + j]) * __powf(d, -beta - 1.f);
    }
    float g = own - 2.f * alpha * beta * xc * acc;
    if (accumulate) g += __half2float(dx[i]);
    dx[i] = __float2half_rn(g);
  }
}

// ---------------------------------------------------------------------------
// Softmax cross-entropy with sparse labels, fused with the backward gradient, the
// loss-scale multiply and top-1 / top-5 hits. One warp per row.
//   loss_row[r] = logsumexp(z) - z[label]
//   dlogits[r]  = (softmax(z) - onehot) * grad_scale       (fp16, row stride lddl)
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    softmax_xent_kernel(const float* __restrict__ logits, long long ldl,
                        const int* __restrict__ labels, int rows, int classes,
                        float grad_scale, const float* __restrict__ dyn_scale,
                        __half* __restrict__ dlogits, long long lddl,
                        float* __restrict__ loss_rows, int* __restrict__ top1,
                        int* __restrict__ top5) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  const float* z = logits + static_cast<long long>(warp) * ldl;
  const int label = labels[warp];
  float m = -INFINITY;
  for (int j = lane; j < classes; j += 32) m = fmaxf(m, z[j]);
  m = warp_max(m);
  float s = 0.f;
  for (int j = lane; j < classes; j += 32) s += __expf(z[j] - m);
  s = warp_sum(s);
  const float zl = (label >= 0 && label < classes) ? z[label] : 0.f;
  const float lse = m + __logf(s);
  // rank of the label logit = number of classes strictly greater (tf.nn.in_top_k counts ties
  // as hits)
  float greater = 0.f;
  for (int j = lane; j < classes; j += 32) greater += (z[j] > zl) ? 1.f : 0.f;
  greater = warp_sum(greater);
  if (lane == 0) {
    loss_rows[warp] = lse - zl;
    if (top1) top1[warp] = greater < 1.f ? 1 : 0;
    if (top5) top5[warp] = greater < 5.f ? 1 : 0;
  }
  if (dlogits != nullptr) {
    if (dyn_scale != nullptr) grad_scale *= *dyn_scale;
    const float inv = 1.f / s;
    __half* d = dlogits + static_cast<long long>(warp) * lddl;
    for (int j = lane; j < lddl; j += 32) {
      float g = 0.f;
      if (j < classes) g = (__expf(z[j] - m) * inv - (j == label ? 1.f : 0.f)) * grad_scale;
      d[j] = __float2half_rn(g);
    }
  }
}

// out[0] = sum(loss_rows)/rows ; out[1] = sum(top1) ; out[2] = sum(top5)  (single block)
__global__ void loss_finalize_kernel(const float* __restrict__ loss_rows,
                                     const int* __restrict__ top1, const int* __restrict__ top5,
                                     int rows, float* __restrict__ out) {
  __shared__ float sh[3][32];
  float a = 0.f, b = 0.f, c = 0.f;
  for (int i = threadIdx.x; i < rows; i += blockDim.x) {
    a += loss_rows[i];
    if (top1) b += top1[i];
    if (top5) c += top5[i];
  }
  a = warp_sum(a);
  b = warp_sum(b);
  c = warp_sum(c);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) {
    sh[0][w] = a;
    sh[1][w] = b;
    sh[2][w] = c;
  }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    a = l < nw ? sh[0][l] : 0.f;
    b = l < nw ? sh[1][l] : 0.f;
    c = l < nw ? sh[2][l] : 0.f;
    a = warp_sum(a);
    b = warp_sum(b);
    c = warp_sum(c);
    if (l == 0) {
      out[0] = a / rows;
      out[1] = b;
      out[2] = c;
    }
  }
}

// fp32 -> fp16 cast (weight shadow refresh when the optimizer is not the fused one)
__global__ void __launch_bounds__(256)
    cast_f32_f16_kernel(const float* __restrict__ in, __half* __restrict__ out, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = __float2half_rn(in[i]);
}

// dst[r][0..cols) = src[r][0..cols) for arbitrary cols / strides (flatten into padded rows).
__global__ void __launch_bounds__(256)
    copy_rows_kernel(const __half* __restrict__ src, long long lds, __half* __restrict__ dst,
                     long long ldd, long long rows, long long cols, int accumulate) {
  const long long total = rows * cols;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / cols;
    const long long c = i - r * cols;
    const __half v = src[r * lds + c];
    __half* o = dst + r * ldd + c;
    *o = accumulate ? __float2half_rn(__half2float(*o) + __half2float(v)) : v;
  }
}

static inline int grid_for(long long work_items, int threads = 256, int max_blocks = 148 * 16) {
  long long b = (work_items + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return static_cast<int>(b);
}

// 2-D launch shape for the per-row apply kernels: x = channel vectors, y = rows.
// Launch-shape knobs for sweeps on the device without a rebuild (read once; the defaults are the
// measured configuration): KFB_APPLY_CTAS_PER_SM (default 16) caps the grid of the BN apply
// kernels, KFB_REDUCE_BLOCKS_PER_SM (default 4) the grid of the channel reductions.
static int env_int(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  if (e == nullptr || *e == 0) return dflt;
  const int v = atoi(e);
  return v < lo ? lo : (v > hi ? hi : v);
}
static int apply_ctas_per_sm() {
  static const int v = env_int("KFB_APPLY_CTAS_PER_SM", 4, 1, 64);  // one resident wave (measured: 4 > 8 > 16 > 6)
  return v;
}
static int apply_unroll() {
  static const int v = env_int("KFB_APPLY_UNROLL", 2, 1, 8);  // measured: 2 beats 1, 4 and 8
  return v;
}
static int reduce_blocks_per_sm() {
  static const int v = env_int("KFB_REDUCE_BLOCKS_PER_SM", 4, 1, 16);
  return v;
}

static inline void apply_launch_shape(long long rows, int C, dim3* grid, dim3* block) {
  const int vpr = C >> 3;
  int vx = 1;
  while (vx < vpr && vx < 256) vx <<= 1;  // power of two >= vpr (capped at 256)
  if (vx > vpr && (vx >> 1) >= vpr) vx >>= 1;
  const int vy = 256 / vx > 0 ? 256 / vx : 1;
  const int gx = (vpr + vx - 1) / vx;
  long long gy = (rows + vy - 1) / vy;
  const long long cap = (148LL * apply_ctas_per_sm() + gx - 1) / gx;
  if (gy > cap) gy = cap;
  if (gy < 1) gy = 1;
  *grid = dim3((unsigned)gx, (unsigned)gy, 1);
  *block = dim3((unsigned)vx, (unsigned)vy, 1);
}

static inline int reduce_blocks(long long rows, int C) {
  // enough blocks to fill the machine, but each block should own >= 64 rows
  long long b = rows / 64;
  const long long cap = 148LL * reduce_blocks_per_sm();
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace kfb

using namespace kfb;

#define KFB_REQ_C8(C, who) KFB_REQUIRE((C) > 0 && (C) % 8 == 0, who ": C (%d) must be a multiple of 8", (C))

extern "C" int kfb_reduce_blocks(long long rows, int C) { return reduce_blocks(rows, C); }

// scratch must hold kfb_reduce_blocks(rows,C) * 2 * C floats.
extern "C" int kfb_bn_forward_train(const void* x, long long ldx, long long rows, int C,
                                    const float* gamma, const float* beta, float* moving_mean,
                                    float* moving_var, float eps, float decay, int update_moving,
                                    float* mean_rstd, float* scale_shift, float* scratch,
                                    int stats_nblocks, const void* residual, long long ldr,
                                    int relu, void* y, long long ldy, void* stream) {
  KFB_REQ_C8(C, "bn_forward_train");
  KFB_REQUIRE(stats_nblocks == 0,
              "bn_forward_train: conv-epilogue statistics go through kfb_bn_forward_train_fused");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = reduce_blocks(rows, C);
  channel_reduce_kernel<0><<<nb, kRedThreads, kRedThreads * 16 * sizeof(float), st>>>(
      (const __half*)x, ldx, nullptr, 0, nullptr, 0, nullptr, nullptr, nullptr, 0, rows, C,
      scratch);
  const float unbias = rows > 1 ? (float)((double)rows / (double)(rows - 1)) : 1.f;
  bn_finalize_kernel<<<(C + 31) / 32, dim3(32, kFoldLanes), 0, st>>>(
      scratch, nb, C, (float)(1.0 / (double)rows), unbias, eps, decay, gamma, beta, moving_mean,
      moving_var, mean_rstd, scale_shift, update_moving);
  dim3 ag, ab;
  apply_launch_shape(rows, C, &ag, &ab);
  switch (apply_unroll()) {
    case 8: bn_apply_kernel<8><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    case 4: bn_apply_kernel<4><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    case 2: bn_apply_kernel<2><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    default: bn_apply_kernel<1><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_bn_forward_inference(const void* x, long long ldx, long long rows, int C,
                                        const float* gamma, const float* beta,
                                        const float* moving_mean, const float* moving_var,
                                        float eps, float* scale_shift, const void* residual,
                                        long long ldr, int relu, void* y, long long ldy,
                                        void* stream) {
  KFB_REQ_C8(C, "bn_forward_inference");
  cudaStream_t st = (cudaStream_t)stream;
  bn_inference_affine_kernel<<<(C + 127) / 128, 128, 0, st>>>(C, eps, gamma, beta, moving_mean,
                                                             moving_var, scale_shift);
  dim3 ag, ab;
  apply_launch_shape(rows, C, &ag, &ab);
  switch (apply_unroll()) {
    case 8: bn_apply_kernel<8><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    case 4: bn_apply_kernel<4><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    case 2: bn_apply_kernel<2><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
    default: bn_apply_kernel<1><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, nullptr); break;
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Two-kernel training forward on conv-epilogue statistics: finalize (fold of the partial rows)
// and apply. For small layers this beats the single-launch form, whose two CTAs per SM and grid
// barrier cost more than the second launch (measured, profiles/README.md).
extern "C" int kfb_bn_forward_train_stats(const void* x, long long ldx, long long rows, int C,
                                          const float* gamma, const float* beta,
                                          float* moving_mean, float* moving_var, float eps,
                                          float decay, int update_moving, float* mean_rstd,
                                          float* scale_shift, const float* stats, int stats_rows,
                                          int stats_bn, const void* residual, long long ldr,
                                          int relu, void* y, long long ldy,
                                          unsigned char* bitmask, void* stream) {
  KFB_REQ_C8(C, "bn_forward_train_stats");
  KFB_REQUIRE(stats != nullptr && stats_rows > 0 && stats_bn > 0,
              "bn_forward_train_stats: needs the conv-epilogue statistics");
  cudaStream_t st = (cudaStream_t)stream;
  const float unbias = rows > 1 ? (float)((double)rows / (double)(rows - 1)) : 1.f;
  bn_finalize_stats_kernel<<<(C + 31) / 32, dim3(32, kFoldLanes), 0, st>>>(
      stats, stats_rows, stats_bn, C, (float)(1.0 / (double)rows), unbias, eps, decay, gamma, beta,
      moving_mean, moving_var, mean_rstd, scale_shift, update_moving);
  dim3 ag, ab;
  apply_launch_shape(rows, C, &ag, &ab);
  switch (apply_unroll()) {
    case 8: bn_apply_kernel<8><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, bitmask); break;
    case 4: bn_apply_kernel<4><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, bitmask); break;
    case 2: bn_apply_kernel<2><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, bitmask); break;
    default: bn_apply_kernel<1><<<ag, ab, 0, st>>>((const __half*)x, ldx, scale_shift,
                                     (const __half*)residual, ldr, (__half*)y, ldy, rows, C,
                                     relu, bitmask); break;
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Backward of y = act(bn(x) (+ residual)). ReLU mask: `mask` = y when given; else, if
// `relu_affine` (the forward scale/shift, [2][C]) is given and there was no residual, the mask
// is recomputed from x (saves re-reading y twice); both null = no activation.
// coef: 3*C floats of scratch; scratch: kfb_reduce_blocks * 2 * C floats.
extern "C" int kfb_bn_backward(const void* dy, long long lddy, const void* mask, long long ldm,
                               const float* relu_affine, const void* x, long long ldx,
                               long long rows, int C, const float* gamma,
                               const float* mean_rstd, float* dgamma,
                               float* dbeta, int accumulate_param_grads, float grad_scale,
                               float* coef, float* scratch, void* dz, long long lddz, void* dres,
                               long long lddr, int dres_accumulate, void* stream) {
  KFB_REQ_C8(C, "bn_backward");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = reduce_blocks(rows, C);
  channel_reduce_kernel<1><<<nb, kRedThreads, kRedThreads * 16 * sizeof(float), st>>>(
      (const __half*)x, ldx, (const __half*)dy, lddy, (const __half*)mask, ldm, mean_rstd,
      mask ? nullptr : relu_affine, nullptr, 0, rows, C, scratch);
  bn_bwd_finalize_kernel<<<(C + 31) / 32, dim3(32, kFoldLanes), 0, st>>>(
      scratch, nb, C, (float)(1.0 / (double)rows), gamma, mean_rstd, dgamma, dbeta,
      accumulate_param_grads, grad_scale, coef);
  dim3 ag, ab;
  apply_launch_shape(rows, C, &ag, &ab);
  bn_bwd_apply_kernel<<<ag, ab, 0, st>>>((const __half*)dy, lddy, (const __half*)mask, ldm,
                                         (const __half*)x, ldx, coef,
                                         mask ? nullptr : relu_affine, (__half*)dz, lddz,
                                         (__half*)dres, lddr, dres_accumulate, rows, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dz = dy * (mask > 0) (dz may be null or alias dy); dbias (+)= scale * colsum(dz).
extern "C" int kfb_bias_act_backward(const void* dy, long long lddy, const void* mask,
                                     long long ldm, long long rows, int C, void* dz,
                                     long long lddz, float* dbias, int accumulate, float scale,
                                     float* scratch, void* stream) {
  KFB_REQ_C8(C, "bias_act_backward");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = reduce_blocks(rows, C);
  channel_reduce_kernel<2><<<nb, kRedThreads, kRedThreads * 16 * sizeof(float), st>>>(
      nullptr, 0, (const __half*)dy, lddy, (const __half*)mask, ldm, nullptr, nullptr,
      (__half*)dz, lddz, rows, C, scratch);
  if (dbias != nullptr)
    colsum_finalize_kernel<<<(C + 31) / 32, dim3(32, kFoldLanes), 0, st>>>(scratch, nb, C, dbias,
                                                                           accumulate, scale);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_add_act(const void* a, long long lda, const void* b, long long ldb, void* y,
                           long long ldy, long long rows, int C, int relu, void* stream) {
  KFB_REQ_C8(C, "add_act");
  add_act_kernel<<<grid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)a, lda, (const __half*)b, ldb, (__half*)y, ldy, rows, C, relu);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_relu_backward(const void* dy, long long lddy, const void* mask, long long ldm,
                                 void* dx, long long lddx, int accumulate, long long rows, int C,
                                 void* stream) {
  KFB_REQ_C8(C, "relu_backward");
  relu_bwd_kernel<<<grid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, lddy, (const __half*)mask, ldm, (__half*)dx, lddx, accumulate, rows, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int fill_pool(PoolGeom* g, int n, int h, int w, int c, int p, int q, int kh, int kw, int sh,
                     int sw, int pad_h, int pad_w) {
  KFB_REQ_C8(c, "pool");
  KFB_REQUIRE(kh * kw <= 255, "pool: window %dx%d too large", kh, kw);
  *g = PoolGeom{n, h, w, c, p, q, kh, kw, sh, sw, pad_h, pad_w};
  return 0;
}

// Grid of the pooling kernels: x = column / channel-vector chunks, y = rows, z = images, with the
// loops of the kernels striding over what the grid does not cover. KFB_POOL_CTAS_PER_SM caps the
// total number of CTAs (0 = one CTA per (chunk, row, image)).
static dim3 pool_grid(int cols_x_vecs, int rows, int images) {
  static const int per_sm = env_int("KFB_POOL_CTAS_PER_SM", 0, 0, 4096);
  unsigned bx = (unsigned)((cols_x_vecs + 255) / 256);
  unsigned by = (unsigned)(rows < 65535 ? rows : 65535);
  unsigned bz = (unsigned)(images < 65535 ? images : 65535);
  if (per_sm > 0) {
    const unsigned long long target = 148ull * per_sm;
    while ((unsigned long long)bx * by * bz > target && bz > 1) bz = (bz + 1) / 2;
    while ((unsigned long long)bx * by * bz > target && by > 1) by = (by + 1) / 2;
  }
  return dim3(bx, by, bz);
}

extern "C" int kfb_pool_forward(const void* x, long long ldx, void* y, long long ldy,
                                unsigned char* argmax, int n, int h, int w, int c, int p, int q,
                                int kh, int kw, int sh, int sw, int pad_h, int pad_w, int is_max,
                                void* stream) {
  PoolGeom g;
  if (fill_pool(&g, n, h, w, c, p, q, kh, kw, sh, sw, pad_h, pad_w)) return 2;
  const dim3 blocks = pool_grid(q * (c / 8), p, n);
  cudaStream_t st = (cudaStream_t)stream;
  static const int packed_form = env_int("KFB_POOL_FWD_PACKED", 1, 0, 1);
  if (is_max && packed_form && kh == kw && sh == sw &&
      ((kh == 3 && sh == 2) || (kh == 2 && sh == 2) || (kh == 3 && sh == 1))) {
    if (kh == 3 && sh == 2)
      pool_fwd_max_packed_kernel<3, 2><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y,
                                                               ldy, argmax, g);
    else if (kh == 2)
      pool_fwd_max_packed_kernel<2, 2><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y,
                                                               ldy, argmax, g);
    else
      pool_fwd_max_packed_kernel<3, 1><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y,
                                                               ldy, argmax, g);
  } else if (kh == 3 && kw == 3 && sh == 2 && sw == 2)
    pool_fwd_kernel<3, 2><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y, ldy, argmax,
                                                  g, is_max);
  else if (kh == 2 && kw == 2 && sh == 2 && sw == 2)
    pool_fwd_kernel<2, 2><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y, ldy, argmax,
                                                  g, is_max);
  else if (kh == 3 && kw == 3 && sh == 1 && sw == 1)
    pool_fwd_kernel<3, 1><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y, ldy, argmax,
                                                  g, is_max);
  else
    pool_fwd_kernel<0, 0><<<blocks, 256, 0, st>>>((const __half*)x, ldx, (__half*)y, ldy, argmax,
                                                  g, is_max);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_pool_backward(const void* dy, long long lddy, const unsigned char* argmax,
                                 void* dx, long long lddx, int accumulate, int n, int h, int w,
                                 int c, int p, int q, int kh, int kw, int sh, int sw, int pad_h,
                                 int pad_w, int is_max, void* stream) {
  PoolGeom g;
  if (fill_pool(&g, n, h, w, c, p, q, kh, kw, sh, sw, pad_h, pad_w)) return 2;
  const dim3 blocks = pool_grid(w * (c / 8), h, n);
  cudaStream_t st = (cudaStream_t)stream;
  static const int block_form = env_int("KFB_POOL_BWD_BLOCK", 1, 0, 1);
  if (kh == 3 && kw == 3 && sh == 2 && sw == 2 && is_max && pad_h >= 0 && pad_w >= 0 &&
      block_form) {
    const dim3 bb = pool_grid(((w + pad_w + 1) / 2) * (c / 8), (h + pad_h + 1) / 2, n);
    pool_bwd_max3s2_block_kernel<<<bb, 256, 0, st>>>((const __half*)dy, lddy, argmax,
                                                      (__half*)dx, lddx, accumulate, g);
  } else if (kh == 3 && kw == 3 && sh == 2 && sw == 2)
    pool_bwd_kernel<3, 2><<<blocks, 256, 0, st>>>((const __half*)dy, lddy, argmax, (__half*)dx,
                                                  lddx, accumulate, g, is_max);
  else if (kh == 2 && kw == 2 && sh == 2 && sw == 2)
    pool_bwd_kernel<2, 2><<<blocks, 256, 0, st>>>((const __half*)dy, lddy, argmax, (__half*)dx,
                                                  lddx, accumulate, g, is_max);
  else if (kh == 3 && kw == 3 && sh == 1 && sw == 1)
    pool_bwd_kernel<3, 1><<<blocks, 256, 0, st>>>((const __half*)dy, lddy, argmax, (__half*)dx,
                                                  lddx, accumulate, g, is_max);
  else
    pool_bwd_kernel<0, 0><<<blocks, 256, 0, st>>>((const __half*)dy, lddy, argmax, (__half*)dx,
                                                  lddx, accumulate, g, is_max);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_spatial_mean_forward(const void* x, long long ldx, void* y, long long ldy, int n,
                                        int hw, int C, void* stream) {
  KFB_REQ_C8(C, "spatial_mean_forward");
  const long long total = (long long)n * (C / 8);
  spatial_mean_fwd_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (__half*)y, ldy, n, hw, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_spatial_mean_backward(const void* dy, long long lddy, void* dx, long long lddx,
                                         int accumulate, int n, int hw, int C, void* stream) {
  KFB_REQ_C8(C, "spatial_mean_backward");
  spatial_mean_bwd_kernel<<<grid_for((long long)n * hw * (C / 8)), 256, 0,
                            (cudaStream_t)stream>>>((const __half*)dy, lddy, (__half*)dx, lddx,
                                                    accumulate, n, hw, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_copy2d(const void* src, long long lds, void* dst, long long ldd, long long rows,
                          int C, int accumulate, void* stream) {
  KFB_REQ_C8(C, "copy2d");
  copy2d_kernel<<<grid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)src, lds, (__half*)dst, ldd, rows, C, accumulate);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_copy_rows(const void* src, long long lds, void* dst, long long ldd,
                             long long rows, long long cols, int accumulate, void* stream) {
  copy_rows_kernel<<<grid_for(rows * cols), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)src, lds, (__half*)dst, ldd, rows, cols, accumulate);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_im2col_small_c(const void* x, void* col, int n, int h, int w, int c, int p,
                                  int q, int kh, int kw, int sh, int sw, int pad_h, int pad_w,
                                  int kpad, void* stream) {
  KFB_REQUIRE(kpad % 8 == 0 && kpad >= kh * kw * c, "im2col_small_c: bad kpad %d", kpad);
  const dim3 blocks((unsigned)((q * (kpad / 8) + 255) / 256), (unsigned)(p < 65535 ? p : 65535),
                    (unsigned)(n < 65535 ? n : 65535));
  cudaStream_t st = (cudaStream_t)stream;
#define KFB_IM2COL(TC, TKH, TKW)                                                              \
  im2col_small_c_kernel<TC, TKH, TKW><<<blocks, 256, 0, st>>>(                                \
      (const __half*)x, (__half*)col, n, h, w, c, p, q, kh, kw, sh, sw, pad_h, pad_w, kpad)
  const bool tight = kpad == (kh * kw * c + 7) / 8 * 8;
  if (tight && c == 3 && kh == 7 && kw == 7) KFB_IM2COL(3, 7, 7);
  else if (tight && c == 3 && kh == 3 && kw == 3) KFB_IM2COL(3, 3, 3);
  else if (tight && c == 3 && kh == 11 && kw == 11) KFB_IM2COL(3, 11, 11);
  else if (tight && c == 3 && kh == 5 && kw == 5) KFB_IM2COL(3, 5, 5);
  else KFB_IM2COL(0, 0, 0);
#undef KFB_IM2COL
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_dropout(const void* x, void* y, long long n, float keep_prob,
                           unsigned long long seed, unsigned long long offset, void* stream) {
  KFB_REQUIRE(n % 8 == 0, "dropout: element count must be a multiple of 8");
  KFB_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "dropout: bad keep_prob %f", keep_prob);
  dropout_kernel<<<grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, (__half*)y, n / 8, keep_prob, seed, offset, nullptr, 0);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// kfb_dropout with the step part of the Philox offset taken from device memory:
// offset + (*step << step_shift). The same call serves forward and backward of a step.
extern "C" int kfb_dropout_dev(const void* x, void* y, long long n, float keep_prob,
                               unsigned long long seed, unsigned long long offset,
                               const unsigned long long* step, int step_shift, void* stream) {
  KFB_REQUIRE(n % 8 == 0, "dropout: element count must be a multiple of 8");
  KFB_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "dropout: bad keep_prob %f", keep_prob);
  dropout_kernel<<<grid_for(n / 8), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, (__half*)y, n / 8, keep_prob, seed, offset, step, step_shift);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

__global__ void step_advance_kernel(unsigned long long* step) { *step += 1ull; }

// *step += 1 on `stream` (last node of a training step; captured with it in a CUDA graph)
extern "C" int kfb_step_advance(unsigned long long* step, void* stream) {
  KFB_REQUIRE(step != nullptr, "step_advance: null counter");
  step_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

struct FloatPack {
  float v[128];
};
__global__ void set_floats_kernel(float* dst, FloatPack pack, int n) {
  if (threadIdx.x < n) dst[threadIdx.x] = pack.v[threadIdx.x];
}

// dst[0..n) <- host values, carried in kernel arguments (no pinned staging buffer that a host
// running ahead of the device could overwrite): per-step scalars of captured graphs
extern "C" int kfb_set_floats(float* dst, const float* host_values, int n, void* stream) {
  KFB_REQUIRE(dst != nullptr && host_values != nullptr && n >= 0, "set_floats: bad arguments");
  for (int i = 0; i < n; i += 128) {
    FloatPack pack;
    const int m = n - i < 128 ? n - i : 128;
    for (int k = 0; k < m; ++k) pack.v[k] = host_values[i + k];
    set_floats_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(dst + i, pack, m);
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_synth_images(void* out, long long n, float mean, float stddev,
                                unsigned long long seed, void* stream) {
  synth_images_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>((__half*)out, n, mean,
                                                                     stddev, seed);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_synth_labels(int* out, int n, int lo, int hi, unsigned long long seed,
                                void* stream) {
  KFB_REQUIRE(hi > lo, "synth_labels: empty range");
  synth_labels_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(out, n, lo, hi, seed);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_lrn_forward(const void* x, void* y, long long rows, int C, int radius,
                               float bias, float alpha, float beta, void* stream) {
  lrn_fwd_kernel<<<grid_for(rows * C), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, (__half*)y, rows, C, radius, bias, alpha, beta);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_lrn_backward(const void* x, const void* dy, void* dx, int accumulate,
                                long long rows, int C, int radius, float bias, float alpha,
                                float beta, void* stream) {
  lrn_bwd_kernel<<<grid_for(rows * C), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, (const __half*)dy, (__half*)dx, accumulate, rows, C, radius, bias, alpha,
      beta);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// out3 = {mean loss, #top1 hits, #top5 hits}; loss_rows/top1/top5 are per-row scratch outputs.
extern "C" int kfb_softmax_xent(const float* logits, long long ldl, const int* labels, int rows,
                                int classes, float grad_scale, const float* dyn_scale,
                                void* dlogits, long long lddl, float* loss_rows, int* top1,
                                int* top5, float* out3, void* stream) {
  KFB_REQUIRE(rows > 0 && classes > 0, "softmax_xent: bad shape");
  cudaStream_t st = (cudaStream_t)stream;
  const int warps_per_block = 8;
  softmax_xent_kernel<<<(rows + warps_per_block - 1) / warps_per_block, 256, 0, st>>>(
      logits, ldl, labels, rows, classes, grad_scale, dyn_scale, (__half*)dlogits, lddl,
      loss_rows, top1, top5);
  if (out3 != nullptr) loss_finalize_kernel<<<1, 256, 0, st>>>(loss_rows, top1, top5, rows, out3);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Allocation-free variant of kfb_softmax_xent.
//
// kfb_softmax_xent needs three per-row scratch arrays (row losses and the top-1 / top-5 hit
// flags that the finalize kernel folds). Some callers have no way to provide them: code that
// runs where cudaMalloc is off limits (inside a latency-critical serving or training loop
// where the implicit device synchronisation of an allocation is not acceptable, or embedded
// in a host application that owns all device memory and hands this library none). For those, the scratch lives in the fixed-size device
// array below, which every GPU that loads this library keeps until the process exits
// (kXentStaticMaxRows * 12 bytes = 48 KiB).
//
// In host code `g_xent_static_scratch` is only a host-side handle for the symbol, not a device
// address; the per-device address is looked up with cudaGetSymbolAddress the first time the
// entry point runs on a device (with that device current) and cached.
// ---------------------------------------------------------------------------
namespace kfb {
constexpr int kXentStaticMaxRows = 4096;
static __device__ __align__(256) unsigned char
    g_xent_static_scratch[kXentStaticMaxRows * 3 * sizeof(float)];

constexpr int kMaxDevices = 64;
static unsigned char* g_xent_static_addr[kMaxDevices] = {nullptr};
}  // namespace kfb

extern "C" int kfb_softmax_xent_static_max_rows(void) { return kfb::kXentStaticMaxRows; }

extern "C" int kfb_softmax_xent_static(const float* logits, long long ldl, const int* labels,
                                       int rows, int classes, float grad_scale,
                                       const float* dyn_scale, void* dlogits, long long lddl,
                                       float* out3, void* stream) {
  KFB_REQUIRE(rows > 0 && classes > 0, "softmax_xent_static: bad shape");
  KFB_REQUIRE(rows <= kXentStaticMaxRows,
              "softmax_xent_static: %d rows exceed the built-in scratch (max %d rows); use "
              "kfb_softmax_xent with caller-provided scratch",
              rows, kXentStaticMaxRows);
  KFB_REQUIRE(out3 != nullptr, "softmax_xent_static: out3 is required");
  int dev = -1;
  KFB_CHECK_CUDA(cudaGetDevice(&dev));
  KFB_REQUIRE(dev >= 0 && dev < kMaxDevices, "softmax_xent_static: device ordinal %d out of range",
              dev);
  if (g_xent_static_addr[dev] == nullptr) {
    // one-time set-up for this device (the current one): resolve the symbol's device address
    unsigned char* p = nullptr;
    KFB_CHECK_CUDA(cudaGetSymbolAddress((void**)&p, g_xent_static_scratch));
    KFB_REQUIRE(p != nullptr, "softmax_xent_static: symbol address lookup returned null");
    g_xent_static_addr[dev] = p;
  }
  unsigned char* base = g_xent_static_addr[dev];
  float* loss_rows = reinterpret_cast<float*>(base);
  int* top1 = reinterpret_cast<int*>(base + kXentStaticMaxRows * sizeof(float));
  int* top5 = reinterpret_cast<int*>(base + 2 * kXentStaticMaxRows * sizeof(float));
  cudaStream_t st = (cudaStream_t)stream;
  const int warps_per_block = 8;
  // Rows [0, rows) of all three arrays are rewritten by the first kernel before the second
  // one reads them, so nothing left over from an earlier call has to be cleared.
  softmax_xent_kernel<<<(rows + warps_per_block - 1) / warps_per_block, 256, 0, st>>>(
      logits, ldl, labels, rows, classes, grad_scale, dyn_scale, (__half*)dlogits, lddl,
      loss_rows, top1, top5);
  loss_finalize_kernel<<<1, 256, 0, st>>>(loss_rows, top1, top5, rows, out3);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Zero fill of a device range on a stream (a memset node: nothing of the step goes through a
// framework fill kernel).
extern "C" int kfb_fill_zero(void* p, long long bytes, void* stream) {
  KFB_REQUIRE(p != nullptr && bytes >= 0, "fill_zero: bad arguments");
  KFB_CHECK_CUDA(cudaMemsetAsync(p, 0, (size_t)bytes, (cudaStream_t)stream));
  return 0;
}

extern "C" int kfb_cast_f32_f16(const float* in, void* out, long long n, void* stream) {
  cast_f32_f16_kernel<<<grid_for(n), 256, 0, (cudaStream_t)stream>>>(in, (__half*)out, n);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
