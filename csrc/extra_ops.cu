// Layer kernels of the wider model zoo (sm_100a): depthwise convolution (MobileNet-v2
// expanded_conv, NASNet separable convs), activations beyond ReLU (ReLU6 / tanh / sigmoid),
// and per-sample drop-path. All are bandwidth-bound, so they are plain vectorised CUDA-core
// kernels on NHWC fp16 rows (16-byte accesses, fp32 math); the GEMM-shaped halves of a
// separable convolution (the 1x1 pointwise convs) run on the tcgen05 kernel.
//
// Reference behaviour: slim.separable_conv2d as used by T/models/mobilenet_conv_blocks.py:163-313
// and T/models/nasnet_utils.py:141-260; tf.nn.relu6; drop_path T/models/nasnet_utils.py:113-138.
#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

static inline int xgrid_for(long long work_items, int threads = 256, int max_blocks = 148 * 16) {
  long long b = (work_items + threads - 1) / threads;
  if (b < 1) b = 1;
  if (b > max_blocks) b = max_blocks;
  return static_cast<int>(b);
}

// ---------------------------------------------------------------------------
// Activations. kind: 1 relu, 2 relu6, 3 tanh, 4 sigmoid. The backward pass is written in
// terms of the stored output y, so no pre-activation copy is kept.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float act_fwd(float v, int kind) {
  switch (kind) {
    case 1: return fmaxf(v, 0.f);
    case 2: return fminf(fmaxf(v, 0.f), 6.f);
    case 3: return tanhf(v);
    case 4: return 1.f / (1.f + __expf(-v));
    default: return v;
  }
}
__device__ __forceinline__ float act_bwd(float g, float y, int kind) {
  switch (kind) {
    case 1: return y > 0.f ? g : 0.f;
    case 2: return (y > 0.f && y < 6.f) ? g : 0.f;
    case 3: return g * (1.f - y * y);
    case 4: return g * y * (1.f - y);
    default: return g;
  }
}

__global__ void __launch_bounds__(256)
    act_fwd_kernel(const __half* __restrict__ x, long long ldx, __half* __restrict__ y,
                   long long ldy, long long rows, int C, int kind) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float v[8];
    load8(x + r * ldx + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = act_fwd(v[k], kind);
    store8(y + r * ldy + cv * 8, v);
  }
}

__global__ void __launch_bounds__(256)
    act_bwd_kernel(const __half* __restrict__ dy, long long lddy, const __half* __restrict__ y,
                   long long ldy, __half* __restrict__ dx, long long lddx, int accumulate,
                   long long rows, int C, int kind) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float g[8], o[8];
    load8(dy + r * lddy + cv * 8, g);
    load8(y + r * ldy + cv * 8, o);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = act_bwd(g[k], o[k], kind);
    if (accumulate) {
      load8(dx + r * lddx + cv * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += o[k];
    }
    store8(dx + r * lddx + cv * 8, g);
  }
}

// ---------------------------------------------------------------------------
// Depthwise convolution, channel multiplier 1. Filters are [R][S][C] (fp16 shadow of the
// fp32 master for fprop/dgrad, fp32 gradient out of wgrad).
// ---------------------------------------------------------------------------
struct DwGeom {
  int n, h, w, c, p, q, r, s, sh, sw, pad_h, pad_w;
};

// one thread = one output pixel x 8 channels
__global__ void __launch_bounds__(256)
    dw_fprop_kernel(const __half* __restrict__ x, long long ldx, const __half* __restrict__ wt,
                    __half* __restrict__ y, long long ldy, DwGeom g) {
  const int vpr = g.c >> 3;
  const long long total = static_cast<long long>(g.n) * g.p * g.q * vpr;
  const bool fits32 = total < 0xFFFFFFFFll;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    int cv, q, p, n;
    split_index(i, fits32, vpr, g.q, g.p, cv, q, p, n);
    const int h0 = p * g.sh - g.pad_h, w0 = q * g.sw - g.pad_w;
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int r = 0; r < g.r; ++r) {
      const int hh = h0 + r;
      if (hh < 0 || hh >= g.h) continue;
      for (int s = 0; s < g.s; ++s) {
        const int ww = w0 + s;
        if (ww < 0 || ww >= g.w) continue;
        float v[8], f[8];
        load8(x + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * ldx + cv * 8, v);
        load8(wt + static_cast<long long>(r * g.s + s) * g.c + cv * 8, f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(v[k], f[k], acc[k]);
      }
    }
    const long long orow = (static_cast<long long>(n) * g.p + p) * g.q + q;
    store8(y + orow * ldy + cv * 8, acc);
  }
}

// one thread = one input pixel x 8 channels; gathers the output pixels whose window covers it
__global__ void __launch_bounds__(256)
    dw_dgrad_kernel(const __half* __restrict__ dy, long long lddy, const __half* __restrict__ wt,
                    __half* __restrict__ dx, long long lddx, int accumulate, DwGeom g) {
  const int vpr = g.c >> 3;
  const long long total = static_cast<long long>(g.n) * g.h * g.w * vpr;
  const bool fits32 = total < 0xFFFFFFFFll;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    int cv, ww, hh, n;
    split_index(i, fits32, vpr, g.w, g.h, cv, ww, hh, n);
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int r = 0; r < g.r; ++r) {
      const int ph = hh + g.pad_h - r;
      if (ph < 0 || ph % g.sh != 0) continue;
      const int p = ph / g.sh;
      if (p >= g.p) continue;
      for (int s = 0; s < g.s; ++s) {
        const int pw = ww + g.pad_w - s;
        if (pw < 0 || pw % g.sw != 0) continue;
        const int q = pw / g.sw;
        if (q >= g.q) continue;
        float v[8], f[8];
        load8(dy + ((static_cast<long long>(n) * g.p + p) * g.q + q) * lddy + cv * 8, v);
        load8(wt + static_cast<long long>(r * g.s + s) * g.c + cv * 8, f);
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = fmaf(v[k], f[k], acc[k]);
      }
    }
    const long long irow = (static_cast<long long>(n) * g.h + hh) * g.w + ww;
    if (accumulate) {
      float o[8];
      load8(dx + irow * lddx + cv * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += o[k];
    }
    store8(dx + irow * lddx + cv * 8, acc);
  }
}

// wgrad stage 1. grid = (slabs, taps, channel chunks of 32 vectors); block = 32 x 8:
// threadIdx.x = channel vector, threadIdx.y = pixel lane. partial[slab][tap][C].
constexpr int kDwLanes = 8;
__global__ void __launch_bounds__(32 * kDwLanes)
    dw_wgrad_kernel(const __half* __restrict__ x, long long ldx, const __half* __restrict__ dy,
                    long long lddy, float* __restrict__ partial, DwGeom g, long long pix_per_slab) {
  __shared__ float red[kDwLanes][32][8];
  const int vpr = g.c >> 3;
  const int cv = blockIdx.z * 32 + threadIdx.x;
  const int tap = blockIdx.y;
  const int r = tap / g.s, s = tap - r * g.s;
  const long long npix = static_cast<long long>(g.n) * g.p * g.q;
  const long long lo = blockIdx.x * pix_per_slab;
  long long hi = lo + pix_per_slab;
  if (hi > npix) hi = npix;
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  if (cv < vpr) {
    for (long long pix = lo + threadIdx.y; pix < hi; pix += kDwLanes) {
      const int q = static_cast<int>(pix % g.q);
      long long t = pix / g.q;
      const int p = static_cast<int>(t % g.p);
      const int n = static_cast<int>(t / g.p);
      const int hh = p * g.sh - g.pad_h + r, ww = q * g.sw - g.pad_w + s;
      if (hh < 0 || hh >= g.h || ww < 0 || ww >= g.w) continue;
      float a[8], b[8];
      load8(x + ((static_cast<long long>(n) * g.h + hh) * g.w + ww) * ldx + cv * 8, a);
      load8(dy + pix * lddy + cv * 8, b);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] = fmaf(a[k], b[k], acc[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) red[threadIdx.y][threadIdx.x][k] = acc[k];
  __syncthreads();
  if (threadIdx.y == 0 && cv < vpr) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float sum = 0.f;
      for (int l = 0; l < kDwLanes; ++l) sum += red[l][threadIdx.x][k];
      partial[(static_cast<long long>(blockIdx.x) * gridDim.y + tap) * g.c + cv * 8 + k] = sum;
    }
  }
}

// wgrad stage 2: dw[tap][c] (+)= scale * sum_slab partial[slab][tap][c] (fixed order).
__global__ void dw_wgrad_fold_kernel(const float* __restrict__ partial, int slabs, int elems,
                                     float* __restrict__ dw, int accumulate, float scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= elems) return;
  float s = 0.f;
  for (int b = 0; b < slabs; ++b) s += partial[static_cast<long long>(b) * elems + i];
  dw[i] = (accumulate ? dw[i] : 0.f) + s * scale;
}

static inline int dw_slabs(const DwGeom& g) {
  const long long npix = static_cast<long long>(g.n) * g.p * g.q;
  const int chunks = ((g.c >> 3) + 31) / 32;
  long long want = (148 * 8) / (static_cast<long long>(g.r) * g.s * chunks);
  if (want < 1) want = 1;
  const long long max_by_work = (npix + 255) / 256;  // at least 256 pixels per slab
  if (want > max_by_work) want = max_by_work;
  if (want < 1) want = 1;
  return static_cast<int>(want);
}

// ---------------------------------------------------------------------------
// Drop-path (NASNet): every sample of the batch is kept with probability keep_prob and
// rescaled by 1/keep_prob, or zeroed. The decision is a pure function of (seed, offset, n),
// so the backward pass regenerates it instead of storing a mask.
// ---------------------------------------------------------------------------
__device__ __forceinline__ float sample_uniform(unsigned long long seed, unsigned long long offset,
                                                unsigned int n) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ull * (offset + n + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return static_cast<float>(z >> 40) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(256)
    drop_path_kernel(const __half* __restrict__ x, long long ldx, __half* __restrict__ y,
                     long long ldy, int accumulate, long long rows, long long rows_per_sample,
                     int C, float keep_prob, unsigned long long seed, unsigned long long offset,
                     const unsigned long long* __restrict__ step, int step_shift,
                     const float* __restrict__ keep_dev) {
  // device-resident step counter / keep probability (CUDA-graph replay)
  if (step != nullptr) offset += (*step) << step_shift;
  if (keep_dev != nullptr) keep_prob = *keep_dev;
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  const float inv = 1.f / keep_prob;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    const unsigned int n = static_cast<unsigned int>(r / rows_per_sample);
    const float scale = sample_uniform(seed, offset, n) < keep_prob ? inv : 0.f;
    float v[8];
    load8(x + r * ldx + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] *= scale;
    if (accumulate) {
      float o[8];
      load8(y + r * ldy + cv * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += o[k];
    }
    store8(y + r * ldy + cv * 8, v);
  }
}

// ---------------------------------------------------------------------------
// Channel gather: y[r][c] (+)= idx[c] >= 0 ? x[r][idx[c]] : 0. Converts between two phantom-
// channel layouts of the same logical tensor (nn/legalize.py); its backward is the same
// kernel with the inverse index.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    channel_gather_kernel(const __half* __restrict__ x, long long ldx, __half* __restrict__ y,
                          long long ldy, int accumulate, long long rows, int C,
                          const int* __restrict__ idx) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int src = idx[cv * 8 + k];
      v[k] = src >= 0 ? __half2float(x[r * ldx + src]) : 0.f;
    }
    if (accumulate) {
      float o[8];
      load8(y + r * ldy + cv * 8, o);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] += o[k];
    }
    store8(y + r * ldy + cv * 8, v);
  }
}

// ---------------------------------------------------------------------------
// Embedding lookup (NCF, T/models/experimental/official_ncf_model.py:62-83): one row of the
// table per example, optionally only a column range of it (the MF / MLP halves of NeuMF share
// one table). The gradient is dense (the reference's --sparse_to_dense_grads): the executor
// zeroes the table's gradient and this kernel scatter-adds into it.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    embedding_gather_kernel(const __half* __restrict__ table, long long ldt, int col_lo,
                            const int* __restrict__ ids, long long num_rows_table,
                            __half* __restrict__ y, long long ldy, long long n, int cols) {
  const int vpr = cols >> 3;
  const long long total = n * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    long long id = ids[r];
    id = id < 0 ? 0 : (id >= num_rows_table ? num_rows_table - 1 : id);
    *reinterpret_cast<Half8*>(y + r * ldy + cv * 8) =
        *reinterpret_cast<const Half8*>(table + id * ldt + col_lo + cv * 8);
  }
}

__global__ void __launch_bounds__(256)
    embedding_scatter_kernel(const __half* __restrict__ dy, long long lddy,
                             const int* __restrict__ ids, long long num_rows_table,
                             float* __restrict__ dtable, long long ldt, int col_lo, long long n,
                             int cols) {
  const int vpr = cols >> 3;
  const long long total = n * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    long long id = ids[r];
    id = id < 0 ? 0 : (id >= num_rows_table ? num_rows_table - 1 : id);
    float g[8];
    load8(dy + r * lddy + cv * 8, g);
    float* dst = dtable + id * ldt + col_lo + cv * 8;
#pragma unroll
    for (int k = 0; k < 8; ++k) atomicAdd(dst + k, g[k]);
  }
}

// y = a * b ; backward: da (+)= dy * b, db (+)= dy * a
__global__ void __launch_bounds__(256)
    mul_fwd_kernel(const __half* __restrict__ a, long long lda, const __half* __restrict__ b,
                   long long ldb, __half* __restrict__ y, long long ldy, long long rows, int C) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float u[8], v[8];
    load8(a + r * lda + cv * 8, u);
    load8(b + r * ldb + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) u[k] *= v[k];
    store8(y + r * ldy + cv * 8, u);
  }
}

__global__ void __launch_bounds__(256)
    mul_bwd_kernel(const __half* __restrict__ dy, long long lddy, const __half* __restrict__ other,
                   long long ldo, __half* __restrict__ dx, long long lddx, int accumulate,
                   long long rows, int C) {
  const int vpr = C >> 3;
  const long long total = rows * vpr;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / vpr;
    const int cv = static_cast<int>(i - r * vpr);
    float g[8], v[8];
    load8(dy + r * lddy + cv * 8, g);
    load8(other + r * ldo + cv * 8, v);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] *= v[k];
    if (accumulate) {
      load8(dx + r * lddx + cv * 8, v);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += v[k];
    }
    store8(dx + r * lddx + cv * 8, g);
  }
}

// NCF loss: 2-class softmax cross-entropy over [1, z] (the reference prepends a column of ones,
// T/models/experimental/official_ncf_model.py:85-97). One block; deterministic tree reduction.
// out[0] = mean loss, out[1] = number of correct predictions (as float), dz = scale*(p1 - y).
__global__ void __launch_bounds__(1024)
    ncf_loss_kernel(const float* __restrict__ z, long long ldz, const int* __restrict__ labels,
                    int n, float scale, const float* __restrict__ dyn_scale,
                    __half* __restrict__ dz, long long lddz, float* __restrict__ out) {
  __shared__ float red[2][32];
  const float sc = scale * (dyn_scale != nullptr ? *dyn_scale : 1.f);
  float loss = 0.f, hits = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float v = z[i * ldz];
    const int y = labels[i] != 0;
    const float m = fmaxf(v, 1.f);
    const float lse = m + __logf(__expf(1.f - m) + __expf(v - m));
    loss += lse - (y ? v : 1.f);
    const float p1 = __expf(v - lse);
    hits += ((v > 1.f) == (y != 0)) ? 1.f : 0.f;
    if (dz != nullptr) dz[i * lddz] = __float2half_rn(sc * (p1 - static_cast<float>(y)));
  }
  loss = warp_sum(loss);
  hits = warp_sum(hits);
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = loss;
    red[1][threadIdx.x >> 5] = hits;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = (blockDim.x + 31) >> 5;
    float a = threadIdx.x < nw ? red[0][threadIdx.x] : 0.f;
    float b = threadIdx.x < nw ? red[1][threadIdx.x] : 0.f;
    a = warp_sum(a);
    b = warp_sum(b);
    if (threadIdx.x == 0) {
      out[0] = a / static_cast<float>(n);
      out[1] = b;
      out[2] = b;
      out[3] = 0.f;
    }
  }
}

}  // namespace kfb

using namespace kfb;

#define KFB_REQ_C8(C, who) \
  KFB_REQUIRE((C) > 0 && (C) % 8 == 0, who ": C (%d) must be a multiple of 8", (C))

extern "C" int kfb_act_forward(const void* x, long long ldx, void* y, long long ldy,
                               long long rows, int C, int kind, void* stream) {
  KFB_REQ_C8(C, "act_forward");
  KFB_REQUIRE(kind >= 1 && kind <= 4, "act_forward: unknown kind %d", kind);
  act_fwd_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (__half*)y, ldy, rows, C, kind);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_act_backward(const void* dy, long long lddy, const void* y, long long ldy,
                                void* dx, long long lddx, int accumulate, long long rows, int C,
                                int kind, void* stream) {
  KFB_REQ_C8(C, "act_backward");
  KFB_REQUIRE(kind >= 1 && kind <= 4, "act_backward: unknown kind %d", kind);
  act_bwd_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, lddy, (const __half*)y, ldy, (__half*)dx, lddx, accumulate, rows, C, kind);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int fill_dw(DwGeom* g, const kfb_conv_desc* d, const char* who) {
  KFB_REQUIRE(d->c > 0 && d->c % 8 == 0, "%s: C (%d) must be a multiple of 8", who, d->c);
  KFB_REQUIRE(d->k == d->c, "%s: depthwise needs k == c (channel multiplier 1)", who);
  KFB_REQUIRE(d->r > 0 && d->s > 0 && d->stride_h > 0 && d->stride_w > 0, "%s: bad geometry", who);
  *g = DwGeom{d->n, d->h, d->w, d->c, d->p, d->q, d->r, d->s, d->stride_h, d->stride_w,
              d->pad_h_lo, d->pad_w_lo};
  return 0;
}

extern "C" int kfb_dwconv_fprop(const kfb_conv_desc* d, const void* x, long long ldx,
                                const void* w, void* y, long long ldy, void* stream) {
  DwGeom g;
  if (fill_dw(&g, d, "dwconv_fprop")) return 2;
  dw_fprop_kernel<<<xgrid_for((long long)g.n * g.p * g.q * (g.c / 8)), 256, 0,
                    (cudaStream_t)stream>>>((const __half*)x, ldx, (const __half*)w, (__half*)y,
                                            ldy, g);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_dwconv_dgrad(const kfb_conv_desc* d, const void* dy, long long lddy,
                                const void* w, void* dx, long long lddx, int accumulate,
                                void* stream) {
  DwGeom g;
  if (fill_dw(&g, d, "dwconv_dgrad")) return 2;
  dw_dgrad_kernel<<<xgrid_for((long long)g.n * g.h * g.w * (g.c / 8)), 256, 0,
                    (cudaStream_t)stream>>>((const __half*)dy, lddy, (const __half*)w, (__half*)dx,
                                            lddx, accumulate, g);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" long long kfb_dwconv_wgrad_workspace_bytes(const kfb_conv_desc* d) {
  DwGeom g;
  if (fill_dw(&g, d, "dwconv_wgrad_workspace_bytes")) return -1;
  return static_cast<long long>(dw_slabs(g)) * g.r * g.s * g.c * sizeof(float);
}

extern "C" int kfb_dwconv_wgrad(const kfb_conv_desc* d, const void* x, long long ldx,
                                const void* dy, long long lddy, float* dw, void* workspace,
                                long long workspace_bytes, int accumulate, float scale,
                                void* stream) {
  DwGeom g;
  if (fill_dw(&g, d, "dwconv_wgrad")) return 2;
  const int slabs = dw_slabs(g);
  const long long need = static_cast<long long>(slabs) * g.r * g.s * g.c * sizeof(float);
  KFB_REQUIRE(workspace != nullptr && workspace_bytes >= need,
              "dwconv_wgrad: workspace %lld bytes < %lld", workspace_bytes, need);
  const long long npix = static_cast<long long>(g.n) * g.p * g.q;
  const long long per = (npix + slabs - 1) / slabs;
  dim3 grid(slabs, g.r * g.s, ((g.c >> 3) + 31) / 32), block(32, kDwLanes);
  dw_wgrad_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (const __half*)dy, lddy, (float*)workspace, g, per);
  KFB_CHECK_CUDA(cudaGetLastError());
  const int elems = g.r * g.s * g.c;
  dw_wgrad_fold_kernel<<<(elems + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      (const float*)workspace, slabs, elems, dw, accumulate, scale);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_drop_path(const void* x, long long ldx, void* y, long long ldy, int accumulate,
                             long long rows, long long rows_per_sample, int C, float keep_prob,
                             unsigned long long seed, unsigned long long offset, void* stream) {
  KFB_REQ_C8(C, "drop_path");
  KFB_REQUIRE(keep_prob > 0.f && keep_prob <= 1.f, "drop_path: keep_prob %f out of (0,1]",
              (double)keep_prob);
  KFB_REQUIRE(rows_per_sample > 0 && rows % rows_per_sample == 0, "drop_path: bad row split");
  drop_path_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (__half*)y, ldy, accumulate, rows, rows_per_sample, C, keep_prob,
      seed, offset, nullptr, 0, nullptr);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// kfb_drop_path with the step part of the Philox offset (offset + (*step << step_shift)) and,
// optionally, the keep probability read from device memory.
extern "C" int kfb_drop_path_dev(const void* x, long long ldx, void* y, long long ldy,
                                 int accumulate, long long rows, long long rows_per_sample, int C,
                                 float keep_prob, unsigned long long seed,
                                 unsigned long long offset, const unsigned long long* step,
                                 int step_shift, const float* keep_dev, void* stream) {
  KFB_REQ_C8(C, "drop_path");
  KFB_REQUIRE(rows_per_sample > 0 && rows % rows_per_sample == 0, "drop_path: bad row split");
  drop_path_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (__half*)y, ldy, accumulate, rows, rows_per_sample, C, keep_prob,
      seed, offset, step, step_shift, keep_dev);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_channel_gather(const void* x, long long ldx, void* y, long long ldy,
                                  int accumulate, long long rows, int C, const int* idx,
                                  void* stream) {
  KFB_REQ_C8(C, "channel_gather");
  KFB_REQUIRE(idx != nullptr, "channel_gather: null index");
  channel_gather_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)x, ldx, (__half*)y, ldy, accumulate, rows, C, idx);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_embedding_gather(const void* table, long long ldt, int col_lo, const int* ids,
                                    long long num_rows_table, void* y, long long ldy, long long n,
                                    int cols, void* stream) {
  KFB_REQ_C8(cols, "embedding_gather");
  KFB_REQUIRE(col_lo % 8 == 0 && ldt % 8 == 0, "embedding_gather: col_lo/ld must be multiples of 8");
  embedding_gather_kernel<<<xgrid_for(n * (cols / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)table, ldt, col_lo, ids, num_rows_table, (__half*)y, ldy, n, cols);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_embedding_scatter_add(const void* dy, long long lddy, const int* ids,
                                         long long num_rows_table, float* dtable, long long ldt,
                                         int col_lo, long long n, int cols, void* stream) {
  KFB_REQ_C8(cols, "embedding_scatter_add");
  embedding_scatter_kernel<<<xgrid_for(n * (cols / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, lddy, ids, num_rows_table, dtable, ldt, col_lo, n, cols);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_mul_forward(const void* a, long long lda, const void* b, long long ldb, void* y,
                               long long ldy, long long rows, int C, void* stream) {
  KFB_REQ_C8(C, "mul_forward");
  mul_fwd_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)a, lda, (const __half*)b, ldb, (__half*)y, ldy, rows, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_mul_backward(const void* dy, long long lddy, const void* other, long long ldo,
                                void* dx, long long lddx, int accumulate, long long rows, int C,
                                void* stream) {
  KFB_REQ_C8(C, "mul_backward");
  mul_bwd_kernel<<<xgrid_for(rows * (C / 8)), 256, 0, (cudaStream_t)stream>>>(
      (const __half*)dy, lddy, (const __half*)other, ldo, (__half*)dx, lddx, accumulate, rows, C);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_ncf_loss(const float* z, long long ldz, const int* labels, int n, float scale,
                            const float* dyn_scale, void* dz, long long lddz, float* out,
                            void* stream) {
  KFB_REQUIRE(n > 0, "ncf_loss: empty batch");
  ncf_loss_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(z, ldz, labels, n, scale, dyn_scale,
                                                        (__half*)dz, lddz, out);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// LSTM cell (tf.nn.rnn_cell.BasicLSTMCell, forget_bias 1, gate order i, j, f, o -
// T/models/experimental/deepspeech.py:121-126,232-268). The matrix products of a layer are
// tcgen05 GEMMs (input projection for all steps at once, recurrent projection per step); these
// kernels are the pointwise part of one time step for the whole batch.
// ---------------------------------------------------------------------------
namespace kfb {

__device__ __forceinline__ float sigmoidf_(float v) { return 1.f / (1.f + __expf(-v)); }

// gates_io: in = x-projection pre-activations of this step, out = activated gates (saved for
// backward). gh = recurrent projection. Writes c, h (into the layer output) and h again into
// the "previous hidden state" slot of the next step (operand of the recurrent wgrad GEMM).
__global__ void __launch_bounds__(256)
    lstm_cell_fwd_kernel(__half* __restrict__ gates_io, long long ldg,
                         const __half* __restrict__ gh, long long ldgh,
                         const float* __restrict__ bias, const float* __restrict__ c_prev,
                         long long ldcp, float* __restrict__ c_out, long long ldc,
                         __half* __restrict__ h_out, long long ldh,
                         __half* __restrict__ h_next, long long ldhn, int n, int H,
                         float forget_bias) {
  const int total = n * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / H, u = i - b * H;
    __half* g = gates_io + static_cast<long long>(b) * ldg;
    const __half* r = gh + static_cast<long long>(b) * ldgh;
    const float gi = __half2float(g[u]) + __half2float(r[u]) + bias[u];
    const float gj = __half2float(g[H + u]) + __half2float(r[H + u]) + bias[H + u];
    const float gf = __half2float(g[2 * H + u]) + __half2float(r[2 * H + u]) + bias[2 * H + u];
    const float go = __half2float(g[3 * H + u]) + __half2float(r[3 * H + u]) + bias[3 * H + u];
    const float ai = sigmoidf_(gi), aj = tanhf(gj), af = sigmoidf_(gf + forget_bias),
                ao = sigmoidf_(go);
    const float cp = c_prev != nullptr ? c_prev[static_cast<long long>(b) * ldcp + u] : 0.f;
    const float c = cp * af + ai * aj;
    const float h = tanhf(c) * ao;
    g[u] = __float2half_rn(ai);
    g[H + u] = __float2half_rn(aj);
    g[2 * H + u] = __float2half_rn(af);
    g[3 * H + u] = __float2half_rn(ao);
    c_out[static_cast<long long>(b) * ldc + u] = c;
    const __half hh = __float2half_rn(h);
    h_out[static_cast<long long>(b) * ldh + u] = hh;
    if (h_next != nullptr) h_next[static_cast<long long>(b) * ldhn + u] = hh;
  }
}

// gates_io: in = activated gates of this step, out = gradients w.r.t. the pre-activations.
// dh = dy (this step's slice of the output gradient) + dh_rec (from the following step).
// dc_io: in = dL/dc from the following step, out = dL/dc for the preceding step.
__global__ void __launch_bounds__(256)
    lstm_cell_bwd_kernel(__half* __restrict__ gates_io, long long ldg,
                         const __half* __restrict__ dy, long long lddy,
                         const __half* __restrict__ dh_rec, long long lddh,
                         const float* __restrict__ c, long long ldc,
                         const float* __restrict__ c_prev, long long ldcp,
                         float* __restrict__ dc_io, int n, int H) {
  const int total = n * H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int b = i / H, u = i - b * H;
    __half* g = gates_io + static_cast<long long>(b) * ldg;
    const float ai = __half2float(g[u]), aj = __half2float(g[H + u]),
                af = __half2float(g[2 * H + u]), ao = __half2float(g[3 * H + u]);
    float dh = __half2float(dy[static_cast<long long>(b) * lddy + u]);
    if (dh_rec != nullptr) dh += __half2float(dh_rec[static_cast<long long>(b) * lddh + u]);
    const float tc = tanhf(c[static_cast<long long>(b) * ldc + u]);
    const float cp = c_prev != nullptr ? c_prev[static_cast<long long>(b) * ldcp + u] : 0.f;
    const float dc = dc_io[i] + dh * ao * (1.f - tc * tc);
    g[u] = __float2half_rn(dc * aj * ai * (1.f - ai));
    g[H + u] = __float2half_rn(dc * ai * (1.f - aj * aj));
    g[2 * H + u] = __float2half_rn(dc * cp * af * (1.f - af));
    g[3 * H + u] = __float2half_rn(dh * tc * ao * (1.f - ao));
    dc_io[i] = dc * af;
  }
}

}  // namespace kfb

extern "C" int kfb_lstm_cell_forward(void* gates_io, long long ldg, const void* gh, long long ldgh,
                                     const float* bias, const float* c_prev, long long ldcp,
                                     float* c_out, long long ldc, void* h_out, long long ldh,
                                     void* h_next, long long ldhn, int n, int H,
                                     float forget_bias, void* stream) {
  KFB_REQUIRE(n > 0 && H > 0, "lstm_cell_forward: bad sizes");
  lstm_cell_fwd_kernel<<<xgrid_for((long long)n * H), 256, 0, (cudaStream_t)stream>>>(
      (__half*)gates_io, ldg, (const __half*)gh, ldgh, bias, c_prev, ldcp, c_out, ldc,
      (__half*)h_out, ldh, (__half*)h_next, ldhn, n, H, forget_bias);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_lstm_cell_backward(void* gates_io, long long ldg, const void* dy, long long lddy,
                                      const void* dh_rec, long long lddh, const float* c,
                                      long long ldc, const float* c_prev, long long ldcp,
                                      float* dc_io, int n, int H, void* stream) {
  KFB_REQUIRE(n > 0 && H > 0, "lstm_cell_backward: bad sizes");
  lstm_cell_bwd_kernel<<<xgrid_for((long long)n * H), 256, 0, (cudaStream_t)stream>>>(
      (__half*)gates_io, ldg, (const __half*)dy, lddy, (const __half*)dh_rec, lddh, c, ldc, c_prev,
      ldcp, dc_io, n, H);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
