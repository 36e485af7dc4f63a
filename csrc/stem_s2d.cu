// Space-to-depth form of a tiny-C stride-2 stem convolution (ResNet's 7x7/2 on 3 channels;
// the reference leaves it to cuDNN, T/models/resnet_model.py:283-286 + convnet_builder conv).
//
// The implicit-GEMM kernel wants 64-channel K blocks; a 3-channel input fills 3/64 of them, and
// the explicit im2col this repo used before writes (and the GEMM re-reads) a 976 MB column
// buffer per step at batch 256. Instead:
//   * the image is regrouped into 2x2 blocks: s2d[n, A, B, (bh*2+bw)*C + c] = x[n, 2A-eh+bh,
//     2B-ew+bw, c] (16 channels per pixel, 4*C used), which turns the RxS/2 convolution into a
//     ceil(R/2) x ceil(S/2) unit-stride one;
//   * the four horizontal taps of that convolution are four NEIGHBOURING 16-channel pixels, i.e.
//     64 contiguous halves: a TMA im2col map whose pixel stride is 16 elements (32 B) but whose
//     pixel size is 64 channels reads them as one dense 64-channel K block. The overlap exists
//     only in the tensor map - the s2d buffer (103 MB at batch 256) is stored once, with the
//     horizontal zero padding materialised so that the overlapping reads never cross a row end.
// The convolution and its weight gradient then run on the ordinary FPROP / WGRAD tcgen05 paths
// as a ceil(R/2) x 1 convolution over 64 "channels"; kfb_stem_pack_w / kfb_stem_unpack_dw map
// the [K, R*S*C] weights and gradients to and from that form.
#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

// One thread per (n, A, jp) pixel of the padded s2d buffer. expand = 0: y[n, A, jp, 16];
// expand = 1 (fallback without overlapping strides): y[n, A, q, t*16 + ch] for q = jp - t.
template <int C>
__global__ void __launch_bounds__(256)
    stem_s2d_kernel(const __half* __restrict__ x, __half* __restrict__ y, int N, int H, int W,
                    int H2, int W2p, int eh, int ew, int lw, int expand, int Q) {
  const long long total = (long long)N * H2 * W2p;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int jp = (int)(idx % W2p);
  const long long t2 = idx / W2p;
  const int A = (int)(t2 % H2);
  const int n = (int)(t2 / H2);
  __align__(16) __half v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __float2half(0.f);
#pragma unroll
  for (int bh = 0; bh < 2; ++bh) {
    const int i = 2 * A - eh + bh;
    if (i < 0 || i >= H) continue;
#pragma unroll
    for (int bw = 0; bw < 2; ++bw) {
      const int j = 2 * (jp - lw) - ew + bw;
      if (j < 0 || j >= W) continue;
      const __half* src = x + (((long long)n * H + i) * W + j) * C;
#pragma unroll
      for (int c = 0; c < C; ++c) v[(bh * 2 + bw) * C + c] = src[c];
    }
  }
  const uint4* vv = reinterpret_cast<const uint4*>(v);
  if (!expand) {
    uint4* dst = reinterpret_cast<uint4*>(y + idx * 16);
    dst[0] = vv[0];
    dst[1] = vv[1];
  } else {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int q = jp - t;
      if (q < 0 || q >= Q) continue;
      uint4* dst = reinterpret_cast<uint4*>(y + ((((long long)n * H2 + A) * Q + q) * 64 + t * 16));
      dst[0] = vv[0];
      dst[1] = vv[1];
    }
  }
}

// w[K, kpad] (columns (r*S + s)*C + c) -> w4[K, TA, 64] (columns aw*16 + (bh*2+bw)*C + c of
// vertical tap ah, r = 2*ah + bh, s = 2*aw + bw)
__global__ void stem_pack_w_kernel(const __half* __restrict__ w, __half* __restrict__ w4, int K,
                                   int kpad, int R, int S, int C, int TA) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= K * TA * 64) return;
  const int col = idx & 63;
  const int ah = (idx >> 6) % TA;
  const int k = (idx >> 6) / TA;
  const int aw = col >> 4, ch = col & 15;
  __half val = __float2half(0.f);
  if (ch < 4 * C) {
    const int blk = ch / C, c = ch - blk * C;
    const int r = 2 * ah + (blk >> 1), s = 2 * aw + (blk & 1);
    if (r < R && s < S) val = w[(long long)k * kpad + (r * S + s) * C + c];
  }
  w4[idx] = val;
}

// dw[K, kpad] (+)= scale * dw4[K, TA, 64] gathered back; pad columns are zeroed (or kept)
__global__ void stem_unpack_dw_kernel(const float* __restrict__ dw4, float* __restrict__ dw, int K,
                                      int kpad, int R, int S, int C, int TA, int accumulate,
                                      float scale) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= K * kpad) return;
  const int col = idx % kpad, k = idx / kpad;
  float val = 0.f;
  if (col < R * S * C) {
    const int tap = col / C, c = col - tap * C;
    const int r = tap / S, s = tap - r * S;
    const int ch = ((r & 1) * 2 + (s & 1)) * C + c;
    val = scale * dw4[((long long)k * TA + (r >> 1)) * 64 + (s >> 1) * 16 + ch];
  }
  dw[idx] = accumulate ? dw[idx] + val : val;
}

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_stem_s2d(const void* x, void* y, int n, int h, int w, int c, int h2, int w2p,
                            int eh, int ew, int lw, int expand, int q, void* stream) {
  KFB_REQUIRE(n > 0 && h > 0 && w > 0 && c >= 1 && c <= 4 && h2 > 0 && w2p > 0,
              "stem_s2d: bad geometry (C = %d must be 1..4)", c);
  KFB_REQUIRE((reinterpret_cast<uintptr_t>(y) & 15) == 0, "stem_s2d: output must be 16-byte aligned");
  const long long total = (long long)n * h2 * w2p;
  const long long blocks = (total + 255) / 256;
  KFB_REQUIRE(blocks < (1ll << 31), "stem_s2d: too many pixels");
#define KFB_S2D(TC)                                                                     \
  stem_s2d_kernel<TC><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(              \
      static_cast<const __half*>(x), static_cast<__half*>(y), n, h, w, h2, w2p, eh, ew, lw, \
      expand, q)
  switch (c) {
    case 1: KFB_S2D(1); break;
    case 2: KFB_S2D(2); break;
    case 3: KFB_S2D(3); break;
    default: KFB_S2D(4); break;
  }
#undef KFB_S2D
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_stem_pack_w(const void* w, void* w4, int k, int kpad, int r, int s, int c,
                               int ta, void* stream) {
  KFB_REQUIRE(k > 0 && ta > 0 && (s + 1) / 2 <= 4 && c >= 1 && c <= 4 && kpad >= r * s * c,
              "stem_pack_w: bad shape");
  const int total = k * ta * 64;
  stem_pack_w_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      static_cast<const __half*>(w), static_cast<__half*>(w4), k, kpad, r, s, c, ta);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int kfb_stem_unpack_dw(const float* dw4, float* dw, int k, int kpad, int r, int s, int c,
                                  int ta, int accumulate, float scale, void* stream) {
  KFB_REQUIRE(k > 0 && ta > 0 && (s + 1) / 2 <= 4 && c >= 1 && c <= 4 && kpad >= r * s * c,
              "stem_unpack_dw: bad shape");
  const int total = k * kpad;
  stem_unpack_dw_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      dw4, dw, k, kpad, r, s, c, ta, accumulate, scale);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
