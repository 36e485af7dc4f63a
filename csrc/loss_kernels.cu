// Loss heads beyond softmax cross-entropy (sm_100a): the SSD-300 multibox loss with hard
// negative mining, reading the twelve head convolutions in place and writing their
// gradients in place (no transpose / concat of the [N, 8732, 85] "logits" tensor the
// reference materialises, T/models/ssd_model.py:138-221, 299-384).
//
// Anchor order (T/ssd_dataloader.py:35-79, the NCHW reshape of T/models/ssd_model.py:196-210):
// feature maps 38,19,10,5,3,1 in turn; within a map default-box-major, then row, then column.
// Anchor (d, y, x) of a map with D default boxes reads channels [d*4, d*4+4) of the loc head
// and [d*L, d*L+L) of the conf head at pixel (y, x) - contiguous in NHWC.
#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

constexpr int kSsdMaps = 6;

struct SsdHeads {
  const __half* loc[kSsdMaps];
  const __half* conf[kSsdMaps];
  __half* dloc[kSsdMaps];
  __half* dconf[kSsdMaps];
  long long ld_loc[kSsdMaps], ld_conf[kSsdMaps];
  int size[kSsdMaps];     // feature map side S
  int ndef[kSsdMaps];     // default boxes per pixel D
  int first[kSsdMaps + 1];  // first anchor index of each map
  int num_anchors, num_classes, batch;
};

struct AnchorRef {
  int map, d;
  long long pixel;  // (n*S + y)*S + x
};

__device__ __forceinline__ AnchorRef locate(const SsdHeads& h, int n, int j) {
  int m = 0;
#pragma unroll
  for (int i = 1; i < kSsdMaps; ++i)
    if (j >= h.first[i]) m = i;
  const int local = j - h.first[m];
  const int ss = h.size[m] * h.size[m];
  AnchorRef a;
  a.map = m;
  a.d = local / ss;
  a.pixel = static_cast<long long>(n) * ss + (local - a.d * ss);
  return a;
}

// One warp per anchor: per-anchor cross-entropy and log-sum-exp.
__global__ void __launch_bounds__(256)
    ssd_xent_kernel(SsdHeads h, const float* __restrict__ gt_label, float* __restrict__ xent,
                    float* __restrict__ lse_out) {
  const int lane = threadIdx.x & 31;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long total = static_cast<long long>(h.batch) * h.num_anchors;
  if (warp >= total) return;
  const int n = static_cast<int>(warp / h.num_anchors);
  const int j = static_cast<int>(warp - static_cast<long long>(n) * h.num_anchors);
  const AnchorRef a = locate(h, n, j);
  const __half* p = h.conf[a.map] + a.pixel * h.ld_conf[a.map] + a.d * h.num_classes;
  const int L = h.num_classes;
  float mx = -3.0e38f;
  for (int c = lane; c < L; c += 32) mx = fmaxf(mx, __half2float(p[c]));
  mx = warp_max(mx);
  float sum = 0.f;
  for (int c = lane; c < L; c += 32) sum += __expf(__half2float(p[c]) - mx);
  sum = warp_sum(sum);
  if (lane == 0) {
    const float lse = mx + __logf(sum);
    int y = static_cast<int>(gt_label[warp]);
    y = y < 0 ? 0 : (y >= L ? L - 1 : y);
    xent[warp] = lse - __half2float(p[y]);
    lse_out[warp] = lse;
  }
}

// One block per sample: hard-negative mining by a bitonic sort of (loss, index) keys, the
// per-sample losses and the per-anchor weight (positives + selected negatives).
constexpr int kSortN = 16384;  // next power of two above 8732
__global__ void __launch_bounds__(1024)
    ssd_mine_kernel(SsdHeads h, const float* __restrict__ gt_loc,
                    const float* __restrict__ gt_label, const float* __restrict__ num_matched,
                    const float* __restrict__ xent, int negs_per_positive,
                    float* __restrict__ weight, float* __restrict__ sample_loss) {
  extern __shared__ unsigned long long keys[];  // kSortN
  __shared__ float red[2][32];
  const int n = blockIdx.x;
  const int A = h.num_anchors;
  const float* xe = xent + static_cast<long long>(n) * A;
  const float* gl = gt_label + static_cast<long long>(n) * A;
  for (int i = threadIdx.x; i < kSortN; i += blockDim.x) {
    unsigned long long k = 0ull;  // padding sorts last
    if (i < A) {
      const bool pos = static_cast<int>(gl[i]) > 0;
      const float v = pos ? 0.f : fmaxf(xe[i], 0.f);  // positives take part with loss 0
      // descending by value, ties by ascending index (stable argsort)
      k = (static_cast<unsigned long long>(__float_as_uint(v)) << 32) |
          static_cast<unsigned long long>(0xFFFFFFFFu - static_cast<unsigned>(i));
      k += 1ull << 63;  // keep real entries above the zero padding
    }
    keys[i] = k;
  }
  __syncthreads();
  for (int size = 2; size <= kSortN; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = threadIdx.x; t < kSortN / 2; t += blockDim.x) {
        const int lo = (t / stride) * (stride << 1) + (t % stride);
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const unsigned long long a = keys[lo], b = keys[hi];
        if ((a < b) == desc) {
          keys[lo] = b;
          keys[hi] = a;
        }
      }
      __syncthreads();
    }
  }
  const float nm = num_matched[n];
  int k = static_cast<int>(nm) * negs_per_positive;
  if (k > A) k = A;
  float* wn = weight + static_cast<long long>(n) * A;
  for (int i = threadIdx.x; i < A; i += blockDim.x)
    wn[i] = static_cast<int>(gl[i]) > 0 ? 1.f : 0.f;
  __syncthreads();
  // ranks 0..k-1 of the sorted order are the selected negatives (a positive that lands there
  // already has weight 1, exactly as mask + top_k_mask would give 2 only if its loss were
  // among the largest negatives' - the reference masks positives to 0 loss, so they rank last
  // unless fewer than k negatives have positive loss; reproduce by adding, not setting)
  for (int r = threadIdx.x; r < k; r += blockDim.x) {
    const unsigned idx = 0xFFFFFFFFu - static_cast<unsigned>(keys[r] & 0xFFFFFFFFull);
    if (idx < static_cast<unsigned>(A)) atomicAdd(&wn[idx], 1.f);
  }
  __syncthreads();
  // per-sample losses: class = sum xent*w ; box = sum_{pos} huber(loc - gt)
  float cls = 0.f, box = 0.f;
  for (int j = threadIdx.x; j < A; j += blockDim.x) {
    cls += xe[j] * wn[j];
    if (static_cast<int>(gl[j]) > 0) {
      const AnchorRef a = locate(h, n, j);
      const __half* p = h.loc[a.map] + a.pixel * h.ld_loc[a.map] + a.d * 4;
      const float* g = gt_loc + (static_cast<long long>(n) * A + j) * 4;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float d = fabsf(__half2float(p[c]) - g[c]);
        box += d <= 1.f ? 0.5f * d * d : d - 0.5f;
      }
    }
  }
  cls = warp_sum(cls);
  box = warp_sum(box);
  if ((threadIdx.x & 31) == 0) {
    red[0][threadIdx.x >> 5] = cls;
    red[1][threadIdx.x >> 5] = box;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    float a = red[0][threadIdx.x], b = red[1][threadIdx.x];
    a = warp_sum(a);
    b = warp_sum(b);
    if (threadIdx.x == 0) {
      sample_loss[2 * n] = a / nm;
      sample_loss[2 * n + 1] = b / nm;
    }
  }
}

// One warp per anchor: gradients of both heads, written in place (every element of every head
// gradient is covered exactly once). scale already holds loss_scale / batch.
__global__ void __launch_bounds__(256)
    ssd_grad_kernel(SsdHeads h, const float* __restrict__ gt_loc,
                    const float* __restrict__ gt_label, const float* __restrict__ num_matched,
                    const float* __restrict__ lse, const float* __restrict__ weight, float scale,
                    const float* __restrict__ dyn_scale) {
  const int lane = threadIdx.x & 31;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long total = static_cast<long long>(h.batch) * h.num_anchors;
  if (warp >= total) return;
  const int n = static_cast<int>(warp / h.num_anchors);
  const int j = static_cast<int>(warp - static_cast<long long>(n) * h.num_anchors);
  const AnchorRef a = locate(h, n, j);
  const float sc = scale * (dyn_scale != nullptr ? *dyn_scale : 1.f) / num_matched[n];
  const int L = h.num_classes;
  int y = static_cast<int>(gt_label[warp]);
  y = y < 0 ? 0 : (y >= L ? L - 1 : y);
  const float w = weight[warp] * sc;
  const float l = lse[warp];
  const long long coff = a.pixel * h.ld_conf[a.map] + a.d * L;
  const __half* p = h.conf[a.map] + coff;
  __half* dp = h.dconf[a.map] + coff;
  for (int c = lane; c < L; c += 32) {
    const float sm = __expf(__half2float(p[c]) - l);
    dp[c] = __float2half_rn(w * (sm - (c == y ? 1.f : 0.f)));
  }
  if (lane < 4) {
    const long long loff = a.pixel * h.ld_loc[a.map] + a.d * 4;
    float g = 0.f;
    if (y > 0) {
      const float d = __half2float(h.loc[a.map][loff + lane]) -
                      gt_loc[warp * 4 + lane];
      g = (fabsf(d) <= 1.f ? d : (d > 0.f ? 1.f : -1.f)) * sc;
    }
    h.dloc[a.map][loff + lane] = __float2half_rn(g);
  }
}

// out[0] = mean over samples of (class + box), out[4] = class part, out[5] = box part
__global__ void ssd_finalize_kernel(const float* __restrict__ sample_loss, int n,
                                    float* __restrict__ out) {
  float c = 0.f, b = 0.f;
  for (int i = threadIdx.x; i < n; i += 32) {
    c += sample_loss[2 * i];
    b += sample_loss[2 * i + 1];
  }
  c = warp_sum(c);
  b = warp_sum(b);
  if (threadIdx.x == 0) {
    out[0] = (c + b) / n;
    out[1] = 0.f;
    out[2] = 0.f;
    out[3] = 0.f;
    out[4] = c / n;
    out[5] = b / n;
  }
}

}  // namespace kfb

using namespace kfb;

extern "C" long long kfb_ssd_loss_workspace_floats(int batch, int num_anchors) {
  // xent, lse, weight per anchor + 2 per sample
  return 3ll * batch * num_anchors + 2ll * batch;
}

extern "C" int kfb_ssd_loss(int batch, int num_classes, const int* map_size, const int* map_ndef,
                            void* const* loc, const long long* ld_loc, void* const* conf,
                            const long long* ld_conf, void* const* dloc, void* const* dconf,
                            const float* gt_loc, const float* gt_label, const float* num_matched,
                            int negs_per_positive, float scale, const float* dyn_scale,
                            float* workspace, long long workspace_floats, float* out,
                            void* stream) {
  SsdHeads h;
  int first = 0;
  for (int i = 0; i < kSsdMaps; ++i) {
    h.loc[i] = (const __half*)loc[i];
    h.conf[i] = (const __half*)conf[i];
    h.dloc[i] = dloc ? (__half*)dloc[i] : nullptr;
    h.dconf[i] = dconf ? (__half*)dconf[i] : nullptr;
    h.ld_loc[i] = ld_loc[i];
    h.ld_conf[i] = ld_conf[i];
    h.size[i] = map_size[i];
    h.ndef[i] = map_ndef[i];
    h.first[i] = first;
    first += map_size[i] * map_size[i] * map_ndef[i];
  }
  h.first[kSsdMaps] = first;
  h.num_anchors = first;
  h.num_classes = num_classes;
  h.batch = batch;
  KFB_REQUIRE(first <= kSortN, "ssd_loss: %d anchors exceed the sort capacity %d", first, kSortN);
  const long long need = kfb_ssd_loss_workspace_floats(batch, first);
  KFB_REQUIRE(workspace != nullptr && workspace_floats >= need,
              "ssd_loss: workspace %lld floats < %lld", workspace_floats, need);
  cudaStream_t st = (cudaStream_t)stream;
  const long long total = static_cast<long long>(batch) * first;
  float* xent = workspace;
  float* lse = xent + total;
  float* weight = lse + total;
  float* sample = weight + total;
  const int blocks = static_cast<int>((total * 32 + 255) / 256);
  ssd_xent_kernel<<<blocks, 256, 0, st>>>(h, gt_label, xent, lse);
  KFB_CHECK_CUDA(cudaGetLastError());
  static bool attr_set = false;
  const int smem = kSortN * sizeof(unsigned long long);
  if (!attr_set) {
    KFB_CHECK_CUDA(cudaFuncSetAttribute(ssd_mine_kernel,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  ssd_mine_kernel<<<batch, 1024, smem, st>>>(h, gt_loc, gt_label, num_matched, xent,
                                             negs_per_positive, weight, sample);
  KFB_CHECK_CUDA(cudaGetLastError());
  if (dloc != nullptr && dconf != nullptr) {
    ssd_grad_kernel<<<blocks, 256, 0, st>>>(h, gt_loc, gt_label, num_matched, lse, weight, scale,
                                            dyn_scale);
    KFB_CHECK_CUDA(cudaGetLastError());
  }
  ssd_finalize_kernel<<<1, 32, 0, st>>>(sample, batch, out);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
