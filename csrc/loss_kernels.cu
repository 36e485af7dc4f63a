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


// ---------------------------------------------------------------------------
// CTC loss (DeepSpeech2, T/models/experimental/deepspeech.py:359-396): the reference feeds
// log(softmax(z) + eps) to tf.nn.ctc_loss, which normalises again, so the emission
// probabilities are p = (softmax(z) + eps) / (1 + C*eps); blank is the last class; samples whose
// label does not fit their input length contribute zero loss and gradient
// (ignore_longer_outputs_than_inputs). One block per sample, log-space alpha/beta recursions
// over the 2L+1 extended-label states, alpha kept in global memory for the gradient pass.
// ---------------------------------------------------------------------------
constexpr float kCtcNegInf = -1.0e30f;

__device__ __forceinline__ float log_add(float a, float b) {
  const float m = fmaxf(a, b);
  if (m <= kCtcNegInf) return kCtcNegInf;
  return m + __logf(__expf(a - m) + __expf(b - m));
}

// lp[k] = log p_t(k) for the C classes of one time step; executed by warp 0.
__device__ __forceinline__ void ctc_log_probs(const float* __restrict__ z, int C, float eps,
                                              float* __restrict__ lp, float* __restrict__ q) {
  const int lane = threadIdx.x;
  float mx = -3.0e38f;
  for (int k = lane; k < C; k += 32) mx = fmaxf(mx, z[k]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int k = lane; k < C; k += 32) sum += __expf(z[k] - mx);
  sum = warp_sum(sum);
  const float norm = 1.f + C * eps;
  for (int k = lane; k < C; k += 32) {
    const float qk = __expf(z[k] - mx) / sum;
    if (q != nullptr) q[k] = qk;
    lp[k] = __logf((qk + eps) / norm);
  }
}

__global__ void __launch_bounds__(1024)
    ctc_alpha_kernel(const float* __restrict__ z, long long ldz, int T, int C,
                     const int* __restrict__ labels, int lmax, const int* __restrict__ in_len,
                     const int* __restrict__ lab_len, float eps, float* __restrict__ alpha,
                     float* __restrict__ sample_loss) {
  extern __shared__ float sm[];
  const int S_max = 2 * lmax + 1;
  float* a0 = sm;
  float* a1 = a0 + S_max;
  float* lp = a1 + S_max;
  int* ext = reinterpret_cast<int*>(lp + C);
  const int n = blockIdx.x;
  const int blank = C - 1;
  int L = lab_len[n];
  L = L < 0 ? 0 : (L > lmax ? lmax : L);
  int Tn = in_len[n];
  Tn = Tn < 0 ? 0 : (Tn > T ? T : Tn);
  const int S = 2 * L + 1;
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    int l = (s & 1) ? labels[n * lmax + (s >> 1)] : blank;
    ext[s] = l < 0 ? 0 : (l >= C ? C - 1 : l);
  }
  __syncthreads();
  float* prev = a0;
  float* cur = a1;
  float* an = alpha + static_cast<long long>(n) * T * S_max;
  for (int t = 0; t < Tn; ++t) {
    if (threadIdx.x < 32)
      ctc_log_probs(z + (static_cast<long long>(n) * T + t) * ldz, C, eps, lp, nullptr);
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
      float v;
      if (t == 0) {
        v = s < 2 ? 0.f : kCtcNegInf;
      } else {
        v = prev[s];
        if (s >= 1) v = log_add(v, prev[s - 1]);
        if (s >= 2 && ext[s] != blank && ext[s] != ext[s - 2]) v = log_add(v, prev[s - 2]);
      }
      v = v <= kCtcNegInf ? kCtcNegInf : v + lp[ext[s]];
      cur[s] = v;
      an[static_cast<long long>(t) * S_max + s] = v;
    }
    __syncthreads();
    float* tmp = prev;
    prev = cur;
    cur = tmp;
  }
  if (threadIdx.x == 0) {
    float logp = kCtcNegInf;
    if (Tn > 0) {
      logp = prev[S - 1];
      if (S >= 2) logp = log_add(logp, prev[S - 2]);
    }
    // infeasible alignment -> ignored (loss 0, no gradient); flagged with a +inf marker
    sample_loss[n] = logp <= kCtcNegInf * 0.5f ? 3.0e38f : -logp;
  }
}

__global__ void __launch_bounds__(1024)
    ctc_beta_grad_kernel(const float* __restrict__ z, long long ldz, int T, int C,
                         const int* __restrict__ labels, int lmax, const int* __restrict__ in_len,
                         const int* __restrict__ lab_len, float eps,
                         const float* __restrict__ alpha, const float* __restrict__ sample_loss,
                         float scale, const float* __restrict__ dyn_scale,
                         __half* __restrict__ dz, long long lddz) {
  extern __shared__ float sm[];
  const int S_max = 2 * lmax + 1;
  float* b0 = sm;
  float* b1 = b0 + S_max;
  float* lp = b1 + S_max;
  float* q = lp + C;
  float* occ = q + C;
  int* ext = reinterpret_cast<int*>(occ + C);
  const int n = blockIdx.x;
  const int blank = C - 1;
  int L = lab_len[n];
  L = L < 0 ? 0 : (L > lmax ? lmax : L);
  int Tn = in_len[n];
  Tn = Tn < 0 ? 0 : (Tn > T ? T : Tn);
  const int S = 2 * L + 1;
  const float loss = sample_loss[n];
  const bool feasible = loss < 1.0e38f;
  const float logp = -loss;
  const float sc = scale * (dyn_scale != nullptr ? *dyn_scale : 1.f);
  for (int s = threadIdx.x; s < S; s += blockDim.x) {
    int l = (s & 1) ? labels[n * lmax + (s >> 1)] : blank;
    ext[s] = l < 0 ? 0 : (l >= C ? C - 1 : l);
  }
  // time steps past the input length (or an ignored sample) get zero gradient
  for (int t = feasible ? Tn : 0; t < T; ++t)
    for (int k = threadIdx.x; k < C; k += blockDim.x)
      dz[(static_cast<long long>(n) * T + t) * lddz + k] = __float2half_rn(0.f);
  if (!feasible) return;
  __syncthreads();
  float* next = b0;
  float* cur = b1;
  const float* an = alpha + static_cast<long long>(n) * T * S_max;
  for (int t = Tn - 1; t >= 0; --t) {
    const long long row = static_cast<long long>(n) * T + t;
    if (threadIdx.x < 32) ctc_log_probs(z + row * ldz, C, eps, lp, q);
    for (int k = threadIdx.x; k < C; k += blockDim.x) occ[k] = 0.f;
    __syncthreads();
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
      float v;
      if (t == Tn - 1) {
        v = s >= S - 2 ? 0.f : kCtcNegInf;
      } else {
        v = next[s];
        if (s + 1 < S) v = log_add(v, next[s + 1]);
        if (s + 2 < S && ext[s] != blank && ext[s] != ext[s + 2]) v = log_add(v, next[s + 2]);
      }
      const float e = lp[ext[s]];
      v = v <= kCtcNegInf ? kCtcNegInf : v + e;
      cur[s] = v;
      const float a = an[static_cast<long long>(t) * S_max + s];
      if (a > kCtcNegInf && v > kCtcNegInf)
        atomicAdd(&occ[ext[s]], __expf(a + v - e - logp));  // alpha*beta / p, normalised
    }
    __syncthreads();
    if (threadIdx.x < 32) {
      // g_k = p_k - occ_k w.r.t. the ctc "logits" u = log(q + eps); chain to z through
      // du_k/dz_j = q_k (delta_kj - q_j) / (q_k + eps)
      float rsum = 0.f;
      for (int k = threadIdx.x; k < C; k += 32) {
        const float g = __expf(lp[k]) - occ[k];
        const float r = g * q[k] / (q[k] + eps);
        lp[k] = r;  // reuse as r_k
        rsum += r;
      }
      rsum = warp_sum(rsum);
      for (int k = threadIdx.x; k < C; k += 32)
        dz[row * lddz + k] = __float2half_rn(sc * (lp[k] - q[k] * rsum));
    }
    __syncthreads();
    float* tmp = next;
    next = cur;
    cur = tmp;
  }
}

// out[0] = mean over the batch of the per-sample losses (ignored samples count as 0)
__global__ void ctc_finalize_kernel(const float* __restrict__ sample_loss, int n,
                                    float* __restrict__ out) {
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += 32) {
    const float v = sample_loss[i];
    s += v < 1.0e38f ? v : 0.f;
  }
  s = warp_sum(s);
  if (threadIdx.x == 0) {
    out[0] = s / n;
    out[1] = 0.f;
    out[2] = 0.f;
    out[3] = 0.f;
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

extern "C" long long kfb_ctc_loss_workspace_floats(int batch, int T, int max_label_len) {
  return static_cast<long long>(batch) * T * (2 * max_label_len + 1) + batch;
}

extern "C" int kfb_ctc_loss(const float* z, long long ldz, int batch, int T, int C,
                            const int* labels, int max_label_len, const int* input_len,
                            const int* label_len, float eps, float scale, const float* dyn_scale,
                            void* dz, long long lddz, float* workspace, long long workspace_floats,
                            float* out, void* stream) {
  KFB_REQUIRE(batch > 0 && T > 0 && C > 1 && max_label_len > 0, "ctc_loss: bad sizes");
  const long long need = kfb_ctc_loss_workspace_floats(batch, T, max_label_len);
  KFB_REQUIRE(workspace != nullptr && workspace_floats >= need,
              "ctc_loss: workspace %lld floats < %lld", workspace_floats, need);
  cudaStream_t st = (cudaStream_t)stream;
  const int S = 2 * max_label_len + 1;
  float* alpha = workspace;
  float* sample = workspace + static_cast<long long>(batch) * T * S;
  const int threads = S >= 1024 ? 1024 : ((S + 31) / 32 * 32 < 64 ? 64 : (S + 31) / 32 * 32);
  const size_t smem_a = (2 * S + C) * sizeof(float) + S * sizeof(int);
  const size_t smem_b = (2 * S + 3 * C) * sizeof(float) + S * sizeof(int);
  KFB_REQUIRE(smem_b <= 200 * 1024, "ctc_loss: label length %d needs too much shared memory",
              max_label_len);
  KFB_CHECK_CUDA(cudaFuncSetAttribute(ctc_alpha_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_a));
  KFB_CHECK_CUDA(cudaFuncSetAttribute(ctc_beta_grad_kernel,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
  ctc_alpha_kernel<<<batch, threads, smem_a, st>>>(z, ldz, T, C, labels, max_label_len, input_len,
                                                   label_len, eps, alpha, sample);
  KFB_CHECK_CUDA(cudaGetLastError());
  if (dz != nullptr) {
    ctc_beta_grad_kernel<<<batch, threads, smem_b, st>>>(z, ldz, T, C, labels, max_label_len,
                                                         input_len, label_len, eps, alpha, sample,
                                                         scale, dyn_scale, (__half*)dz, lddz);
    KFB_CHECK_CUDA(cudaGetLastError());
  }
  ctc_finalize_kernel<<<1, 32, 0, st>>>(sample, batch, out);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
