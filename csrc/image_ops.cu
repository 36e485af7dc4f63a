// Batch image transform on the GPU: decoded uint8 RGB images (uploaded raw from pinned host
// memory) -> crop, resize, mirror, colour jitter, normalise, fp16 NHWC, one launch per batch
// (plus one small launch for the channel means the contrast adjustment needs).
//
// Device counterpart of csrc/host/image_ops.cpp (same kfb_image_op descriptor, same TF1
// arithmetic: tf.image.resize_bilinear / resize_nearest_neighbor with align_corners=False,
// tf.image.crop_and_resize, adjust_brightness / saturation / hue / contrast in the two orders of
// T/preprocessing.py:281-303), so that the host threads only decode JPEG: the reference does this
// stage inside tf.data's C++ runtime (T/preprocessing.py:192-307, 368-399), here it is one
// memory-bound kernel on the device that consumes the batch. Methods 2 (bicubic) and 3 (area)
// stay on the host path.
#include "kfb_common.cuh"
#include "../include/kfb200.h"
#include "../include/kfb200_host.h"

namespace kfb {

__device__ __forceinline__ void img_rgb_to_hsv(const float* p, float* h, float* s, float* v) {
  const float r = p[0], g = p[1], b = p[2];
  const float vv = fmaxf(r, fmaxf(g, b));
  const float range = vv - fminf(r, fminf(g, b));
  *v = vv;
  *s = vv > 0.f ? __fdiv_rn(range, vv) : 0.f;
  float hh = 0.f;
  if (range > 0.f) {
    const float norm = __fdiv_rn(1.f, 6.f * range);
    if (r == vv) {
      hh = __fmul_rn(norm, g - b);
    } else if (g == vv) {
      hh = __fadd_rn(__fmul_rn(norm, b - r), 2.f / 6.f);
    } else {
      hh = __fadd_rn(__fmul_rn(norm, r - g), 4.f / 6.f);
    }
    if (hh < 0.f) hh += 1.f;
  }
  *h = hh;
}

__device__ __forceinline__ void img_hsv_to_rgb(float h, float s, float v, float* p) {
  const float c = __fmul_rn(s, v), m = v - c;
  const float dh = __fmul_rn(h, 6.f);
  const float fmodu = __fadd_rn(dh, -__fmul_rn(2.f, floorf(__fmul_rn(dh, 0.5f))));
  const float x = __fmul_rn(c, 1.f - fabsf(fmodu - 1.f));
  float r = 0.f, g = 0.f, b = 0.f;
  switch (static_cast<int>(dh)) {
    case 0: r = c; g = x; break;
    case 1: r = x; g = c; break;
    case 2: g = c; b = x; break;
    case 3: g = x; b = c; break;
    case 4: r = x; b = c; break;
    default: r = c; b = x; break;
  }
  p[0] = r + m;
  p[1] = g + m;
  p[2] = b + m;
}

__device__ __forceinline__ void img_sat_hue(const kfb_image_op& op, float* v) {
  float h, s, val;
  img_rgb_to_hsv(v, &h, &s, &val);
  s = fminf(1.f, fmaxf(0.f, __fmul_rn(s, op.saturation)));
  h = __fadd_rn(h, op.hue);
  h -= floorf(h);
  img_hsv_to_rgb(h, s, val, v);
}

// pixel (y, x) of the (optionally mirrored) crop
__device__ __forceinline__ const uint8_t* img_px(const uint8_t* src, const kfb_image_op& op, int y,
                                                 int x) {
  const int sx = op.flip ? op.cw - 1 - x : x;
  return src + static_cast<size_t>(op.top + y) * op.src_stride +
         static_cast<size_t>(op.left + sx) * 3;
}

// Resized value of output pixel (y, x): methods 0 (bilinear), 1 (nearest), 4 (crop_and_resize).
// Explicit round-to-nearest multiplies / adds keep the result identical to the host's
// non-contracted float arithmetic.
__device__ __forceinline__ void img_sample(const uint8_t* src, const kfb_image_op& op, int y, int x,
                                           float* o) {
  if (op.method == 4) {
    const int H = op.ih, W = op.iw;
    const float y0 = op.box[0], x0 = op.box[1], y1 = op.box[2], x1 = op.box[3];
    const float hs = op.oh > 1 ? __fdiv_rn(__fmul_rn(y1 - y0, (float)(H - 1)), (float)(op.oh - 1))
                               : 0.f;
    const float ws = op.ow > 1 ? __fdiv_rn(__fmul_rn(x1 - x0, (float)(W - 1)), (float)(op.ow - 1))
                               : 0.f;
    const int xx = op.flip ? op.ow - 1 - x : x;
    const float inx = op.ow > 1 ? __fadd_rn(__fmul_rn(x0, (float)(W - 1)), __fmul_rn((float)xx, ws))
                                : __fmul_rn(__fmul_rn(0.5f, x0 + x1), (float)(W - 1));
    const float iny = op.oh > 1 ? __fadd_rn(__fmul_rn(y0, (float)(H - 1)), __fmul_rn((float)y, hs))
                                : __fmul_rn(__fmul_rn(0.5f, y0 + y1), (float)(H - 1));
    if (iny < 0.f || iny > (float)(H - 1) || inx < 0.f || inx > (float)(W - 1)) {
      o[0] = o[1] = o[2] = 0.f;
      return;
    }
    const int top = (int)floorf(iny), bot = (int)ceilf(iny);
    const int lo = (int)floorf(inx), hi = (int)ceilf(inx);
    const float wy = iny - floorf(iny), wx = inx - floorf(inx);
    const uint8_t* rt = src + static_cast<size_t>(top) * op.src_stride;
    const uint8_t* rb = src + static_cast<size_t>(bot) * op.src_stride;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float a = rt[3 * lo + k], b = rt[3 * hi + k], c = rb[3 * lo + k], d = rb[3 * hi + k];
      const float t = __fadd_rn(a, __fmul_rn(b - a, wx));
      const float u = __fadd_rn(c, __fmul_rn(d - c, wx));
      o[k] = __fadd_rn(t, __fmul_rn(u - t, wy));
    }
    return;
  }
  const float sy = __fdiv_rn((float)op.ch, (float)op.vh), sx = __fdiv_rn((float)op.cw, (float)op.vw);
  if (op.method == 1) {
    const int iy = min((int)floorf(__fmul_rn((float)(op.oy + y), sy)), op.ch - 1);
    const int ix = min((int)floorf(__fmul_rn((float)(op.ox + x), sx)), op.cw - 1);
    const uint8_t* p = img_px(src, op, iy, ix);
    o[0] = p[0];
    o[1] = p[1];
    o[2] = p[2];
    return;
  }
  const float inx = __fmul_rn((float)(op.ox + x), sx), iny = __fmul_rn((float)(op.oy + y), sy);
  const int x0 = min((int)floorf(inx), op.cw - 1), x1 = min(x0 + 1, op.cw - 1);
  const int y0 = min((int)floorf(iny), op.ch - 1), y1 = min(y0 + 1, op.ch - 1);
  const float wx = inx - floorf(inx), wy = iny - floorf(iny);
  const uint8_t *a = img_px(src, op, y0, x0), *b = img_px(src, op, y0, x1);
  const uint8_t *c = img_px(src, op, y1, x0), *d = img_px(src, op, y1, x1);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float fa = a[k], fb = b[k], fc = c[k], fd = d[k];
    const float top = __fadd_rn(fa, __fmul_rn(fb - fa, wx));
    const float bot = __fadd_rn(fc, __fmul_rn(fd - fc, wx));
    o[k] = __fadd_rn(top, __fmul_rn(bot - top, wy));
  }
}

// Channel means at the point where the contrast adjustment is applied (see Jitter::prepare of
// the host code): one block per image, double accumulation, fixed-order tree - deterministic.
__global__ void __launch_bounds__(256)
    image_means_kernel(const unsigned long long* __restrict__ srcs,
                       const kfb_image_op* __restrict__ ops, float* __restrict__ means) {
  const kfb_image_op op = ops[blockIdx.x];
  if (op.distort_order == 0) return;
  const uint8_t* src = reinterpret_cast<const uint8_t*>(srcs[blockIdx.x]);
  const int npix = op.oh * op.ow;
  double m[3] = {0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < npix; i += blockDim.x) {
    float p[3];
    img_sample(src, op, i / op.ow, i % op.ow, p);
    if (op.distort_order == 1) {
      p[0] = __fadd_rn(p[0], op.brightness);
      p[1] = __fadd_rn(p[1], op.brightness);
      p[2] = __fadd_rn(p[2], op.brightness);
      img_sat_hue(op, p);
    }
    m[0] += p[0];
    m[1] += p[1];
    m[2] += p[2];
  }
  __shared__ double sh[3][256];
  for (int k = 0; k < 3; ++k) sh[k][threadIdx.x] = m[k];
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w)
      for (int k = 0; k < 3; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x < 3) {
    const float mean = static_cast<float>(sh[threadIdx.x][0] / static_cast<double>(npix));
    means[blockIdx.x * 3 + threadIdx.x] = mean + (op.distort_order == 1 ? 0.f : op.brightness);
  }
}

__global__ void __launch_bounds__(256)
    image_transform_kernel(const unsigned long long* __restrict__ srcs,
                           const kfb_image_op* __restrict__ ops, const float* __restrict__ means,
                           void* __restrict__ out, long long image_stride) {
  const kfb_image_op op = ops[blockIdx.y];
  const uint8_t* src = reinterpret_cast<const uint8_t*>(srcs[blockIdx.y]);
  const int npix = op.oh * op.ow;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npix; i += gridDim.x * blockDim.x) {
    float v[3];
    img_sample(src, op, i / op.ow, i % op.ow, v);
    if (op.distort_order != 0) {
      const float* mean = means + blockIdx.y * 3;
      if (op.distort_order == 1) {
        v[0] = __fadd_rn(v[0], op.brightness);
        v[1] = __fadd_rn(v[1], op.brightness);
        v[2] = __fadd_rn(v[2], op.brightness);
        img_sat_hue(op, v);
#pragma unroll
        for (int k = 0; k < 3; ++k)
          v[k] = __fadd_rn(mean[k], __fmul_rn(v[k] - mean[k], op.contrast));
      } else {
#pragma unroll
        for (int k = 0; k < 3; ++k)
          v[k] = __fadd_rn(mean[k], __fmul_rn(__fadd_rn(v[k], op.brightness) - mean[k],
                                              op.contrast));
        img_sat_hue(op, v);
      }
      if (op.no_clamp == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] = fminf(fmaxf(v[k], 0.f), 255.f);
      }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
      v[k] = op.per_channel ? __fadd_rn(__fmul_rn(v[k], op.scale3[k]), op.bias3[k])
                            : __fadd_rn(__fmul_rn(v[k], op.scale), op.bias);
    const long long o = blockIdx.y * image_stride + static_cast<long long>(i) * 3;
    if (op.out_f16) {
      __half* d = static_cast<__half*>(out) + o;
      d[0] = __float2half_rn(v[0]);
      d[1] = __float2half_rn(v[1]);
      d[2] = __float2half_rn(v[2]);
    } else {
      float* d = static_cast<float*>(out) + o;
      d[0] = v[0];
      d[1] = v[1];
      d[2] = v[2];
    }
  }
}

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_image_op_bytes_device(void) { return (int)sizeof(kfb_image_op); }

// srcs: n device addresses of the raw images (device array); ops: n descriptors (device array,
// methods 0 / 1 / 4 only - the caller checks); means: 3 * n floats of scratch; out: n images of
// `image_stride` elements each (>= oh * ow * 3), fp16 or fp32 per op.out_f16.
extern "C" int kfb_image_transform_batch(const void* srcs, const void* ops, int n, int max_pixels,
                                         float* means, void* out, long long image_stride,
                                         void* stream) {
  KFB_REQUIRE(srcs != nullptr && ops != nullptr && out != nullptr && means != nullptr && n > 0 &&
                  max_pixels > 0,
              "image_transform_batch: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  image_means_kernel<<<n, 256, 0, st>>>((const unsigned long long*)srcs,
                                        (const kfb_image_op*)ops, means);
  int bx = (max_pixels + 255) / 256;
  if (bx > 64) bx = 64;
  image_transform_kernel<<<dim3(bx, n), 256, 0, st>>>((const unsigned long long*)srcs,
                                                      (const kfb_image_op*)ops, means, out,
                                                      image_stride);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
