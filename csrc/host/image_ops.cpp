// Host-side image transforms of the real-data input pipeline (plain C++17, no CUDA): crop +
// resize + horizontal flip + colour jitter + normalisation + cast, one call per decoded image,
// written straight into the caller's (pinned) batch buffer. Counterpart of the TensorFlow image
// kernels the reference's preprocessing graph runs per example on the tf.data thread pool
// (crop + resize_{bilinear,nearest_neighbor,bicubic,area}, flip, distort_color, normalise:
// T/preprocessing.py:130-307, 653-737).
//
// The resize kernels follow the semantics the reference gets from TF1-style
// tf.image.resize_* with align_corners=False: source coordinate = destination index *
// (in_size / out_size), no half-pixel shift, no antialiasing.
//
// The call releases the GIL (ctypes), so a Python thread pool scales across cores: JPEG decode
// (Pillow, also GIL-free) and this function are the two per-image costs.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#if defined(__F16C__)
#include <immintrin.h>
#endif

#include "../../include/kfb200_host.h"

namespace {

thread_local std::vector<float> g_plane;   // oh * ow * 3 floats
thread_local char g_err[256];

inline uint16_t to_half(float f) {
#if defined(__F16C__)
  return static_cast<uint16_t>(_cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT));
#else
  uint32_t x;
  std::memcpy(&x, &f, 4);
  const uint32_t sign = (x >> 16) & 0x8000u;
  const int32_t exp = static_cast<int32_t>((x >> 23) & 0xFF) - 127 + 15;
  uint32_t man = x & 0x7FFFFFu;
  if (((x >> 23) & 0xFF) == 0xFF) return static_cast<uint16_t>(sign | 0x7C00u | (man ? 0x200u : 0));
  if (exp >= 31) return static_cast<uint16_t>(sign | 0x7C00u);
  if (exp <= 0) {
    if (exp < -10) return static_cast<uint16_t>(sign);
    man |= 0x800000u;
    const int shift = 14 - exp;
    uint32_t h = man >> shift;
    const uint32_t rem = man & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (h & 1))) ++h;
    return static_cast<uint16_t>(sign | h);
  }
  uint32_t h = (static_cast<uint32_t>(exp) << 10) | (man >> 13);
  const uint32_t rem = man & 0x1FFFu;
  if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) ++h;
  return static_cast<uint16_t>(sign | h);
#endif
}

struct Tap2 {
  int i0, i1;
  float w1;  // weight of i1; i0 gets 1 - w1
};

// Keys cubic convolution, A = -0.75, fraction quantised to 1/1024 like TF's coefficient table
inline void cubic_weights(float frac, float w[4]) {
  const float a = -0.75f;
  const float x = std::lrint(frac * 1024.f) / 1024.f;
  auto near = [a](float t) { return ((a + 2.f) * t - (a + 3.f)) * t * t + 1.f; };
  auto far = [a](float t) { return ((a * t - 5.f * a) * t + 8.f * a) * t - 4.f * a; };
  w[0] = far(x + 1.f);
  w[1] = near(x);
  w[2] = near(1.f - x);
  w[3] = far(2.f - x);
}

// pixel (y, x) of the (optionally mirrored) crop
inline const uint8_t* px(const uint8_t* src, const kfb_image_op& op, int y, int x) {
  const int sx = op.flip ? op.cw - 1 - x : x;
  return src + static_cast<size_t>(op.top + y) * op.src_stride +
         static_cast<size_t>(op.left + sx) * 3;
}

// tf.image.crop_and_resize, bilinear, extrapolation value 0 (T/ssd_dataloader.py:218-223)
void crop_and_resize(const uint8_t* src, const kfb_image_op& op, float* out) {
  const int oh = op.oh, ow = op.ow, H = op.ih, W = op.iw;
  const float y0 = op.box[0], x0 = op.box[1], y1 = op.box[2], x1 = op.box[3];
  const float hs = oh > 1 ? (y1 - y0) * (H - 1) / (oh - 1) : 0.f;
  const float ws = ow > 1 ? (x1 - x0) * (W - 1) / (ow - 1) : 0.f;
  struct XT {
    int lo, hi;
    float w;
    bool ok;
  };
  std::vector<XT> xt(ow);
  for (int x = 0; x < ow; ++x) {
    const int xx = op.flip ? ow - 1 - x : x;
    const float in = ow > 1 ? x0 * (W - 1) + xx * ws : 0.5f * (x0 + x1) * (W - 1);
    xt[x].ok = in >= 0.f && in <= static_cast<float>(W - 1);
    xt[x].lo = static_cast<int>(std::floor(in));
    xt[x].hi = static_cast<int>(std::ceil(in));
    xt[x].w = in - std::floor(in);
  }
  for (int y = 0; y < oh; ++y) {
    float* o = out + static_cast<size_t>(y) * ow * 3;
    const float in = oh > 1 ? y0 * (H - 1) + y * hs : 0.5f * (y0 + y1) * (H - 1);
    if (in < 0.f || in > static_cast<float>(H - 1)) {
      std::memset(o, 0, sizeof(float) * ow * 3);
      continue;
    }
    const int top = static_cast<int>(std::floor(in)), bot = static_cast<int>(std::ceil(in));
    const float wy = in - std::floor(in);
    const uint8_t* rt = src + static_cast<size_t>(top) * op.src_stride;
    const uint8_t* rb = src + static_cast<size_t>(bot) * op.src_stride;
    for (int x = 0; x < ow; ++x) {
      if (!xt[x].ok) {
        o[3 * x] = o[3 * x + 1] = o[3 * x + 2] = 0.f;
        continue;
      }
      const uint8_t *a = rt + 3 * xt[x].lo, *b = rt + 3 * xt[x].hi;
      const uint8_t *c = rb + 3 * xt[x].lo, *d = rb + 3 * xt[x].hi;
      for (int k = 0; k < 3; ++k) {
        const float t = a[k] + (b[k] - a[k]) * xt[x].w;
        const float u = c[k] + (d[k] - c[k]) * xt[x].w;
        o[3 * x + k] = t + (u - t) * wy;
      }
    }
  }
}

void resize_window(const uint8_t* src, const kfb_image_op& op, float* out) {
  const float sy = static_cast<float>(op.ch) / op.vh, sx = static_cast<float>(op.cw) / op.vw;
  const int oh = op.oh, ow = op.ow;
  if (op.method == 1) {  // nearest
    std::vector<int> xs(ow);
    for (int x = 0; x < ow; ++x)
      xs[x] = std::min(static_cast<int>(std::floor((op.ox + x) * sx)), op.cw - 1);
    for (int y = 0; y < oh; ++y) {
      const int iy = std::min(static_cast<int>(std::floor((op.oy + y) * sy)), op.ch - 1);
      float* o = out + static_cast<size_t>(y) * ow * 3;
      for (int x = 0; x < ow; ++x) {
        const uint8_t* p = px(src, op, iy, xs[x]);
        o[3 * x] = p[0];
        o[3 * x + 1] = p[1];
        o[3 * x + 2] = p[2];
      }
    }
    return;
  }
  if (op.method == 0) {  // bilinear
    std::vector<Tap2> xt(ow);
    for (int x = 0; x < ow; ++x) {
      const float in = (op.ox + x) * sx;
      const int lo = std::min(static_cast<int>(std::floor(in)), op.cw - 1);
      xt[x] = {lo, std::min(lo + 1, op.cw - 1), in - std::floor(in)};
    }
    for (int y = 0; y < oh; ++y) {
      const float in = (op.oy + y) * sy;
      const int y0 = std::min(static_cast<int>(std::floor(in)), op.ch - 1);
      const int y1 = std::min(y0 + 1, op.ch - 1);
      const float wy = in - std::floor(in);
      float* o = out + static_cast<size_t>(y) * ow * 3;
      for (int x = 0; x < ow; ++x) {
        const Tap2& t = xt[x];
        const uint8_t *a = px(src, op, y0, t.i0), *b = px(src, op, y0, t.i1);
        const uint8_t *c = px(src, op, y1, t.i0), *d = px(src, op, y1, t.i1);
        for (int k = 0; k < 3; ++k) {
          const float top = a[k] + (b[k] - a[k]) * t.w1;
          const float bot = c[k] + (d[k] - c[k]) * t.w1;
          o[3 * x + k] = top + (bot - top) * wy;
        }
      }
    }
    return;
  }
  if (op.method == 2) {  // bicubic
    struct Tap4 {
      int i[4];
      float w[4];
    };
    std::vector<Tap4> xt(ow);
    for (int x = 0; x < ow; ++x) {
      const float in = (op.ox + x) * sx;
      const int base = static_cast<int>(std::floor(in));
      cubic_weights(in - base, xt[x].w);
      for (int k = 0; k < 4; ++k) xt[x].i[k] = std::min(std::max(base - 1 + k, 0), op.cw - 1);
    }
    for (int y = 0; y < oh; ++y) {
      const float in = (op.oy + y) * sy;
      const int base = static_cast<int>(std::floor(in));
      float wy[4];
      int iy[4];
      cubic_weights(in - base, wy);
      for (int k = 0; k < 4; ++k) iy[k] = std::min(std::max(base - 1 + k, 0), op.ch - 1);
      float* o = out + static_cast<size_t>(y) * ow * 3;
      for (int x = 0; x < ow; ++x) {
        float acc[3] = {0.f, 0.f, 0.f};
        for (int r = 0; r < 4; ++r) {
          float row[3] = {0.f, 0.f, 0.f};
          for (int c = 0; c < 4; ++c) {
            const uint8_t* p = px(src, op, iy[r], xt[x].i[c]);
            row[0] += p[0] * xt[x].w[c];
            row[1] += p[1] * xt[x].w[c];
            row[2] += p[2] * xt[x].w[c];
          }
          acc[0] += row[0] * wy[r];
          acc[1] += row[1] * wy[r];
          acc[2] += row[2] * wy[r];
        }
        o[3 * x] = acc[0];
        o[3 * x + 1] = acc[1];
        o[3 * x + 2] = acc[2];
      }
    }
    return;
  }
  // area: average of the source box [o*s, (o+1)*s) with fractional edge weights
  struct Span {
    int begin, end;            // source indices [begin, end)
    std::vector<float> w;
  };
  auto spans = [](int off, int n, float s) {
    std::vector<Span> v(n);
    for (int o = 0; o < n; ++o) {
      const float a = (off + o) * s, b = (off + o + 1) * s;
      Span& sp = v[o];
      sp.begin = static_cast<int>(std::floor(a));
      sp.end = static_cast<int>(std::ceil(b));
      for (int i = sp.begin; i < sp.end; ++i)
        sp.w.push_back(i < a ? (i + 1 > b ? b - a : i + 1 - a) : (i + 1 > b ? b - i : 1.f));
    }
    return v;
  };
  const std::vector<Span> ys = spans(op.oy, oh, sy), xs = spans(op.ox, ow, sx);
  const float norm = 1.f / (sy * sx);
  for (int y = 0; y < oh; ++y) {
    float* o = out + static_cast<size_t>(y) * ow * 3;
    for (int x = 0; x < ow; ++x) {
      float acc[3] = {0.f, 0.f, 0.f};
      for (int i = ys[y].begin; i < ys[y].end; ++i) {
        const int iy = std::min(i, op.ch - 1);
        const float wy = ys[y].w[i - ys[y].begin];
        for (int j = xs[x].begin; j < xs[x].end; ++j) {
          const uint8_t* p = px(src, op, iy, std::min(j, op.cw - 1));
          const float w = wy * xs[x].w[j - xs[x].begin];
          acc[0] += p[0] * w;
          acc[1] += p[1] * w;
          acc[2] += p[2] * w;
        }
      }
      o[3 * x] = acc[0] * norm;
      o[3 * x + 1] = acc[1] * norm;
      o[3 * x + 2] = acc[2] * norm;
    }
  }
}

// Colour jitter with the arithmetic of tf.image.adjust_{brightness,saturation,hue,contrast}
// (values on the 0..255 scale here, 0..1 there): brightness adds, saturation and hue go through
// HSV (one round trip for both: they touch different components), contrast scales around the
// per-channel image mean taken at the point where it is applied; one clamp at the very end.
inline void rgb_to_hsv(const float* p, float* h, float* s, float* v) {
  const float r = p[0], g = p[1], b = p[2];
  const float vv = std::max(r, std::max(g, b));
  const float range = vv - std::min(r, std::min(g, b));
  *v = vv;
  *s = vv > 0.f ? range / vv : 0.f;
  float hh = 0.f;
  if (range > 0.f) {
    const float norm = 1.f / (6.f * range);
    if (r == vv) {
      hh = norm * (g - b);
    } else if (g == vv) {
      hh = norm * (b - r) + 2.f / 6.f;
    } else {
      hh = norm * (r - g) + 4.f / 6.f;
    }
    if (hh < 0.f) hh += 1.f;
  }
  *h = hh;
}

inline void hsv_to_rgb(float h, float s, float v, float* p) {
  const float c = s * v, m = v - c;
  const float dh = h * 6.f;
  const float fmodu = dh - 2.f * std::floor(dh * 0.5f);  // dh mod 2
  const float x = c * (1.f - std::fabs(fmodu - 1.f));
  float r = 0.f, g = 0.f, b = 0.f;
  switch (static_cast<int>(dh)) {
    case 0: r = c; g = x; break;
    case 1: r = x; g = c; break;
    case 2: g = c; b = x; break;
    case 3: g = x; b = c; break;
    case 4: r = x; b = c; break;
    default: r = c; b = x; break;  // 5 (and h == 1 after rounding)
  }
  p[0] = r + m;
  p[1] = g + m;
  p[2] = b + m;
}

struct Jitter {
  bool on = false, clamp = true;
  int order = 0;
  float b = 0.f, fs = 1.f, dh = 0.f, fc = 1.f;
  float mean[3] = {0.f, 0.f, 0.f};  // channel means at the point where contrast is applied

  inline void sat_hue(float* v) const {
    float h, s, val;
    rgb_to_hsv(v, &h, &s, &val);
    s = std::min(1.f, std::max(0.f, s * fs));
    h += dh;
    h -= std::floor(h);  // wrap into [0, 1)
    hsv_to_rgb(h, s, val, v);
  }

  // Order 1 applies brightness + saturation/hue in place here (contrast needs the channel means
  // of that result); order 2 only needs the means of the resized image.
  void prepare(const kfb_image_op& op, float* img, size_t npix) {
    on = op.distort_order != 0;
    if (!on) return;
    clamp = op.no_clamp == 0;
    order = op.distort_order;
    b = op.brightness;
    fs = op.saturation;
    dh = op.hue;
    fc = op.contrast;
    double m[3] = {0, 0, 0};
    for (size_t i = 0; i < npix; ++i) {
      float* p = img + 3 * i;
      if (order == 1) {
        p[0] += b;
        p[1] += b;
        p[2] += b;
        sat_hue(p);
      }
      m[0] += p[0];
      m[1] += p[1];
      m[2] += p[2];
    }
    for (int k = 0; k < 3; ++k)
      mean[k] = static_cast<float>(m[k] / static_cast<double>(npix)) + (order == 1 ? 0.f : b);
  }

  inline void apply(const float* p, float v[3]) const {
    v[0] = p[0];
    v[1] = p[1];
    v[2] = p[2];
    if (!on) return;
    if (order == 1) {
      for (int k = 0; k < 3; ++k) v[k] = mean[k] + (v[k] - mean[k]) * fc;
    } else {
      for (int k = 0; k < 3; ++k) v[k] = mean[k] + (v[k] + b - mean[k]) * fc;
      sat_hue(v);
    }
    if (clamp)
      for (int k = 0; k < 3; ++k) v[k] = std::min(std::max(v[k], 0.f), 255.f);
  }
};

}  // namespace

extern "C" {

const char* kfb_image_last_error() { return g_err; }

int kfb_image_op_bytes() { return static_cast<int>(sizeof(kfb_image_op)); }

// 0 on success; -1 on invalid arguments (kfb_image_last_error()).
int kfb_image_transform(const uint8_t* src, const kfb_image_op* opp, void* out) {
  if (src == nullptr || opp == nullptr || out == nullptr) {
    std::snprintf(g_err, sizeof(g_err), "image_transform: null argument");
    return -1;
  }
  const kfb_image_op& op = *opp;
  const bool car = op.method == 4;
  if (car) {
    if (op.ih <= 0 || op.iw <= 0 || op.src_stride < op.iw * 3 || op.oh <= 0 || op.ow <= 0) {
      std::snprintf(g_err, sizeof(g_err), "image_transform: bad geometry (image %dx%d, output %dx%d)",
                    op.ih, op.iw, op.oh, op.ow);
      return -1;
    }
  } else if (op.ih <= 0 || op.iw <= 0 || op.src_stride < op.iw * 3 || op.ch <= 0 || op.cw <= 0 ||
      op.top < 0 || op.left < 0 || op.top + op.ch > op.ih || op.left + op.cw > op.iw ||
      op.vh <= 0 || op.vw <= 0 || op.oh <= 0 || op.ow <= 0 || op.oy < 0 || op.ox < 0 ||
      op.oy + op.oh > op.vh || op.ox + op.ow > op.vw || op.method < 0 || op.method > 3) {
    std::snprintf(g_err, sizeof(g_err),
                  "image_transform: bad geometry (image %dx%d, crop %d,%d %dx%d, resize %dx%d, "
                  "window %d,%d %dx%d, method %d)",
                  op.ih, op.iw, op.top, op.left, op.ch, op.cw, op.vh, op.vw, op.oy, op.ox, op.oh,
                  op.ow, op.method);
    return -1;
  }
  const size_t npix = static_cast<size_t>(op.oh) * op.ow;
  if (g_plane.size() < npix * 3) g_plane.resize(npix * 3);
  float* plane = g_plane.data();
  if (car) {
    crop_and_resize(src, op, plane);
  } else {
    resize_window(src, op, plane);
  }
  Jitter jitter;
  jitter.prepare(op, plane, npix);
  const int ow = op.ow;
  for (int y = 0; y < op.oh; ++y) {
    const float* row = plane + static_cast<size_t>(y) * ow * 3;
    for (int x = 0; x < ow; ++x) {
      float p[3];
      jitter.apply(row + 3 * x, p);
      const size_t o = (static_cast<size_t>(y) * ow + x) * 3;
      if (op.per_channel) {
        for (int k = 0; k < 3; ++k) p[k] = p[k] * op.scale3[k] + op.bias3[k];
      } else {
        for (int k = 0; k < 3; ++k) p[k] = p[k] * op.scale + op.bias;
      }
      if (op.out_f16) {
        uint16_t* d = static_cast<uint16_t*>(out) + o;
        for (int k = 0; k < 3; ++k) d[k] = to_half(p[k]);
      } else {
        float* d = static_cast<float*>(out) + o;
        for (int k = 0; k < 3; ++k) d[k] = p[k];
      }
    }
  }
  return 0;
}

}  // extern "C"
