// Batched JPEG decode on the device through nvJPEG, feeding kfb_image_transform_batch.
//
// The reference decodes inside tf.data's C++ runtime (tf.image.decode_jpeg,
// T/preprocessing.py:95-104) on host threads; here the host threads only cut the JPEG bytes out of
// the tf.Example and read the frame header, the bitstreams of a batch go to nvJPEG (hardware
// engine where the GPU has one, GPU-assisted Huffman + IDCT kernels otherwise) and the decoded
// interleaved RGB images stay on the device for the transform kernel: no decoded pixel crosses
// PCIe and no host core spends 2 ms per image in libjpeg.
//
// nvJPEG is loaded with dlopen so that the library has no link-time dependency on it:
// kfb_jpeg_available() says whether it could be loaded.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvjpeg.h>
#include <stdint.h>

#include <mutex>
#include <vector>

#include "kfb_common.cuh"
#include "../include/kfb200.h"

namespace kfb {

struct NvjpegApi {
  void* lib = nullptr;
  decltype(&nvjpegCreateEx) create_ex = nullptr;
  decltype(&nvjpegCreateSimple) create_simple = nullptr;
  decltype(&nvjpegDestroy) destroy = nullptr;
  decltype(&nvjpegJpegStateCreate) state_create = nullptr;
  decltype(&nvjpegJpegStateDestroy) state_destroy = nullptr;
  decltype(&nvjpegGetImageInfo) image_info = nullptr;
  decltype(&nvjpegDecodeBatchedInitialize) batched_init = nullptr;
  decltype(&nvjpegDecodeBatched) batched = nullptr;
  decltype(&nvjpegDecode) decode = nullptr;
  bool ok = false;
};

static NvjpegApi* nvjpeg_api() {
  static NvjpegApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnvjpeg.so.12", "libnvjpeg.so", "/usr/local/cuda/lib64/libnvjpeg.so.12"};
    for (const char* n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
      if (api.lib != nullptr) break;
    }
    if (api.lib == nullptr) return;
#define KFB_SYM(field, name) \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name))
    KFB_SYM(create_ex, "nvjpegCreateEx");
    KFB_SYM(create_simple, "nvjpegCreateSimple");
    KFB_SYM(destroy, "nvjpegDestroy");
    KFB_SYM(state_create, "nvjpegJpegStateCreate");
    KFB_SYM(state_destroy, "nvjpegJpegStateDestroy");
    KFB_SYM(image_info, "nvjpegGetImageInfo");
    KFB_SYM(batched_init, "nvjpegDecodeBatchedInitialize");
    KFB_SYM(batched, "nvjpegDecodeBatched");
    KFB_SYM(decode, "nvjpegDecode");
#undef KFB_SYM
    api.ok = api.create_ex && api.create_simple && api.destroy && api.state_create &&
             api.state_destroy && api.image_info && api.batched_init && api.batched && api.decode;
  });
  return &api;
}

struct JpegDecoder {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;
  int backend = 0;        // nvjpegBackend_t actually created
  int batch_init = 0;     // batch size of the last nvjpegDecodeBatchedInitialize
  int cpu_threads = 1;
  std::vector<nvjpegImage_t> dst;
};

}  // namespace kfb

using namespace kfb;

extern "C" int kfb_jpeg_available(void) { return nvjpeg_api()->ok ? 1 : 0; }

// backend: 0 = hardware engine if present, else GPU-assisted; 1 = hardware only; 2 = GPU-assisted
// Huffman (NVJPEG_BACKEND_GPU_HYBRID); 3 = nvJPEG's default.
extern "C" int kfb_jpeg_create(int backend, int cpu_threads, void** out) {
  NvjpegApi* api = nvjpeg_api();
  KFB_REQUIRE(api->ok, "jpeg_create: libnvjpeg could not be loaded");
  KFB_REQUIRE(out != nullptr, "jpeg_create: null output");
  JpegDecoder* d = new JpegDecoder();
  d->cpu_threads = cpu_threads > 0 ? cpu_threads : 1;
  const nvjpegBackend_t order_auto[] = {NVJPEG_BACKEND_HARDWARE, NVJPEG_BACKEND_GPU_HYBRID,
                                        NVJPEG_BACKEND_DEFAULT};
  const nvjpegBackend_t order_hw[] = {NVJPEG_BACKEND_HARDWARE};
  const nvjpegBackend_t order_gpu[] = {NVJPEG_BACKEND_GPU_HYBRID, NVJPEG_BACKEND_DEFAULT};
  const nvjpegBackend_t order_def[] = {NVJPEG_BACKEND_DEFAULT};
  const nvjpegBackend_t* order = order_auto;
  int n = 3;
  if (backend == 1) { order = order_hw; n = 1; }
  if (backend == 2) { order = order_gpu; n = 2; }
  if (backend == 3) { order = order_def; n = 1; }
  nvjpegStatus_t st = NVJPEG_STATUS_NOT_INITIALIZED;
  for (int i = 0; i < n; ++i) {
    st = api->create_ex(order[i], nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->handle);
    if (st == NVJPEG_STATUS_SUCCESS) {
      d->backend = (int)order[i];
      break;
    }
    d->handle = nullptr;
  }
  if (d->handle == nullptr) {
    delete d;
    set_error("jpeg_create: nvjpegCreateEx failed (status %d)", (int)st);
    return 1;
  }
  st = api->state_create(d->handle, &d->state);
  if (st != NVJPEG_STATUS_SUCCESS) {
    api->destroy(d->handle);
    delete d;
    set_error("jpeg_create: nvjpegJpegStateCreate failed (status %d)", (int)st);
    return 1;
  }
  *out = d;
  return 0;
}

extern "C" int kfb_jpeg_backend(void* h) {
  return h != nullptr ? static_cast<JpegDecoder*>(h)->backend : -1;
}

extern "C" int kfb_jpeg_destroy(void* h) {
  if (h == nullptr) return 0;
  NvjpegApi* api = nvjpeg_api();
  JpegDecoder* d = static_cast<JpegDecoder*>(h);
  if (d->state != nullptr) api->state_destroy(d->state);
  if (d->handle != nullptr) api->destroy(d->handle);
  delete d;
  return 0;
}

// Frame header of one bitstream (host only): width / height of the luma plane, number of
// components (1 = grayscale, 3 = colour; anything else is not decodable here).
extern "C" int kfb_jpeg_image_info(void* h, const unsigned char* data, unsigned long long len,
                                   int* width, int* height, int* components) {
  KFB_REQUIRE(h != nullptr && data != nullptr && len > 0, "jpeg_image_info: bad arguments");
  JpegDecoder* d = static_cast<JpegDecoder*>(h);
  int nc = 0, ws[NVJPEG_MAX_COMPONENT] = {0}, hs[NVJPEG_MAX_COMPONENT] = {0};
  nvjpegChromaSubsampling_t ss;
  const nvjpegStatus_t st =
      nvjpeg_api()->image_info(d->handle, data, (size_t)len, &nc, &ss, ws, hs);
  KFB_REQUIRE(st == NVJPEG_STATUS_SUCCESS, "jpeg_image_info: nvjpegGetImageInfo status %d", (int)st);
  if (width != nullptr) *width = ws[0];
  if (height != nullptr) *height = hs[0];
  if (components != nullptr) *components = nc;
  return 0;
}

// Decodes n bitstreams (host memory) into n interleaved RGB device buffers dst[i] with row
// pitch pitch[i] bytes (>= 3 * width). Asynchronous on `stream` as far as the backend allows
// (Huffman decoding of the GPU-hybrid backends runs on up to cpu_threads host threads inside
// the call); the input buffers must stay valid until the stream has passed this point.
extern "C" int kfb_jpeg_decode_batch(void* h, const unsigned char* const* data,
                                     const unsigned long long* lens, int n,
                                     unsigned char* const* dst, const int* pitch, void* stream) {
  KFB_REQUIRE(h != nullptr && n > 0 && data != nullptr && lens != nullptr && dst != nullptr &&
                  pitch != nullptr,
              "jpeg_decode_batch: bad arguments");
  NvjpegApi* api = nvjpeg_api();
  JpegDecoder* d = static_cast<JpegDecoder*>(h);
  nvjpegStatus_t st;
  if (d->batch_init != n) {
    st = api->batched_init(d->handle, d->state, n, d->cpu_threads, NVJPEG_OUTPUT_RGBI);
    KFB_REQUIRE(st == NVJPEG_STATUS_SUCCESS,
                "jpeg_decode_batch: nvjpegDecodeBatchedInitialize status %d", (int)st);
    d->batch_init = n;
  }
  d->dst.assign(n, nvjpegImage_t());
  std::vector<size_t> lengths(n);
  for (int i = 0; i < n; ++i) {
    for (int c = 0; c < NVJPEG_MAX_COMPONENT; ++c) {
      d->dst[i].channel[c] = nullptr;
      d->dst[i].pitch[c] = 0;
    }
    d->dst[i].channel[0] = dst[i];
    d->dst[i].pitch[0] = (size_t)pitch[i];
    lengths[i] = (size_t)lens[i];
  }
  st = api->batched(d->handle, d->state, data, lengths.data(), d->dst.data(),
                    (cudaStream_t)stream);
  if (st != NVJPEG_STATUS_SUCCESS) {
    d->batch_init = 0;  // a failed batch must be re-initialised
    set_error("jpeg_decode_batch: nvjpegDecodeBatched status %d (backend %d)", (int)st,
              d->backend);
    return 1;
  }
  return 0;
}
