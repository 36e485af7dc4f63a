/* C ABI of the host-side input runtime, _lib/libkfb200_host.so (plain C++17, no CUDA; sources in
 * csrc/host/). These are the pieces of TensorFlow's C++ input runtime that the reference drives
 * through RecordInput / tf.data / tf.image (T/preprocessing.py:27-72, 130-307, 358-399, 505-548,
 * 601-617). Python bindings: kf_benchmarks_b200/data/native_reader.py and
 * native_image.py. All functions are thread-safe unless noted; errors are reported through the
 * return value plus a thread-local message. */
#ifndef KFB200_HOST_H_
#define KFB200_HOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ */
/* TFRecord streaming (csrc/host/record_io.cpp)                        */
/* ------------------------------------------------------------------ */

/* Message of the last failed call of this thread. */
const char* kfb_host_last_error(void);

/* CRC32C (Castagnoli; hardware instruction when built with SSE4.2) and TFRecord's masked form. */
uint32_t kfb_crc32c(const void* data, uint64_t len);
uint32_t kfb_masked_crc32c(const void* data, uint64_t len);

/* Opens a reader over `npaths` TFRecord files read by `num_threads` background threads (file i
 * belongs to thread i % num_threads - a parallel interleave with that cycle length,
 * T/preprocessing.py:516-519). `capacity` bounds the records queued ahead, `check_crc` validates
 * both CRCs of every record, `repeat` cycles forever, `shuffle_files` reshuffles each thread's
 * file list per pass with `seed`. Records come in whatever order the threads deliver them
 * (datasets_sloppy_parallel_interleave). Returns a handle, or NULL (kfb_host_last_error). */
void* kfb_rr_open(const char* const* paths, int npaths, int num_threads, int capacity,
                  int check_crc, int repeat, int shuffle_files, uint64_t seed);
/* As kfb_rr_open; `deterministic` != 0 takes one record from each reader in turn (tf.data
 * interleave with block length 1), so the order depends on the file list only. */
void* kfb_rr_open2(const char* const* paths, int npaths, int num_threads, int capacity,
                   int check_crc, int repeat, int shuffle_files, uint64_t seed,
                   int deterministic);
/* 1: *data / *len describe the next record (valid until the next call on this handle);
 * 0: end of data; -1: error (kfb_host_last_error). One consumer thread per handle. */
int kfb_rr_next(void* handle, const uint8_t** data, uint64_t* len);
/* Stops the reader threads (also when they are blocked on a full queue) and frees the handle. */
void kfb_rr_close(void* handle);

/* ------------------------------------------------------------------ */
/* tf.Example field scanner (csrc/host/example_scan.cpp)               */
/* ------------------------------------------------------------------ */

/* First element of the BytesList feature `key` of a serialized tf.Example, in place.
 * 1 found, 0 missing / not bytes. Never reads outside [data, data + len). */
int kfb_example_get_bytes(const uint8_t* data, uint64_t len, const char* key, const uint8_t** out,
                          uint64_t* out_len);
/* Values of the FloatList / Int64List feature `key` as doubles. Returns the number of values the
 * feature holds (at most `cap` are written), or -1 if the key is missing. */
int kfb_example_get_numbers(const uint8_t* data, uint64_t len, const char* key, double* out,
                            int cap);

/* ------------------------------------------------------------------ */
/* Per-image transform (csrc/host/image_ops.cpp)                       */
/* ------------------------------------------------------------------ */

/* One decoded uint8 RGB image -> crop, resize, mirror, colour jitter, normalise, cast, written to
 * the caller's buffer. Resize methods 0-3 use the arithmetic of TF1 tf.image.resize_* with
 * align_corners=False (T/preprocessing.py:130-259); method 4 is tf.image.crop_and_resize of a
 * normalised window (SSD, T/ssd_dataloader.py:218-223). The jitter is tf.image.adjust_brightness
 * / adjust_saturation / adjust_hue / adjust_contrast on the 0..255 scale in the two orders of
 * T/preprocessing.py:281-303. */
typedef struct {
  int ih, iw, src_stride;  /* source: uint8 RGB, ih x iw, row stride in bytes */
  int top, left, ch, cw;   /* crop rectangle in source pixels (methods 0-3) */
  int vh, vw;              /* size the crop is resized to */
  int oy, ox, oh, ow;      /* window of that resized image which is produced */
  int method;              /* 0 bilinear, 1 nearest, 2 bicubic, 3 area, 4 crop_and_resize of `box` */
  int flip;                /* mirror horizontally (the crop before resizing; the output, method 4) */
  int distort_order;       /* 0 none, 1 brightness - saturation/hue - contrast,
                              2 brightness - contrast - saturation/hue */
  float brightness;        /* added */
  float saturation;        /* HSV saturation factor, result clamped to [0, 1] */
  float hue;               /* HSV hue shift in turns */
  float contrast;          /* factor around the per-channel image mean */
  float scale, bias;       /* out = value * scale + bias */
  int out_f16;             /* 0: float32 output, 1: IEEE binary16 output */
  float box[4];            /* method 4: normalised (y0, x0, y1, x1); zero outside the image */
  int per_channel;         /* 1: out = value * scale3[k] + bias3[k] */
  float scale3[3], bias3[3];
  int no_clamp;            /* 1: no clamp to [0, 255] after the jitter */
} kfb_image_op;

/* 0 on success; -1 on invalid arguments (kfb_image_last_error). `out` holds oh * ow * 3 elements. */
int kfb_image_transform(const uint8_t* src, const kfb_image_op* op, void* out);
const char* kfb_image_last_error(void);
/* sizeof(kfb_image_op) as this library was built (bindings check their mirror against it). */
int kfb_image_op_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* KFB200_HOST_H_ */
