/* kf-benchmarks B200 native library — C ABI.
 *
 * All entry points return 0 on success; on failure they return non-zero and
 * kfb_last_error() describes the problem. Pointers are device pointers unless
 * stated otherwise; `stream` is a cudaStream_t passed as void*. No entry point
 * synchronises the device. Tensors are dense row-major; activations are NHWC
 * fp16, filters OHWI ([Cout][R][S][Cin]) fp16.
 *
 * Citations name the reference call site (tf_cnn_benchmarks, T/ =
 * scripts/tf_cnn_benchmarks/) whose behaviour the entry point provides.
 */
#ifndef KFB200_H_
#define KFB200_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* kfb_last_error(void);

/* ------------------------------------------------------------------ */
/* tcgen05 GEMM / implicit-GEMM convolution (csrc/conv_gemm_tcgen05.cu) */
/* ------------------------------------------------------------------ */

typedef struct {
  int n, h, w, c;          /* input  NHWC */
  int k, r, s;             /* filter Cout, R, S */
  int pad_h_lo, pad_h_hi;  /* explicit padding (SAME / VALID / SAME_RESNET are resolved by the caller) */
  int pad_w_lo, pad_w_hi;
  int stride_h, stride_w;
  int p, q;                /* output spatial size */
} kfb_conv_desc;

/* C[M,N] = A[M,K] * B[N,K]^T (+bias[N]) (ReLU). fp16 in, fp16 or fp32 out.
 * T/convnet_builder_legacy.py:318 (xw_plus_b) and 1x1 convolutions. */
int kfb_gemm_tn(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                long long ldb, long long ldc, const float* bias, int relu, int out_fp32,
                void* stream);

/* C[M,N] (fp32) (+)= scale * A[K,M]^T * B[K,N]; K is split across CTAs into `workspace`
 * (fp32 slices of M*N) and reduced in a fixed order. FC weight gradient. */
int kfb_gemm_nt(const void* A, const void* B, float* C, int M, int N, int K, long long lda,
                long long ldb, long long ldc, float* workspace, long long workspace_bytes,
                int accumulate, float scale, void* stream);

/* y = conv(x, w) (+bias) (ReLU). T/convnet_builder_legacy.py:123-141. */
int kfb_conv_fprop(const kfb_conv_desc* d, const void* x, const void* w, void* y, long long ldy,
                   const float* bias, int relu, void* stream);

/* dx = conv^T(dy, w) (autodiff of the above, T/benchmark_cnn.py:3104). For stride > 1 only
 * 1x1 un-padded filters are supported and dx must be zero-filled by the caller. */
int kfb_conv_dgrad(const kfb_conv_desc* d, const void* dy, const void* w, void* dx, void* stream);

/* dw (fp32, OHWI) (+)= scale * x^T (*) dy, split over pixels into `workspace`. */
long long kfb_conv_wgrad_workspace_bytes(const kfb_conv_desc* d);
int kfb_conv_wgrad(const kfb_conv_desc* d, const void* x, const void* dy, long long lddy,
                   float* dw, float* workspace, long long workspace_bytes, int accumulate,
                   float scale, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KFB200_H_ */
