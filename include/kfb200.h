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
  /* element strides of the input between neighbouring pixels / rows; 0 = dense (c and w*c). A
   * pixel stride smaller than c makes neighbouring pixels overlap: the space-to-depth stem reads
   * four 16-channel pixels as one 64-channel pixel this way (kfb_stem_s2d). fprop / wgrad only. */
  int x_pix_stride, x_row_stride;
} kfb_conv_desc;

/* C[M,N] = A[M,K] * B[N,K]^T (+bias[N]) (ReLU). fp16 in, fp16 or fp32 out.
 * T/convnet_builder_legacy.py:318 (xw_plus_b) and 1x1 convolutions. */
int kfb_gemm_tn(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                long long ldb, long long ldc, const float* bias, int relu, int out_fp32,
                float* stats, void* stream);
/* `stats` (optional, fp16 output only): fused batch-norm statistics of the output. For every
 * bn-wide column block the kernel writes `rows` partial rows of [2][bn] floats (per-column sum and
 * sum of squares of the fp16-rounded output): stats[((block * rows + r) * 2 + s) * bn + j].
 * kfb_gemm_stats_layout gives rows / bn for a shape (num_kb = 64-wide K blocks; a convolution has
 * R*S*ceil(C/64)); every row is written exactly once with plain stores (no zero fill needed). The
 * format is what kfb_bn_forward_train_fused consumes - the separate statistics pass over the conv
 * output disappears (the reference gets this from cuDNN's fused batch norm,
 * T/convnet_builder_legacy.py:418-431). */
int kfb_gemm_stats_layout(int M, int N, int num_kb, int* rows, int* bn);
/* The same for the forward GEMM of a convolution writing rows of `ldy` elements (3x3 unit-stride
 * layers traverse rows of Q + 1 virtual pixels, so their tile count is not that of N*P*Q rows). */
int kfb_conv_stats_layout(const kfb_conv_desc* d, long long ldy, int* rows, int* bn);

/* C[M,N] (+)= A[M,K] * B[K,N]; B is consumed in place as an MN-major operand (FC dgrad). */
int kfb_gemm_nn(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                long long ldb, long long ldc, int out_fp32, int accumulate, void* stream);

/* Destination of a fused weight-gradient reduce-scatter. The gradient bucket that contains this
 * output is cut into `nranks` shards of `per` elements; element e of the output has bucket
 * index rel + e and belongs to rank (rel + e) / per; dst[o] is the address on rank o (a
 * peer-mapped pointer for o != this rank) where element 0 of the output lands inside this
 * rank's staging slot. The owners sum their nranks slots in rank order (kfb_allreduce_update_
 * staged), so the result is bit-identical to the pull-based all-reduce. No counterpart in the
 * reference (TF schedules NcclAllReduce after the gradient op, T/allreduce.py:298-299). */
typedef struct kfb_push_target {
  int nranks;
  long long per;
  long long rel;
  void* dst[8];
} kfb_push_target;

/* C[M,N] (fp32) (+)= scale * A[K,M]^T * B[K,N]; K is split across CTAs into `workspace`
 * (fp32 slices of M*N) and reduced in a fixed order. FC weight gradient. */
int kfb_gemm_nt(const void* A, const void* B, float* C, int M, int N, int K, long long lda,
                long long ldb, long long ldc, float* workspace, long long workspace_bytes,
                int accumulate, float scale, void* stream);

/* y = conv(x, w) (+bias) (ReLU). T/convnet_builder_legacy.py:123-141. */
int kfb_conv_fprop(const kfb_conv_desc* d, const void* x, const void* w, void* y, long long ldy,
                   const float* bias, int relu, float* stats, void* stream);

/* dx (+)= conv^T(dy, w) (autodiff of the above, T/benchmark_cnn.py:3104). Strided convolutions
 * run one launch per output phase. When kfb_conv_dgrad_needs_zero_fill(d) is non-zero some
 * pixels of dx receive no contribution and, without `accumulate`, the caller zero-fills dx
 * first. */
int kfb_conv_dgrad_needs_zero_fill(const kfb_conv_desc* d);
int kfb_conv_dgrad(const kfb_conv_desc* d, const void* dy, const void* w, void* dx,
                   int accumulate, void* stream);
/* dx = conv^T(dy, w) + res_src * mask in one kernel: the data gradient of the first convolution
 * of a residual block plus the gradient arriving through the identity shortcut (res_src =
 * gradient of the block output [rows][C] with row stride res_ld, mask = packed ReLU mask of the
 * block output, one bit per element, res_mask_ld bytes per row, NULL = unmasked). The old tiles
 * arrive by TMA load while the MMA runs and the sum leaves by TMA store, so the separate shortcut
 * gradient tensor of T/models/resnet_model.py:72-118's add + relu is never materialised.
 * kfb_conv_dgrad_residual_ok tells whether the shape qualifies (unit stride, C >= 64). */
int kfb_conv_dgrad_residual_ok(const kfb_conv_desc* d);
int kfb_conv_dgrad_residual(const kfb_conv_desc* d, const void* dy, const void* w, void* dx,
                            const void* res_src, long long res_ld,
                            const unsigned char* res_mask, long long res_mask_ld, void* stream);

/* dw (fp32, OHWI) (+)= scale * x^T (*) dy, split over pixels into `workspace`. */
long long kfb_conv_wgrad_workspace_bytes(const kfb_conv_desc* d);
int kfb_conv_wgrad(const kfb_conv_desc* d, const void* x, const void* dy, long long lddy,
                   float* dw, float* workspace, long long workspace_bytes, int accumulate,
                   float scale, void* stream);

/* ------------------------------------------------------------------ */
/* Memory-bound layer kernels (csrc/nn_kernels.cu). Tensors are viewed */
/* as [rows][C] fp16 with an explicit row stride; C % 8 == 0.          */
/* ------------------------------------------------------------------ */

/* Number of stage-1 blocks the per-channel reductions use; scratch buffers passed to the
 * entry points below must hold kfb_reduce_blocks(rows, C) * 2 * C floats. */
int kfb_reduce_blocks(long long rows, int C);

/* Training-mode batch norm fused with ReLU and residual add:
 * y = act(gamma * (x - mean) * rstd + beta (+ residual)); updates the moving statistics.
 * stats_nblocks must be 0 (the conv-epilogue statistics are consumed by the single-launch form
 * below).
 * T/convnet_builder_legacy.py:393-459 (fused_batch_norm + assign_moving_average). */
int kfb_bn_forward_train(const void* x, long long ldx, long long rows, int C, const float* gamma,
                         const float* beta, float* moving_mean, float* moving_var, float eps,
                         float decay, int update_moving, float* mean_rstd, float* scale_shift,
                         float* scratch, int stats_nblocks, const void* residual, long long ldr,
                         int relu, void* y, long long ldy, void* stream);
/* Single-launch forms (csrc/bn_fused.cu): one persistent kernel per direction whose phases
 * (statistics fold / finalize / apply; reductions / coefficients / apply) are separated by
 * grid-wide barriers instead of kernel boundaries. `sync`: two zero-initialised words owned by
 * the caller (one pair per stream of such calls); `partial`: kfb_bn_fused_max_grid() * 2 * C
 * floats of scratch. Forward: `stats` / `stats_rows` / `stats_bn` describe conv-epilogue
 * statistics (kfb_gemm_stats_layout) or stats_rows = 0 to reduce x in the kernel; `bitmask`
 * (optional, [rows][C/8] bytes) receives the packed ReLU mask of the output, which the backward
 * accepts instead of `mask` (the forward output) - it spares two reads of the output tensor for
 * residual blocks. T/convnet_builder_legacy.py:393-459. */
int kfb_bn_fused_max_grid(void);
int kfb_bn_forward_train_fused(const void* x, long long ldx, long long rows, int C,
                               const float* gamma, const float* beta, float* moving_mean,
                               float* moving_var, float eps, float decay, int update_moving,
                               float* mean_rstd, float* scale_shift, const float* stats,
                               int stats_rows, int stats_bn, float* partial, const void* residual,
                               long long ldr, int relu, void* y, long long ldy,
                               unsigned char* bitmask, unsigned* sync, void* stream);
int kfb_bn_backward_fused(const void* dy, long long lddy, const void* mask, long long ldm,
                          const unsigned char* bitmask, const float* relu_affine, const void* x,
                          long long ldx, long long rows, int C, const float* gamma,
                          const float* mean_rstd, float* dgamma, float* dbeta,
                          int accumulate_param_grads, float grad_scale, float* coef,
                          float* partial, void* dz, long long lddz, void* dres, long long lddr,
                          int dres_accumulate, unsigned* sync, void* stream);
/* Two-kernel training forward on conv-epilogue statistics (finalize + apply): the form small
 * layers take, where the second launch costs less than the grid barrier of the single launch. */
int kfb_bn_forward_train_stats(const void* x, long long ldx, long long rows, int C,
                               const float* gamma, const float* beta, float* moving_mean,
                               float* moving_var, float eps, float decay, int update_moving,
                               float* mean_rstd, float* scale_shift, const float* stats,
                               int stats_rows, int stats_bn, const void* residual, long long ldr,
                               int relu, void* y, long long ldy, unsigned char* bitmask,
                               void* stream);
int kfb_bn_forward_inference(const void* x, long long ldx, long long rows, int C,
                             const float* gamma, const float* beta, const float* moving_mean,
                             const float* moving_var, float eps, float* scale_shift,
                             const void* residual, long long ldr, int relu, void* y,
                             long long ldy, void* stream);
/* Backward of the fused op above: dgamma/dbeta, dz (grad wrt x) and optionally the residual
 * gradient. ReLU mask: `mask` (the forward output) if given, else recomputed from x through
 * `relu_affine` (the forward scale_shift; only valid without a residual). coef: 3*C floats. */
int kfb_bn_backward(const void* dy, long long lddy, const void* mask, long long ldm,
                    const float* relu_affine, const void* x, long long ldx, long long rows, int C,
                    const float* gamma,
                    const float* mean_rstd, float* dgamma, float* dbeta,
                    int accumulate_param_grads, float grad_scale, float* coef, float* scratch,
                    void* dz, long long lddz, void* dres, long long lddr, int dres_accumulate,
                    void* stream);
/* dz = dy * (mask > 0); dbias (+)= scale * column sums of dz (conv/FC bias + ReLU backward). */
int kfb_bias_act_backward(const void* dy, long long lddy, const void* mask, long long ldm,
                          long long rows, int C, void* dz, long long lddz, float* dbias,
                          int accumulate, float scale, float* scratch, void* stream);
int kfb_add_act(const void* a, long long lda, const void* b, long long ldb, void* y, long long ldy,
                long long rows, int C, int relu, void* stream);
int kfb_relu_backward(const void* dy, long long lddy, const void* mask, long long ldm, void* dx,
                      long long lddx, int accumulate, long long rows, int C, void* stream);
/* Max / average pooling, NHWC (T/convnet_builder_legacy.py:228-286). argmax: uint8 [N,P,Q,C]. */
int kfb_pool_forward(const void* x, long long ldx, void* y, long long ldy, unsigned char* argmax,
                     int n, int h, int w, int c, int p, int q, int kh, int kw, int sh, int sw,
                     int pad_h, int pad_w, int is_max, void* stream);
int kfb_pool_backward(const void* dy, long long lddy, const unsigned char* argmax, void* dx,
                      long long lddx, int accumulate, int n, int h, int w, int c, int p, int q,
                      int kh, int kw, int sh, int sw, int pad_h, int pad_w, int is_max,
                      void* stream);
int kfb_spatial_mean_forward(const void* x, long long ldx, void* y, long long ldy, int n, int hw,
                             int C, void* stream);
int kfb_spatial_mean_backward(const void* dy, long long lddy, void* dx, long long lddx,
                              int accumulate, int n, int hw, int C, void* stream);
/* Strided 2-D copy / accumulate (concat, split). */
int kfb_copy2d(const void* src, long long lds, void* dst, long long ldd, long long rows, int C,
               int accumulate, void* stream);
/* Row copy with arbitrary column count / strides (flatten into 8-padded rows). */
int kfb_copy_rows(const void* src, long long lds, void* dst, long long ldd, long long rows,
                  long long cols, int accumulate, void* stream);
/* Explicit im2col for stems whose input channel count is not a multiple of 8. */
int kfb_im2col_small_c(const void* x, void* col, int n, int h, int w, int c, int p, int q, int kh,
                       int kw, int sh, int sw, int pad_h, int pad_w, int kpad, void* stream);
/* Batched JPEG decode on the device through nvJPEG (csrc/jpeg_decode.cu; libnvjpeg is loaded with
 * dlopen: kfb_jpeg_available() == 0 when it is missing). backend: 0 = hardware engine if the GPU
 * has one, else GPU-assisted Huffman; 1 = hardware only; 2 = GPU-assisted; 3 = nvJPEG's default.
 * kfb_jpeg_image_info reads the frame header on the host; kfb_jpeg_decode_batch decodes n host
 * bitstreams into n interleaved-RGB device buffers (row pitch >= 3 * width bytes) on `stream`. */
int kfb_jpeg_available(void);
int kfb_jpeg_create(int backend, int cpu_threads, void** handle);
int kfb_jpeg_backend(void* handle);
int kfb_jpeg_destroy(void* handle);
int kfb_jpeg_image_info(void* handle, const unsigned char* data, unsigned long long len, int* width,
                        int* height, int* components);
int kfb_jpeg_decode_batch(void* handle, const unsigned char* const* data,
                          const unsigned long long* lens, int n, unsigned char* const* dst,
                          const int* pitch, void* stream);
/* Space-to-depth stem (csrc/stem_s2d.cu): a stride-2 convolution on C <= 4 channels with
 * ceil(S/2) == 4 runs as a ceil(R/2) x 1 unit-stride convolution over 64-"channel" pixels that
 * are four overlapping 16-channel pixels of the regrouped image (kfb_conv_desc.x_pix_stride = 16).
 * kfb_stem_s2d writes y[n, h2, w2p, 16] (expand = 0; eh / ew = parity of the low padding, lw =
 * zero pixels on the left) or the materialised [n, h2, q, 64] form (expand = 1);
 * kfb_stem_pack_w / kfb_stem_unpack_dw convert [K, kpad] weights / gradients (columns
 * (r*S + s)*C + c) to and from [K, ceil(R/2), 64]. */
int kfb_stem_s2d(const void* x, void* y, int n, int h, int w, int c, int h2, int w2p, int eh,
                 int ew, int lw, int expand, int q, void* stream);
int kfb_stem_pack_w(const void* w, void* w4, int k, int kpad, int r, int s, int c, int ta,
                    void* stream);
int kfb_stem_unpack_dw(const float* dw4, float* dw, int k, int kpad, int r, int s, int c, int ta,
                       int accumulate, float scale, void* stream);
/* Philox dropout; the backward pass calls it again on dy with the same (seed, offset). */
int kfb_dropout(const void* x, void* y, long long n, float keep_prob, unsigned long long seed,
                unsigned long long offset, void* stream);
/* Device-resident step state for CUDA-graph replay of steps with dropout / drop-path: the Philox
 * offset becomes offset + (*step << step_shift), so a captured launch draws a fresh mask in
 * every replay; kfb_step_advance is the step's last node; kfb_set_floats writes per-step host
 * scalars (drop-path keep probabilities) through kernel arguments. */
int kfb_dropout_dev(const void* x, void* y, long long n, float keep_prob, unsigned long long seed,
                    unsigned long long offset, const unsigned long long* step, int step_shift,
                    void* stream);
int kfb_step_advance(unsigned long long* step, void* stream);
int kfb_set_floats(float* dst, const float* host_values, int n, void* stream);
/* Synthetic inputs generated on the device (T/models/model.py:220-237). */
int kfb_synth_images(void* out, long long n, float mean, float stddev, unsigned long long seed,
                     void* stream);
int kfb_synth_labels(int* out, int n, int lo, int hi, unsigned long long seed, void* stream);
int kfb_lrn_forward(const void* x, void* y, long long rows, int C, int radius, float bias,
                    float alpha, float beta, void* stream);
int kfb_lrn_backward(const void* x, const void* dy, void* dx, int accumulate, long long rows,
                     int C, int radius, float bias, float alpha, float beta, void* stream);
/* Sparse softmax cross-entropy, mean loss, fused gradient (x grad_scale) and top-1/top-5 hits
 * (T/models/model.py:285-312, loss-scale multiply of T/benchmark_cnn.py:3102-3103).
 * out3 = {mean loss, #top-1 hits, #top-5 hits}. dyn_scale (device, optional) multiplies
 * grad_scale (device-resident loss scale). */
int kfb_softmax_xent(const float* logits, long long ldl, const int* labels, int rows, int classes,
                     float grad_scale, const float* dyn_scale, void* dlogits, long long lddl,
                     float* loss_rows, int* top1, int* top5, float* out3, void* stream);
/* Allocation-free variant of kfb_softmax_xent (no counterpart in the reference; it exists for
 * callers of this library that cannot allocate device memory: hot loops that must not hit the
 * implicit synchronisation of cudaMalloc, and hosts that own the
 * whole device heap). The per-row scratch lives in a 48 KiB static device array inside the
 * library, so rows <= kfb_softmax_xent_static_max_rows() (4096); larger batches are rejected
 * with an error. No cudaMalloc / cudaFree / allocating call is made. The first call on each
 * device resolves the array's device address (cudaGetSymbolAddress) - make that first call
 * before starting any stream capture. All calls share the one array: it keeps the previous
 * call's contents (harmless - each call rewrites the rows it reads), and two calls that may
 * run concurrently on the same GPU (different streams) race on it and must be serialised by
 * the caller. out3 = {mean loss, #top-1 hits, #top-5 hits} as for kfb_softmax_xent. */
int kfb_softmax_xent_static_max_rows(void);
int kfb_softmax_xent_static(const float* logits, long long ldl, const int* labels, int rows,
                            int classes, float grad_scale, const float* dyn_scale, void* dlogits,
                            long long lddl, float* out3, void* stream);
int kfb_cast_f32_f16(const float* in, void* out, long long n, void* stream);

/* ------------------------------------------------------------------ */
/* Optimizers, P2P collectives, symmetric memory (csrc/optim_comm.cu)  */
/* ------------------------------------------------------------------ */

typedef struct {
  int opt;            /* 0 sgd, 1 momentum (Nesterov), 2 rmsprop, 3 adam */
  float lr;
  float momentum;     /* momentum | rmsprop momentum | adam beta1 */
  float decay;        /* rmsprop decay | adam beta2 */
  float epsilon;
  float weight_decay; /* g += weight_decay * w (the L2 term of T/benchmark_cnn.py:3068-3099) */
  float grad_scale;   /* multiplies the reduced gradient (1/loss_scale, 1/N for mean modes) */
  float clip;         /* elementwise clip_by_value bound, <= 0 disables */
  float adam_t;       /* step count for Adam bias correction */
  /* optional (device): one byte per 128-element block of the arena range the call covers, 0 =
   * no L2 term for that block (T/models/ssd_model.py:91-94: batch-norm variables are not
   * decayed); NULL = every variable is decayed */
  const unsigned char* decay_mask;
} kfb_hyper;

/* Fused optimizer over a flat arena; dyn (device, optional) = {lr, grad_scale, adam_t, -}
 * overrides the by-value fields; skip_flag (device, optional) != 0 skips the update. */
int kfb_optimizer_step(float* w, const float* g, float* m, float* v, void* w16, long long n,
                       const kfb_hyper* hp, const float* dyn, const int* skip_flag, void* stream);
/* SynchronousAveraging from a local copy of the averaged model (kfb_allreduce_update2 mode 3
 * produced it on the exchange stream during backward): w <- (1-alpha) w + alpha avg, then the
 * local gradient step. T/benchmark_cnn.py:1199-1201. */
int kfb_optimizer_step_blend(float* w, const float* avg, const float* g, float* m, float* v,
                             void* w16, long long n, const kfb_hyper* hp, const float* dyn,
                             float alpha, void* stream);
int kfb_check_finite(const float* g, long long n, int* flag, void* stream);
/* cudaMemsetAsync(p, 0, bytes) on `stream` (gradient buffers of strided 1x1 dgrads, scatter
 * tables): the step issues no framework fill kernels. */
int kfb_fill_zero(void* p, long long bytes, void* stream);
/* dyn[i] <- the argument for every bit i of `mask` (lr, grad_scale, adam_t), from a one-thread
 * kernel on `stream`: per-step values for captured graphs / device-resident loss scaling without
 * a pinned staging buffer. */
int kfb_set_hyper_dyn(float* dyn, float lr, float grad_scale, float adam_t, int mask,
                      void* stream);
int kfb_loss_scale_update(float* state, int* flag, int inc_every, float* dyn,
                          float base_grad_scale, void* stream);

/* Two-shot P2P all-reduce over mapped peer arenas fused with the optimizer (mode 0),
 * SynchronousAveraging of the model (mode 1), or reduce-only into `reduced` (mode 2, used by
 * variable_consistency=relaxed where the sum is applied one step later). src / reduced / flags are host arrays of
 * nranks device pointers (peer mappings); seq must increase by one per collective call and
 * be identical on every rank. */
int kfb_flag_page_bytes(void);
/* sizeof of the structures that cross this ABI as the library was built: 0 kfb_conv_desc,
 * 1 kfb_hyper, 2 kfb_push_target (-1 otherwise). Bindings check their mirrors against it. */
int kfb_sizeof(int which);
/* Device-resident sequence offset for multi-rank CUDA graphs (B200 addition; the reference has no
 * graph replay): once set, every collective below adds the uint32 at `counter` (device memory)
 * to its `seq`; kfb_seq_advance bumps it on a stream (the last node of a captured step). */
int kfb_set_seq_base(const void* counter);
int kfb_seq_advance(void* counter, unsigned k, void* stream);
int kfb_allreduce_update(int mode, int rank, int nranks, void* const* src, void* const* reduced,
                         void* const* flags, long long offset, long long count, unsigned seq,
                         float* w, const float* g_local, float* m, float* v, void* w16,
                         const kfb_hyper* hp, const float* dyn, const int* skip_flag,
                         float sma_alpha, int blocks, void* stream);
/* NVLS form of kfb_allreduce_update (--comm_impl=nvls; W10 "nvls" of SURVEY.md 8, the
 * in-switch counterpart of the NcclAllReduce path T/allreduce.py:280-324): the reduction is one
 * multimem.ld_reduce and the broadcast one multimem.st on the multicast aliases `src_mc` /
 * `reduced_mc` of the source arena and of the landing buffer (both must be bound to a multicast
 * object on every rank; parallel/comm.py does it through torch symmetric memory);
 * `reduced_local` is this rank's own landing buffer. Same modes, flags and seq rules. */
int kfb_allreduce_update_nvls(int mode, int rank, int nranks, const void* src_mc,
                              void* reduced_mc, void* reduced_local, void* const* flags,
                              long long offset, long long count, unsigned seq, float* w,
                              const float* g_local, float* m, float* v, void* w16,
                              const kfb_hyper* hp, const float* dyn, const int* skip_flag,
                              float sma_alpha, int blocks, void* stream);
/* Second generation of the two entry points above, one kernel template for both forms. Pull
 * form: src / reduced = per-rank addresses, src_mc = reduced_mc = NULL. NVLS form: src_mc /
 * reduced_mc = multicast aliases (src / reduced may be NULL). Modes: 0 gradients + optimizer,
 * 1 SMA in one kernel, 2 reduce only (sum), 3 average only (the SMA weight exchange launched at
 * the start of the step; kfb_optimizer_step_blend applies it after backward). half = 1 is the
 * compact gradient transfer of the reference (--compact_gradient_transfer: fp16 whenever use_fp16,
 * T/variable_mgr.py:311-313, T/batch_allreduce.py:429-453): sources and landing buffers hold
 * fp16, the sum is taken in fp32; with cast_from / cast_to this rank's fp32 gradients of the
 * range are first copied as fp16 into its source arena on the same stream. offset and count
 * must be multiples of 8. Phase 2 keeps four 16-byte vectors per array in flight per thread, so
 * the few CTAs the kernel runs with (backward wants the SMs) still stream at HBM rate. */
int kfb_allreduce_update2(int mode, int rank, int nranks, void* const* src, void* const* reduced,
                          const void* src_mc, void* reduced_mc, const void* reduced_local,
                          void* const* flags, int half, const float* cast_from, void* cast_to,
                          long long offset, long long count, unsigned seq, float* w,
                          const float* g_local, float* m, float* v, void* w16,
                          const kfb_hyper* hp, const float* dyn, const int* skip_flag,
                          float sma_alpha, int blocks, void* stream);
/* Owner side of the fused weight-gradient reduce-scatter (see kfb_push_target): sums the nranks
 * local staging slots of this rank's shard of [offset, offset+count), pushes the sum to every
 * rank and applies the optimizer - kfb_allreduce_update with local instead of peer sources. */
int kfb_allreduce_update_staged(int rank, int nranks, void* const* reduced, void* const* flags,
                                long long offset, long long count, unsigned seq, float* w,
                                float* m, float* v, void* w16, const kfb_hyper* hp,
                                const float* dyn, const int* skip_flag, const float* stage,
                                long long stage_stride, int blocks, void* stream);
/* Pushes arena ranges [lo, hi) (pairs in device memory, multiples of 4) of `grad` into the
 * owners' staging slots; slot_base[o] = this rank's slot on rank o at arena index 0. */
int kfb_push_ranges(const float* grad, const long long* ranges_dev, int nranges,
                    long long max_range_elems, long long bucket_lo, long long per, int nranks,
                    void* const* slot_base, void* stream);
int kfb_broadcast_pull(int rank, int nranks, void* const* src, void* const* flags, int root,
                       long long offset, long long count, unsigned seq, float* dst, void* dst16,
                       int blocks, void* stream);
int kfb_device_barrier(int rank, int nranks, void* const* flags, unsigned seq, void* stream);
int kfb_flag_or(int rank, int nranks, void* const* flags, unsigned seq, int* flag, void* stream);
/* PairAveraging (AD-PSGD, T/benchmark_cnn.py:1196-1198): pull one peer's newest published model
 * into `pull_buf` (count floats, local) with 16-byte NVLink loads, validate the pull against the
 * peer's version words and repeat it (up to three times; torn_counter counts the repeats) if the
 * peer overwrote the slot meanwhile, then blend 1/2, apply the local gradient and publish.
 * `sync`: four zero-initialised words owned by the caller (grid barrier + retry flag). */
int kfb_pair_average_update(const float* peer_pub, const unsigned* peer_ver, float* my_pub,
                            unsigned* my_ver, long long stride, long long count, float* w,
                            const float* g, float* m, float* v, void* w16, const kfb_hyper* hp,
                            const float* dyn, unsigned my_next_version, int has_peer,
                            int* torn_counter, float* pull_buf, unsigned* sync, int blocks,
                            void* stream);

/* Symmetric memory over CUDA IPC (cudaMalloc'd, zero-filled). */
int kfb_symm_alloc(long long bytes, void** out);
int kfb_symm_free(void* p);
int kfb_ipc_handle_bytes(void);
int kfb_ipc_get_handle(void* p, void* handle_out);
int kfb_ipc_open(const void* handle, void** out);
int kfb_ipc_close(void* p);
int kfb_enable_peer_access(int peer_device);

/* Weight-gradient GEMM / convolution fused with the reduce-scatter of data-parallel training:
 * the fp32 result is not written locally but scattered over NVLink into the owners' staging
 * slots - by the epilogue of the tcgen05 kernel itself when one K split suffices (scale 1), by
 * the split-K reduction kernel otherwise. */
int kfb_gemm_nt_rs(const void* A, const void* B, int M, int N, int K, long long lda,
                   long long ldb, float* workspace, long long workspace_bytes, float scale,
                   const kfb_push_target* push, void* stream);
int kfb_conv_wgrad_rs(const kfb_conv_desc* d, const void* x, const void* dy, long long lddy,
                      float* workspace, long long workspace_bytes, float scale,
                      const kfb_push_target* push, void* stream);

/* ---- wider model zoo (csrc/extra_ops.cu) ------------------------------------------------ */
/* Activations on NHWC fp16 rows. kind: 1 relu, 2 relu6 (MobileNet-v2,
 * T/models/mobilenet_v2.py:42-82), 3 tanh, 4 sigmoid. Backward reads the stored output y. */
int kfb_act_forward(const void* x, long long ldx, void* y, long long ldy, long long rows, int C,
                    int kind, void* stream);
int kfb_act_backward(const void* dy, long long lddy, const void* y, long long ldy, void* dx,
                     long long lddx, int accumulate, long long rows, int C, int kind,
                     void* stream);
/* Depthwise convolution, channel multiplier 1 (slim.separable_conv2d with num_outputs=None,
 * T/models/mobilenet_conv_blocks.py:246-256; the depthwise half of NASNet's separable convs,
 * T/models/nasnet_utils.py:212-260). d->k must equal d->c; w is [R][S][C] fp16, dw fp32. */
int kfb_dwconv_fprop(const kfb_conv_desc* d, const void* x, long long ldx, const void* w, void* y,
                     long long ldy, void* stream);
int kfb_dwconv_dgrad(const kfb_conv_desc* d, const void* dy, long long lddy, const void* w,
                     void* dx, long long lddx, int accumulate, void* stream);
long long kfb_dwconv_wgrad_workspace_bytes(const kfb_conv_desc* d);
int kfb_dwconv_wgrad(const kfb_conv_desc* d, const void* x, long long ldx, const void* dy,
                     long long lddy, float* dw, void* workspace, long long workspace_bytes,
                     int accumulate, float scale, void* stream);
/* Embedding lookup of a column range [col_lo, col_lo+cols) of a [rows][ldt] fp16 table, and the
 * dense scatter-add of its gradient into the fp32 gradient of the table (zeroed by the caller;
 * fp32 atomics, so the summation order is not fixed). NeuMF embeddings,
 * T/models/experimental/official_ncf_model.py:62-83. */
int kfb_embedding_gather(const void* table, long long ldt, int col_lo, const int* ids,
                         long long num_rows_table, void* y, long long ldy, long long n, int cols,
                         void* stream);
int kfb_embedding_scatter_add(const void* dy, long long lddy, const int* ids,
                              long long num_rows_table, float* dtable, long long ldt, int col_lo,
                              long long n, int cols, void* stream);
/* Elementwise product (the GMF branch of NeuMF) and its backward for one operand. */
int kfb_mul_forward(const void* a, long long lda, const void* b, long long ldb, void* y,
                    long long ldy, long long rows, int C, void* stream);
int kfb_mul_backward(const void* dy, long long lddy, const void* other, long long ldo, void* dx,
                     long long lddx, int accumulate, long long rows, int C, void* stream);
/* NCF loss: 2-class softmax cross-entropy over [1, z] with labels in {0,1}
 * (T/models/experimental/official_ncf_model.py:85-97). out[0] mean loss, out[1] correct count;
 * dz (fp16, may be null) = scale * dyn * (p1 - y). */
int kfb_ncf_loss(const float* z, long long ldz, const int* labels, int n, float scale,
                 const float* dyn_scale, void* dz, long long lddz, float* out, void* stream);
/* Pointwise part of one BasicLSTMCell time step for the whole batch (gate order i, j, f, o,
 * forget_bias added to f; T/models/experimental/deepspeech.py:232-268). gates_io holds the
 * input-projection pre-activations on entry and the activated gates on exit (backward: the
 * activated gates on entry, the pre-activation gradients on exit); gh is the recurrent
 * projection h_{t-1}.Wh; c is kept in fp32. The GEMMs of the layer are the kfb_gemm_* calls. */
int kfb_lstm_cell_forward(void* gates_io, long long ldg, const void* gh, long long ldgh,
                          const float* bias, const float* c_prev, long long ldcp, float* c_out,
                          long long ldc, void* h_out, long long ldh, void* h_next, long long ldhn,
                          int n, int H, float forget_bias, void* stream);
int kfb_lstm_cell_backward(void* gates_io, long long ldg, const void* dy, long long lddy,
                           const void* dh_rec, long long lddh, const float* c, long long ldc,
                           const float* c_prev, long long ldcp, float* dc_io, int n, int H,
                           void* stream);
/* y[r][c] (+)= idx[c] >= 0 ? x[r][idx[c]] : 0 over C output channels (idx on the device):
 * converts between two zero-padded channel layouts of one logical tensor (nn/legalize.py). */
int kfb_channel_gather(const void* x, long long ldx, void* y, long long ldy, int accumulate,
                       long long rows, int C, const int* idx, void* stream);
/* Per-sample drop-path (T/models/nasnet_utils.py:113-138): y (+)= x * (u_n < keep ? 1/keep : 0)
 * with u_n a pure function of (seed, offset, sample); the same call replays it for dy. */
int kfb_drop_path(const void* x, long long ldx, void* y, long long ldy, int accumulate,
                  long long rows, long long rows_per_sample, int C, float keep_prob,
                  unsigned long long seed, unsigned long long offset, void* stream);
int kfb_drop_path_dev(const void* x, long long ldx, void* y, long long ldy, int accumulate,
                      long long rows, long long rows_per_sample, int C, float keep_prob,
                      unsigned long long seed, unsigned long long offset,
                      const unsigned long long* step, int step_shift, const float* keep_dev,
                      void* stream);

/* ---- loss heads (csrc/loss_kernels.cu) ---------------------------------------------------- */
/* SSD-300 multibox loss (T/models/ssd_model.py:299-384): smooth-L1 on matched anchors + softmax
 * cross-entropy on positives and the 3:1 hardest negatives, each normalised by the sample's
 * number of matched boxes, mean over the batch. Reads the 6 loc / 6 conf head convolutions
 * (NHWC fp16, row strides ld_*) in place and writes their gradients in place; gt_loc
 * [N,A,4], gt_label [N,A], num_matched [N] are fp32 as in the reference's input pipeline.
 * out[0] total, out[4] class part, out[5] box part. dloc/dconf may be null (forward only). */
long long kfb_ssd_loss_workspace_floats(int batch, int num_anchors);
int kfb_ssd_loss(int batch, int num_classes, const int* map_size, const int* map_ndef,
                 void* const* loc, const long long* ld_loc, void* const* conf,
                 const long long* ld_conf, void* const* dloc, void* const* dconf,
                 const float* gt_loc, const float* gt_label, const float* num_matched,
                 int negs_per_positive, float scale, const float* dyn_scale, float* workspace,
                 long long workspace_floats, float* out, void* stream);

/* CTC loss as the reference computes it (T/models/experimental/deepspeech.py:359-396):
 * emissions (softmax(z)+eps)/(1+C*eps), blank = C-1, mean over the batch, samples whose labels
 * do not fit are ignored. z [batch*T][ldz] fp32 (row n*T+t), labels [batch][max_label_len];
 * dz fp16 (may be null) receives scale*dyn*dL/dz. out[0] = mean loss. */
long long kfb_ctc_loss_workspace_floats(int batch, int T, int max_label_len);
int kfb_ctc_loss(const float* z, long long ldz, int batch, int T, int C, const int* labels,
                 int max_label_len, const int* input_len, const int* label_len, float eps,
                 float scale, const float* dyn_scale, void* dz, long long lddz, float* workspace,
                 long long workspace_floats, float* out, void* stream);

/* ------------------------------------------------------------------ */
/* Input preprocessing on the device (csrc/image_ops.cu)               */
/* ------------------------------------------------------------------ */
/* Batch form of kfb_image_transform (include/kfb200_host.h, same kfb_image_op descriptor and
 * arithmetic) as CUDA kernels: `srcs` = n device addresses of raw uint8 RGB images, `ops` = n
 * descriptors in device memory (methods 0 bilinear, 1 nearest, 4 crop_and_resize; bicubic / area
 * stay on the host), `means` = 3 n floats of scratch, `out` = n images `image_stride` elements
 * apart. The host threads then only decode JPEG; crop / resize / mirror / colour jitter /
 * normalise / fp16 happen where the batch is consumed (the reference runs this stage in tf.data,
 * T/preprocessing.py:192-307). kfb_image_op_bytes_device = sizeof(kfb_image_op) in this library. */
int kfb_image_op_bytes_device(void);
int kfb_image_transform_batch(const void* srcs, const void* ops, int n, int max_pixels,
                              float* means, void* out, long long image_stride, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KFB200_H_ */
