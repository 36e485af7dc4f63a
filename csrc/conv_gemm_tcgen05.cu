// tcgen05 / TMEM / TMA GEMM and implicit-GEMM convolution for sm_100a.
//
// One warp-specialised kernel covers the five GEMM-shaped hot ops of the CNN
// zoo (reference call sites: tf.nn.conv2d / conv2d gradients / xw_plus_b at
// scripts/tf_cnn_benchmarks/convnet_builder_legacy.py:123-141,318 and
// tf.gradients at benchmark_cnn.py:3104; there they are cuDNN / cuBLAS calls):
//
//   GEMM_TN : C[M,N] = A[M,K] * B[N,K]^T             (FC fprop / dgrad, 1x1 conv)
//   GEMM_NT : C[M,N] = A[K,M]^T * B[K,N]             (FC wgrad)
//   GEMM_NN : C[M,N] = A[M,K] * B[K,N]               (FC dgrad; B read MN-major in place)
//   FPROP   : y = conv(x, w)       A = im2col(x) via TMA im2col mode, B = w[Cout][R*S*C]
//   DGRAD   : dx = conv^T(dy, w)   A = im2col(dy), B = w read in place, MN-major, taps flipped
//   WGRAD   : dw = x^T (*) dy      A = dy (MN-major), B = im2col(x) (MN-major), split over pixels
//
// Layout: activations NHWC fp16, filters OHWI (= [Cout][R][S][C]) fp16, fp32
// accumulation in TMEM. Tile = 128 x BN (BN in {64,128,256}), K-block = 64
// fp16 = one 128-byte swizzled row. Roles per CTA (192 threads):
//   warp 0 : TMA producer (one elected lane)
//   warp 1 : TMEM allocator + tcgen05.mma issuer (one elected lane)
//   warps 2-5 : epilogue (tcgen05.ld -> bias/ReLU -> global), one TMEM lane quarter each
#include "kfb_common.cuh"
#include "../include/kfb200.h"

#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

namespace kfb {

// ---------------------------------------------------------------------------
// error string
// ---------------------------------------------------------------------------
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

enum GemmMode : int { GEMM_TN = 0, GEMM_NT = 1, FPROP = 2, DGRAD = 3, WGRAD = 4, GEMM_NN = 5 };

struct GemmParams {
  int mode;
  int M, N;           // logical output extent (rows, cols) of one split
  int num_kb;         // total number of 64-wide K blocks
  int kb_per_split;   // K blocks handled by one blockIdx.z
  // im2col geometry (FPROP / DGRAD: pixel = output row; WGRAD: pixel = K index)
  int P, Q;           // im2col pixel grid (rows x cols) per image
  int lower_h, lower_w;
  int stride_h, stride_w;
  int R, S;           // filter taps
  int CB;             // 64-wide channel blocks per tap (FPROP/DGRAD)
  int Ck;             // K extent per tap in elements (FPROP: C, DGRAD: Cout)
  int Cn;             // WGRAD: channels per tap (N extent per tap)
  int NCB;            // WGRAD: BN-wide channel blocks per tap
  // epilogue
  void* out;
  long long ldc;           // elements between output rows
  long long split_stride;  // elements between split-K slices
  const float* bias;
  int relu;
  int out_fp32;
  int accumulate;          // out += result (gradient accumulation for fan-out tensors)
  int stages;              // smem pipeline depth of this launch (1..kMaxStages)
  int m_tiles, n_tiles, splits;  // tile space walked by the persistent CTAs
  int wg_flat;             // WGRAD: N is the flattened (tap, channel) axis (C % 64 == 0)
  // optional fused batch-norm statistics of the (fp16-rounded) output: per-column sum and sum
  // of squares. Every (n-block, CTA slot, TMEM lane quarter) owns one row of [2][BN] floats:
  // stats[((n_blk * stats_rpn + slot) * 4 + quarter)][2][BN], slot = position of the CTA among
  // the stats_rpn = min(m_tiles, grid) CTAs that visit the n-block. The persistent tile order
  // (m fastest) makes every CTA visit an n-block in one run, so each row is written exactly
  // once with plain stores: no zero fill, no atomics, deterministic. BN = 64: eight rows per
  // slot (the two warps of a quarter take alternate tiles and own a row each).
  float* stats;
  int stats_rpn;
  // optional row scatter (strided dgrad): row (n,i,j) -> (n, h0 + i*sh, w0 + j*sw) of H x W
  int sc_P, sc_Q, sc_H, sc_W, sc_sh, sc_sw, sc_h0, sc_w0;
  // n / d == (n * mul) >> sh for 0 <= n < 2^31 (filled by the launcher from sc_P, sc_Q)
  unsigned long long sc_pq_mul, sc_q_mul;
  int sc_pq_sh, sc_q_sh;
  // DGRAD: filter tap visited at (r, s) of this launch's (sub-)filter
  int wt_r0, wt_rstep, wt_s0, wt_sstep, wt_S;
  // Fused reduce-scatter (weight gradients, data-parallel training): when push_n > 1 the
  // epilogue stores each 16-byte piece of the fp32 result straight into the staging slot of
  // the rank that owns it - a peer-mapped address over NVLink for every owner but this rank -
  // so the gradient transfer overlaps the MMA main loop tile by tile. Element e of the output
  // belongs to owner #{r >= 1 : e >= push_bound[r]}; push_base[o] is where e = 0 lands on o.
  int push_n;
  long long push_bound[8];
  float* push_base[8];
  // TMA-store epilogue (default, KFB_TMA_STORE=0 disables): the staged 32 x 64 fp16 chunks leave
  // through cp.async.bulk.tensor stores on `tmC` instead of per-lane 16-byte stores. Host-side
  // eligibility: fp16 output, no accumulate / row scatter / split-K / push.
  int tma_store;
  // Accumulating TMA-store epilogue with a masked source (fused residual gradient): the values
  // added to the result are acc_src * mask, acc_src being another [M][N] fp16 tensor read
  // through tensor map D and mask one bit per element (byte (row, col / 8), bit col % 8) - the
  // packed ReLU mask the batch-norm forward wrote. Null mask: plain out += result (D = C).
  const unsigned char* acc_mask;
  long long acc_mask_ld;  // bytes per row
  const void* acc_src;    // host side only: the tensor map D is encoded over it
  long long acc_src_ld;   // elements per row
  // SV3: rows are virtual pixels of (sc_Q = Q + 1)-wide rows; column sc_Q - 1 is a dummy
  int sv3;
  int bres_bytes;  // BRES: size of the resident B operand (all K blocks)
};

// PRMT in its generic form: selector nibble bit 3 replicates the sign of the selected byte over
// the result byte (__byte_perm masks that bit away)
__device__ __forceinline__ unsigned prmt_sign(unsigned a, unsigned sel) {
  unsigned d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(0u), "r"(sel));
  return d;
}
__device__ __forceinline__ unsigned hadd2_u32(unsigned a, unsigned b) {
  const __half2 r =
      __hadd2(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
  return *reinterpret_cast<const unsigned*>(&r);
}

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kTileABytes = kBlockM * kBlockK * 2;  // 16 KiB
constexpr int kMaxStages = 8;
constexpr int kStageRowBytes = 144;                 // 128 B payload + 16 B pad (conflict free)
constexpr int kSmemBudget = 227 * 1024;

template <int BN>
struct TileCfg {
  static constexpr int kTileBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kTileABytes + kTileBBytes;
  // Accumulator stages in TMEM (512 columns): four for tiles of at most 128 columns, so that the
  // MMA warp runs up to three tiles ahead of the epilogue (with two stages the tile time of the
  // short-K 64-channel layers was (MMA + epilogue) / 2: each stage's cycle is fill -> drain)
  static constexpr int kAccStages = BN <= 128 ? 4 : 2;
  static constexpr int kTmemCols = kAccStages * BN;  // a power of two >= 128
  // Narrow tiles are picked for short-K, store-bound problems: give them eight epilogue warps
  // (two per TMEM lane quarter, interleaving column chunks); 256-wide tiles keep four and a
  // deeper smem ring instead.
  static constexpr int kEpiWarps = 8;
  static constexpr int kThreads = 64 + 32 * kEpiWarps;
  // per epilogue warp: one 32-row tile with padded rows (per-lane store epilogue), or two dense
  // swizzled 4 KiB tiles (TMA-store epilogue: the store of chunk k reads one tile while chunk
  // k+1 is packed into the other)
  static constexpr int kStagingBytes = kEpiWarps * 8192;
  static constexpr int kFixedBytes = kStagingBytes + 1024 /*align*/ + 512 /*barriers*/;
  static constexpr int kMaxFit = (kSmemBudget - kFixedBytes) / kStageBytes;
  static constexpr int max_stages() { return kMaxFit > kMaxStages ? kMaxStages : kMaxFit; }
  static constexpr int smem_bytes(int stages) { return stages * kStageBytes + kFixedBytes; }
  // SV3 (shifted-view 3-tap rows, see the kernel): one stage = a (128 + 2)-pixel A box (17
  // swizzle atoms reserved) plus the three B tiles of a filter row
  static constexpr int kSv3ABytes = 17 * 1024;
  static constexpr int kSv3StageBytes = kSv3ABytes + 3 * kTileBBytes;
  static constexpr int kSv3MaxFit = (kSmemBudget - kFixedBytes) / kSv3StageBytes;
  static constexpr int max_stages_sv3() { return kSv3MaxFit > kMaxStages ? kMaxStages : kSv3MaxFit; }
  static constexpr int smem_bytes_sv3(int stages) { return stages * kSv3StageBytes + kFixedBytes; }
  // BRES: `bres` bytes of resident B; the ring holds A tiles only
  static constexpr int kBresMax = 72 * 1024;
  static constexpr int stages_bres(int a_bytes, int bres) {
    const int fit = (kSmemBudget - kFixedBytes - bres) / a_bytes;
    return fit > kMaxStages ? kMaxStages : fit;
  }
  static constexpr int smem_bytes_bres(int a_bytes, int bres, int stages) {
    return stages * a_bytes + bres + kFixedBytes;
  }
  // accumulating TMA-store epilogue: two 4 KiB tiles per epilogue warp receive the old values
  // of the output (TMA loads, prefetched while the MMA of the tile still runs)
  static constexpr int kOldBytes = kEpiWarps * 2 * 4096;
  static constexpr int kMaxFitAcc = (kSmemBudget - kFixedBytes - kOldBytes - 512) / kStageBytes;
  static constexpr int max_stages_acc() { return kMaxFitAcc > kMaxStages ? kMaxStages : kMaxFitAcc; }
  static constexpr int smem_bytes_acc(int stages) {
    return stages * kStageBytes + kFixedBytes + kOldBytes + 512;
  }
};

// Persistent kernel: one CTA per SM walks the tile list (m fastest, then n, then split).
// Three pipelines run concurrently: TMA -> smem ring (full/empty mbarriers), MMA -> TMEM
// accumulator pair (tmem_full/tmem_empty), and the epilogue draining accumulator `a` while the
// MMA warp already fills accumulator `a^1` for the next tile.
// FAST = every output piece is a full, 16-byte aligned vector (all conv layers): the scalar
// fallbacks are compiled out, which shrinks the kernel below the instruction-cache size.
// TMA_ST = the TMA-store epilogue (fp16 output without accumulate / scatter / split-K): a
// separate instantiation, so the other variants carry none of its code.
// ACC (with TMA_ST) = out += result: the old 32 x 64 chunks arrive by TMA load into shared memory
// (issued before the wait for the accumulator, so their latency hides behind the MMA), are added
// in fp32 and the sum leaves through the same TMA store; replaces the per-lane read-modify-write.
// SV3 (FPROP / DGRAD, unit stride, three taps per filter row, C % 64 == 0): the im2col re-read
// of the input - nine boxes per 3x3 tile, the L2->SM traffic that bounds these layers - drops
// to three. The output is traversed as rows of Q + 1 virtual pixels (one dummy column after the
// last real one), for which tap s of virtual pixel m is tap 0 of virtual pixel m + s - also
// across a row end, where the wrapped pixel is the left padding column of the next row, i.e.
// the same zeros as the right padding it stands for. So ONE TMA im2col box of 128 + 2 pixels
// per filter row serves all three taps: the MMA reads tap s through a descriptor whose start
// address is s rows (s x 128 B) into the box (tcgen05 applies the 128B swizzle to absolute
// address bits, so a row-shifted start needs no re-layout and no base offset; measured on
// B200). The dummy column is computed and dropped by the epilogue (1.8 % at Q = 56, 3.5 % at
// 28, 7 % at 14).
// BRES (FPROP / DGRAD with one column tile, i.e. N <= BN): the whole B operand (all K blocks of
// the weights, p.bres_bytes <= 72 KiB) is loaded once per CTA and stays resident in shared
// memory; the ring stages carry A only. For the 64-channel layers the weight tiles are 60 % of
// the L2 -> SM traffic of a tile (72 of 122 KiB for a 3x3), and that traffic is their bound.
template <int BN, int MODE, bool FAST, bool TMA_ST = false, bool ACC = false, bool SV3 = false,
          bool BRES = false>
__global__ void __launch_bounds__(TileCfg<BN>::kThreads, 1)
    gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA,
                        const __grid_constant__ CUtensorMap tmB,
                        const __grid_constant__ CUtensorMap tmC,
                        const __grid_constant__ CUtensorMap tmD, const GemmParams p) {
  using Cfg = TileCfg<BN>;
  const int kStages = p.stages;
  constexpr uint32_t kAMajor = (MODE == GEMM_NT || MODE == WGRAD) ? kMajorMN : kMajorK;
  // only the weight-gradient modes carry the fused reduce-scatter epilogue
  constexpr bool kCanPush = (MODE == GEMM_NT || MODE == WGRAD);
  constexpr uint32_t kBMajor =
      (MODE == GEMM_NT || MODE == WGRAD || MODE == DGRAD || MODE == GEMM_NN) ? kMajorMN : kMajorK;

  extern __shared__ uint8_t smem_raw[];
  // 1 KiB aligned; pointer arithmetic on the array (not an integer round trip) so that the
  // compiler keeps the shared address space and emits LDS / STS for the staging tiles
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int kABytes = SV3 ? Cfg::kSv3ABytes : kTileABytes;
  constexpr int kBBytes = SV3 ? 3 * Cfg::kTileBBytes : Cfg::kTileBBytes;   // per stage
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;   // BRES: the resident operand, p.bres_bytes
  uint8_t* staging = BRES ? smem_b + p.bres_bytes : smem + kStages * (kABytes + kBBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + Cfg::kStagingBytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  constexpr int kAcc = Cfg::kAccStages;
  uint64_t* tmem_full_bar = empty_bar + kMaxStages;   // [4]
  uint64_t* tmem_empty_bar = tmem_full_bar + 4;       // [4]
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty_bar + 4);
  uint64_t* old_bar = tmem_empty_bar + 6;  // [kEpiWarps][2] (ACC only; inside the 512 B block)
  uint8_t* old_area = staging + Cfg::kStagingBytes + 1024;  // 1 KiB aligned (ACC only)
  uint64_t* bres_bar = tmem_empty_bar + 22;                 // (BRES only)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  // 64-wide fp16 tiles are one column chunk: instead of leaving half of the epilogue warps idle,
  // the two warps of a TMEM lane quarter take alternate tiles (= alternate accumulator stages)
  const bool alt_tiles = (BN == 64) && !p.out_fp32;
  const int tiles_mn = p.m_tiles * p.n_tiles;
  const int total_tiles = tiles_mn * p.splits;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < kAcc; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      // one arrive per epilogue warp that reads the stage
      mbar_init(&tmem_empty_bar[a], alt_tiles ? Cfg::kEpiWarps / 2 : Cfg::kEpiWarps);
    }
    if (ACC) {
      for (int i = 0; i < 2 * Cfg::kEpiWarps; ++i) mbar_init(&old_bar[i], 1);
    }
    if (BRES) mbar_init(bres_bar, 1);
    mbar_fence_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) {
    tmem_alloc(tmem_ptr_smem, Cfg::kTmemCols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      if (BRES && static_cast<int>(blockIdx.x) < total_tiles) {
        // every K block of the weights, once (n0 = 0: one column tile)
        mbar_expect_tx(bres_bar, static_cast<uint32_t>(p.bres_bytes));
        // in the order the MMA warp walks them: SV3 (r, cb, tap); otherwise (r, s, cb)
        uint8_t* dst = smem_b;
        if (SV3) {
          for (int r = 0; r < p.R; ++r)
            for (int cb = 0; cb < p.CB; ++cb)
              for (int ts = 0; ts < 3; ++ts, dst += Cfg::kTileBBytes) {
                if (MODE == FPROP) {
                  tma_load_2d(dst, &tmB, bres_bar, (r * p.S + ts) * p.Ck + cb * kBlockK, 0);
                } else {
                  const int wtap = (p.wt_r0 + p.wt_rstep * (p.R - 1 - r)) * p.wt_S +
                                   (p.wt_s0 + p.wt_sstep * (p.S - 1 - ts));
#pragma unroll
                  for (int j = 0; j < BN / 64; ++j)
                    tma_load_3d(dst + j * 8192, &tmB, bres_bar, 64 * j, wtap, cb * kBlockK);
                }
              }
        } else {
          for (int r = 0; r < p.R; ++r)
            for (int ss = 0; ss < p.S; ++ss)
              for (int cb = 0; cb < p.CB; ++cb, dst += Cfg::kTileBBytes) {
                if (MODE == FPROP) {
                  tma_load_2d(dst, &tmB, bres_bar, (r * p.S + ss) * p.Ck + cb * kBlockK, 0);
                } else {
                  const int wtap = (p.wt_r0 + p.wt_rstep * (p.R - 1 - r)) * p.wt_S +
                                   (p.wt_s0 + p.wt_sstep * (p.S - 1 - ss));
#pragma unroll
                  for (int j = 0; j < BN / 64; ++j)
                    tma_load_3d(dst + j * 8192, &tmB, bres_bar, 64 * j, wtap, cb * kBlockK);
                }
              }
        }
      }
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int z = p.splits == 1 ? 0 : tile / tiles_mn;
        const int rem_t = tile - z * tiles_mn;
        const int n_blk = p.n_tiles == 1 ? 0 : rem_t / p.m_tiles;
        const int m0 = (rem_t - n_blk * p.m_tiles) * kBlockM;
        const int n0 = n_blk * BN;
        const int kb_begin = z * p.kb_per_split;
        int kb_end = kb_begin + p.kb_per_split;
        if (kb_end > p.num_kb) kb_end = p.num_kb;
        // im2col start coordinates of this tile's first pixel (FPROP/DGRAD)
        int px_n = 0, px_h = 0, px_w = 0;
        if (MODE == FPROP || MODE == DGRAD) {
          const int pq = p.P * p.Q;
          px_n = m0 / pq;
          const int rem = m0 - px_n * pq;
          const int pp = rem / p.Q;
          px_h = p.lower_h + pp * p.stride_h;
          px_w = p.lower_w + (rem - pp * p.Q) * p.stride_w;
        }
        // ---- per-tile constants and incremental counters (no divisions inside the K loop) ----
        // FPROP/DGRAD: kb -> (r, s, cb)
        int it_r = 0, it_s = 0, it_cb = 0;
        if (SV3) {  // K blocks are (filter row, channel block); the taps of a row share a box
          it_r = kb_begin / p.CB;
          it_cb = kb_begin - it_r * p.CB;
        } else if (MODE == FPROP || MODE == DGRAD) {
          const int tap = kb_begin / p.CB;
          it_cb = kb_begin - tap * p.CB;
          it_r = tap / p.S;
          it_s = tap - it_r * p.S;
        }
        // WGRAD: pixel cursor (n, row, col) of the K block, and the (tap, channel) of each
        // 64-wide column block of this n-tile
        int wg_n = 0, wg_p = 0, wg_q = 0;
        int wg_c0[BN / 64], wg_r[BN / 64], wg_s[BN / 64];
        if (MODE == WGRAD) {
          const int pix = kb_begin * kBlockK;
          const int pq = p.P * p.Q;
          wg_n = pix / pq;
          const int rem = pix - wg_n * pq;
          wg_p = rem / p.Q;
          wg_q = rem - wg_p * p.Q;
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) {
            int tap, c0;
            if (p.wg_flat) {
              const int nn = n0 + 64 * j;  // flattened (tap * C + c), C % 64 == 0
              tap = nn / p.Cn;
              c0 = nn - tap * p.Cn;
              if (tap >= p.R * p.S) {      // past the end: out-of-bounds channels read as zeros
                tap = 0;
                c0 = p.Cn;
              }
            } else {
              tap = n_blk / p.NCB;
              c0 = (n_blk - tap * p.NCB) * BN + 64 * j;
            }
            wg_c0[j] = c0;
            wg_r[j] = tap / p.S;
            wg_s[j] = tap - wg_r[j] * p.S;
          }
        }
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem_a + stage * kABytes;
          uint8_t* sb = smem_b + stage * kBBytes;
          mbar_expect_tx(&full_bar[stage], (SV3 ? (kBlockM + 2) * 128 : kTileABytes) +
                                               (BRES ? 0 : kBBytes));
          if (SV3) {
            // one box of 128 + 2 virtual pixels (tap offset 0 along w) for the filter row, and
            // the weight tiles of its three taps
            tma_load_im2col_4d(sa, &tmA, &full_bar[stage], it_cb * kBlockK, px_w, px_h, px_n,
                               static_cast<uint16_t>(0), static_cast<uint16_t>(it_r));
#pragma unroll
            for (int ts = 0; ts < 3; ++ts) {
              if (BRES) break;
              uint8_t* sbt = sb + ts * Cfg::kTileBBytes;
              if (MODE == FPROP) {
                tma_load_2d(sbt, &tmB, &full_bar[stage],
                            (it_r * p.S + ts) * p.Ck + it_cb * kBlockK, n0);
              } else {
                const int wtap = (p.wt_r0 + p.wt_rstep * (p.R - 1 - it_r)) * p.wt_S +
                                 (p.wt_s0 + p.wt_sstep * (p.S - 1 - ts));
#pragma unroll
                for (int j = 0; j < BN / 64; ++j)
                  tma_load_3d(sbt + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, wtap,
                              it_cb * kBlockK);
              }
            }
            if (++it_cb == p.CB) {
              it_cb = 0;
              ++it_r;
            }
          } else if (MODE == GEMM_TN) {
            tma_load_2d(sa, &tmA, &full_bar[stage], kb * kBlockK, m0);
            tma_load_2d(sb, &tmB, &full_bar[stage], kb * kBlockK, n0);
          } else if (MODE == GEMM_NN) {
            tma_load_2d(sa, &tmA, &full_bar[stage], kb * kBlockK, m0);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, kb * kBlockK);
          } else if (MODE == GEMM_NT) {
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j)
              tma_load_2d(sa + j * 8192, &tmA, &full_bar[stage], m0 + 64 * j, kb * kBlockK);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(sb + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, kb * kBlockK);
          } else if (MODE == FPROP || MODE == DGRAD) {
            tma_load_im2col_4d(sa, &tmA, &full_bar[stage], it_cb * kBlockK, px_w, px_h, px_n,
                               static_cast<uint16_t>(it_s), static_cast<uint16_t>(it_r));
            if (BRES) {
            } else if (MODE == FPROP) {
              tma_load_2d(sb, &tmB, &full_bar[stage],
                          (it_r * p.S + it_s) * p.Ck + it_cb * kBlockK, n0);
            } else {
              const int wtap = (p.wt_r0 + p.wt_rstep * (p.R - 1 - it_r)) * p.wt_S +
                               (p.wt_s0 + p.wt_sstep * (p.S - 1 - it_s));
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                tma_load_3d(sb + j * 8192, &tmB, &full_bar[stage], n0 + 64 * j, wtap,
                            it_cb * kBlockK);
            }
            if (++it_cb == p.CB) {
              it_cb = 0;
              if (++it_s == p.S) {
                it_s = 0;
                ++it_r;
              }
            }
          } else {  // WGRAD: K = pixels; N = (tap, channel) flattened or one tap per n-tile
            const int h = p.lower_h + wg_p * p.stride_h;
            const int w = p.lower_w + wg_q * p.stride_w;
#pragma unroll
            for (int j = 0; j < kBlockM / 64; ++j)
              tma_load_2d(sa + j * 8192, &tmA, &full_bar[stage], m0 + 64 * j, kb * kBlockK);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_im2col_4d(sb + j * 8192, &tmB, &full_bar[stage], wg_c0[j], w, h, wg_n,
                                 static_cast<uint16_t>(wg_s[j]), static_cast<uint16_t>(wg_r[j]));
            wg_q += kBlockK;  // advance the pixel cursor by one K block
            while (wg_q >= p.Q) {
              wg_q -= p.Q;
              if (++wg_p == p.P) {
                wg_p = 0;
                ++wg_n;
              }
            }
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // All 32 lanes walk the loops: with warp-uniform control flow the shared-memory descriptors
    // live in uniform registers and advance with one add per operand; one elected lane issues
    // the MMAs and the commits. (Issued from inside an `if (lane == 0)` region every
    // tcgen05.mma cost ~17 dependent instructions - 125 cycles, four times the duration of an
    // N = 64 MMA - and the tensor pipe idled 70 % of the time on narrow tiles.)
    {
      constexpr uint32_t idesc = make_idesc_f16(kBlockM, BN, kAMajor, kBMajor);
      // K-major: 8-row groups 1024 B apart; +32 B per UMMA_K(16) inside the swizzle atom.
      // MN-major: 64-element column groups 8192 B apart (LBO), 8-row K groups 1024 B apart
      //           (SBO); +2048 B per UMMA_K(16 rows).
      constexpr uint32_t a_lbo = (kAMajor == kMajorK) ? 16 : 8192;
      constexpr uint32_t b_lbo = (kBMajor == kMajorK) ? 16 : 8192;
      constexpr uint32_t a_kstep = (kAMajor == kMajorK) ? 32 : 2048;
      constexpr uint32_t b_kstep = (kBMajor == kMajorK) ? 32 : 2048;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      if (BRES && static_cast<int>(blockIdx.x) < total_tiles) mbar_wait(bres_bar, 0);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int z = p.splits == 1 ? 0 : tile / tiles_mn;
        const int kb_begin = z * p.kb_per_split;
        int kb_end = kb_begin + p.kb_per_split;
        if (kb_end > p.num_kb) kb_end = p.num_kb;
        const int acc = it % kAcc;
        const uint32_t acc_phase = (it / kAcc) & 1;
        mbar_wait(&tmem_empty_bar[acc], acc_phase ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = kb_begin; kb < kb_end; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * kABytes);
          const uint32_t b_addr = smem_u32(smem_b + (BRES ? kb : stage) * kBBytes);
          // descriptors of the stage; the start-address field (bits 0-13, units of 16 bytes)
          // takes the per-MMA offsets - it cannot carry out of the field below 256 KiB
          const uint64_t da0 = make_smem_desc_sw128(a_addr, a_lbo, 1024);
          const uint64_t db0 = make_smem_desc_sw128(b_addr, b_lbo, 1024);
          const uint32_t first = kb > kb_begin ? 1u : 0u;
          if (elect_one_sync()) {
            if (SV3) {
              // tap ts reads the box through a start address ts rows (128 B each) further on
#pragma unroll
              for (int ts = 0; ts < 3; ++ts) {
#pragma unroll
                for (int k = 0; k < kBlockK / 16; ++k)
                  umma_f16_ss(tmem_d, da0 + ((ts * 128 + k * a_kstep) >> 4),
                              db0 + ((ts * Cfg::kTileBBytes + k * b_kstep) >> 4), idesc,
                              (ts > 0 || k > 0) ? 1u : first);
              }
            } else {
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k)
                umma_f16_ss(tmem_d, da0 + ((k * a_kstep) >> 4), db0 + ((k * b_kstep) >> 4), idesc,
                            k > 0 ? 1u : first);
            }
            umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs retire
          }
          __syncwarp();
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (elect_one_sync()) umma_commit(&tmem_full_bar[acc]);
        __syncwarp();
      }
    }
  } else {
    // ===================== epilogue =====================
    // TMEM -> registers (thread = row) -> bias/ReLU -> per-warp staging tile in shared memory
    // -> global stores in which every instruction writes four full 128-byte row segments.
    const int quarter = warp & 3;                  // TMEM lanes this warp may read
    const int epi = warp - 2;                      // 0 .. kEpiWarps-1
    constexpr int kColSplit = Cfg::kEpiWarps / 4;  // warps sharing a lane quarter
    const int col_part = epi / 4;                  // which interleaved set of chunks is mine
    // TMA-store variant: dense 128-byte rows with the 128B swizzle of `tmC` (16-byte piece index
    // XOR row % 8), 4 KiB per warp, 1 KiB aligned (the staging area starts 1 KiB aligned)
    constexpr bool kTmaStoreMode = TMA_ST;
    constexpr bool tma_store = TMA_ST;
    bool store_pending = false;  // warp-uniform: a bulk store may still be reading stage_buf
    uint8_t* const stage_base = staging + epi * 8192;
    uint8_t* stage_buf = stage_base;
    constexpr bool kTwoStaging = TMA_ST && !ACC;   // the accumulating variant keeps one tile
    uint32_t chunk_no = 0;
    uint8_t* old_buf = old_area + epi * 8192;          // [2][4096] (ACC only)
    uint64_t* my_old_bar = old_bar + epi * 2;
    uint32_t old_issued = 0, old_used = 0;             // chunk loads issued / consumed so far
    const int esize = p.out_fp32 ? 4 : 2;
    const int cols_per_chunk = 128 / esize;  // 32 (fp32) or 64 (fp16)
    // ACC: a prefetch cursor runs ahead of the chunk being processed - also into the next tiles
    // of this CTA - so that the old values (TMA) and the mask words (one 8-byte load per lane)
    // of a chunk are requested two chunks before they are used; tiles of one K block (1x1
    // layers) have no main loop to hide that latency behind otherwise.
    int pf_tile = blockIdx.x, pf_k = 0, pf_it = 0;
    unsigned long long pf_mask0 = ~0ull, pf_mask1 = ~0ull;
    auto pump_old = [&]() {
      const int cstep = alt_tiles ? 64 : cols_per_chunk * kColSplit;
      const int cfirst = alt_tiles ? 0 : col_part * cols_per_chunk;
      while (old_issued - old_used < 2u && pf_tile < total_tiles) {
        // (the accumulating epilogue has no K splits)
        const int nb = p.n_tiles == 1 ? 0 : pf_tile / p.m_tiles;
        const int pm0 = (pf_tile - nb * p.m_tiles) * kBlockM;
        const int pcol0 = nb * BN;
        int lim = p.N - pcol0;
        if (lim > BN) lim = BN;
        int chunks = 0;
        if (pm0 + quarter * 32 < p.M && cfirst < lim) chunks = (lim - cfirst + cstep - 1) / cstep;
        if (alt_tiles && (pf_it & 1) != col_part) chunks = 0;  // the other warp's tile
        if (pf_k >= chunks) {
          pf_tile += gridDim.x;
          ++pf_it;
          pf_k = 0;
          continue;
        }
        const int slot = old_issued & 1;
        const int ccol = pcol0 + cfirst + pf_k * cstep;
        if (lane == 0) {
          mbar_expect_tx(&my_old_bar[slot], 4096);
          tma_load_2d(old_buf + slot * 4096, &tmD, &my_old_bar[slot], ccol, pm0 + quarter * 32);
        }
        unsigned long long mb = ~0ull;
        const int prow = pm0 + quarter * 32 + lane;
        if (p.acc_mask != nullptr && prow < p.M)
          mb = *reinterpret_cast<const unsigned long long*>(
              p.acc_mask + static_cast<long long>(prow) * p.acc_mask_ld + (ccol >> 3));
        if (slot) pf_mask1 = mb; else pf_mask0 = mb;
        ++old_issued;
        ++pf_k;
      }
    };
    // fused BN statistics: lane owns columns (2*lane, 2*lane+1) of each 64-wide chunk it handles
    constexpr int kStatChunks = (BN / 64) / kColSplit > 0 ? (BN / 64) / kColSplit : 1;
    float st_sum[kStatChunks][2], st_sq[kStatChunks][2];
#pragma unroll
    for (int c = 0; c < kStatChunks; ++c) st_sum[c][0] = st_sum[c][1] = st_sq[c][0] = st_sq[c][1] = 0.f;
    int st_nblk = -1;
    const bool do_stats = p.stats != nullptr && !p.out_fp32;
    auto flush_stats = [&](int nblk) {
      // stores this warp's sums of column block `nblk` into the row it owns (see GemmParams)
      const int first = (nblk * p.m_tiles) % static_cast<int>(gridDim.x);
      int slot = static_cast<int>(blockIdx.x) - first;
      if (slot < 0) slot += gridDim.x;
      const int sub = alt_tiles ? col_part * 4 + quarter : quarter;
      float* st_row = p.stats + (static_cast<long long>(nblk * p.stats_rpn + slot) *
                                     (alt_tiles ? 8 : 4) + sub) * 2 * BN;
#pragma unroll
      for (int c = 0; c < kStatChunks; ++c) {
        // column inside the tile
        const int col = (alt_tiles ? 0 : (col_part + c * kColSplit)) * 64 + 2 * lane;
        if (col < BN) {
          *reinterpret_cast<float2*>(st_row + col) = make_float2(st_sum[c][0], st_sum[c][1]);
          *reinterpret_cast<float2*>(st_row + BN + col) = make_float2(st_sq[c][0], st_sq[c][1]);
        }
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          st_sum[c][e] = 0.f;
          st_sq[c][e] = 0.f;
        }
      }
    };
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int z = p.splits == 1 ? 0 : tile / tiles_mn;
      const int rem_t = tile - z * tiles_mn;
      const int n_blk = p.n_tiles == 1 ? 0 : rem_t / p.m_tiles;
      if (do_stats && n_blk != st_nblk) {
        if (st_nblk >= 0) flush_stats(st_nblk);
        st_nblk = n_blk;
      }
      if (alt_tiles && (it & 1) != col_part) continue;  // the other warp of the quarter's tile
      const int m0 = (rem_t - n_blk * p.m_tiles) * kBlockM;
      int n0 = n_blk * BN;
      int col_mem0 = n0;       // first memory column of this tile
      int n_limit = p.N - n0;  // number of valid columns in this tile
      if (MODE == WGRAD && !p.wg_flat) {
        const int tap = n_blk / p.NCB;
        n0 = (n_blk - tap * p.NCB) * BN;
        col_mem0 = tap * p.Cn + n0;
        n_limit = p.Cn - n0;
      }
      if (n_limit > BN) n_limit = BN;
      const int acc = it % kAcc;
      const uint32_t acc_phase = (it / kAcc) & 1;
      const int row = m0 + quarter * 32 + lane;
      long long row_off = row;
      if (p.sc_P > 0) {
        // exact division by multiply-high (host-computed constants; row < 2^31)
        const int pq = p.sc_P * p.sc_Q;
        const int n_img = static_cast<int>(
            (static_cast<unsigned long long>(static_cast<unsigned>(row)) * p.sc_pq_mul) >> p.sc_pq_sh);
        const int rem = row - n_img * pq;
        const int pp = static_cast<int>(
            (static_cast<unsigned long long>(static_cast<unsigned>(rem)) * p.sc_q_mul) >> p.sc_q_sh);
        const int qq = rem - pp * p.sc_Q;
        row_off = (static_cast<long long>(n_img) * p.sc_H + p.sc_h0 +
                   static_cast<long long>(pp) * p.sc_sh) *
                      p.sc_W +
                  p.sc_w0 + static_cast<long long>(qq) * p.sc_sw;
        if (SV3 && qq == p.sc_Q - 1) row_off = -1;  // dummy column
      }
      if (row >= p.M) row_off = -1;  // marks an invalid row
      // SV3: the dummy rows hold real numbers (not zeros): they must stay out of the statistics
      const unsigned valid_rows = SV3 ? __ballot_sync(0xffffffffu, row_off >= 0) : 0xffffffffu;
      const long long base = static_cast<long long>(z) * p.split_stride + col_mem0;
      const int c_step = alt_tiles ? 64 : cols_per_chunk * kColSplit;
      const int c_first = alt_tiles ? 0 : col_part * cols_per_chunk;
      // rows of this warp beyond M are never stored: nothing is loaded for them either
      const bool acc_on = ACC && (m0 + quarter * 32 < p.M);
      if (ACC) pump_old();
      mbar_wait(&tmem_full_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t tmem_acc = tmem_base + acc * BN + (static_cast<uint32_t>(quarter * 32) << 16);
      if (c_first >= n_limit || c_first >= BN) {
        // nothing to read for this warp in this tile: release the accumulator right away
        if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
      }
#pragma unroll 1
      for (int c0 = c_first; c0 < BN; c0 += c_step) {
        if (c0 >= n_limit) break;  // warp-uniform
        const bool last_chunk = (c0 + c_step >= BN) || (c0 + c_step >= n_limit);
        // ---- registers <- TMEM, epilogue math, -> staging row of this thread ----
#pragma unroll 1
        for (int h = 0; h < cols_per_chunk / 32; ++h) {
          uint32_t v[32];
          const int cc = c0 + h * 32;
          tmem_ld_32x32b_x32(tmem_acc + cc, v);
          tmem_ld_wait();
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (p.bias != nullptr) {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (FAST || cc + j < n_limit) f[j] += __ldg(p.bias + n0 + cc + j);
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.0f);
          }
          if (kTmaStoreMode && tma_store) {
            if (h == 0) {
              if (kTwoStaging) stage_buf = stage_base + (chunk_no & 1) * 4096;
              ++chunk_no;
              if (store_pending) {
                // the bulk store that last read this staging tile (the previous chunk's, or with
                // two tiles the one before it) must be done reading before it is overwritten
                if (lane == 0) {
                  if (kTwoStaging) bulk_wait_group_read_1(); else bulk_wait_group_read_all();
                }
                __syncwarp();
              }
            }
            uint8_t* srow = stage_buf + lane * 128;
            const uint8_t* orow = old_buf + (old_used & 1) * 4096 + lane * 128;
            if (acc_on && h == 0) {
              // the old values of this chunk have landed (phase = use count of the buffer / 2)
              mbar_wait(&my_old_bar[old_used & 1], (old_used >> 1) & 1);
            }
            // masked source: the 64 mask bits of this thread's row for the chunk (requested
            // together with the old values)
            const unsigned long long mbits = (old_used & 1) ? pf_mask1 : pf_mask0;
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              const int piece = h * 4 + (j >> 3);
              uint4 u;
              u.x = pack_half2(f[j], f[j + 1]);
              u.y = pack_half2(f[j + 2], f[j + 3]);
              u.z = pack_half2(f[j + 4], f[j + 5]);
              u.w = pack_half2(f[j + 6], f[j + 7]);
              if (acc_on) {
                // out = fp16(result) + old * mask, packed (the sum of two fp16 tensors, as the
                // reference's add_n forms it). The eight mask bits become four half2 lane masks:
                // a multiply spreads a nibble to the sign bits of four bytes, PRMT replicates
                // the signs - 18 instructions per eight elements instead of 40 scalar ones.
                const uint4 o = *reinterpret_cast<const uint4*>(orow + ((piece ^ (lane & 7)) << 4));
                const unsigned mb = static_cast<unsigned>(mbits >> (8 * piece)) & 0xffu;
                const unsigned sl = ((mb & 0xfu) * 0x10204080u) & 0x80808080u;
                const unsigned sh = ((mb >> 4) * 0x10204080u) & 0x80808080u;
                u.x = hadd2_u32(u.x, o.x & prmt_sign(sl, 0x9988u));
                u.y = hadd2_u32(u.y, o.y & prmt_sign(sl, 0xBBAAu));
                u.z = hadd2_u32(u.z, o.z & prmt_sign(sh, 0x9988u));
                u.w = hadd2_u32(u.w, o.w & prmt_sign(sh, 0xBBAAu));
              }
              *reinterpret_cast<uint4*>(srow + ((piece ^ (lane & 7)) << 4)) = u;
            }
            continue;
          }
          uint8_t* srow = stage_buf + lane * kStageRowBytes;
          if (p.out_fp32) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              *reinterpret_cast<float4*>(srow + j * 4) =
                  make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              uint4 u;
              u.x = pack_half2(f[j], f[j + 1]);
              u.y = pack_half2(f[j + 2], f[j + 3]);
              u.z = pack_half2(f[j + 4], f[j + 5]);
              u.w = pack_half2(f[j + 6], f[j + 7]);
              *reinterpret_cast<uint4*>(srow + h * 64 + j * 2) = u;
            }
          }
        }
        if (last_chunk) {
          // this warp has read everything it needs from the accumulator: hand it back to the
          // MMA warp before the (slow) global stores
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
        }
        __syncwarp();
        if (do_stats) {
          // column sums of the staged (already fp16-rounded) 32 x 64 chunk; rows beyond M and
          // columns beyond N hold exact zeros (zero-filled operands), so no masking is needed
          const int sc = alt_tiles ? 0 : (c0 / 64 - col_part) / kColSplit;
          float a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            if (SV3 && !((valid_rows >> r) & 1u)) continue;
            // the lane's two columns of row r (swizzled 16-byte pieces in the TMA-store layout)
            const uint8_t* sp =
                (kTmaStoreMode && tma_store)
                    ? stage_buf + r * 128 + ((((lane >> 2) ^ (r & 7)) << 4) | ((lane & 3) << 2))
                    : stage_buf + r * kStageRowBytes + lane * 4;
            const float2 v = __half22float2(*reinterpret_cast<const __half2*>(sp));
            a0 += v.x;
            a1 += v.y;
            q0 = fmaf(v.x, v.x, q0);
            q1 = fmaf(v.y, v.y, q1);
          }
#pragma unroll
          for (int c = 0; c < kStatChunks; ++c)
            if (c == sc) {
              st_sum[c][0] += a0;
              st_sum[c][1] += a1;
              st_sq[c][0] += q0;
              st_sq[c][1] += q1;
            }
        }
        if (kTmaStoreMode && tma_store) {
          // generic-proxy writes of every lane -> visible to the async proxy, then one lane
          // stores the 32-row x 64-column box; rows >= M and columns >= N are clipped by TMA
          fence_proxy_async_smem();
          __syncwarp();
          if (acc_on) {
            // every lane has read the old tile (ordered by the fence + syncwarp above): its
            // buffer takes the old values of the chunk after next (of this or a later tile)
            ++old_used;
            pump_old();
          }
          if (lane == 0 && m0 + quarter * 32 < p.M) {
            tma_store_2d(&tmC, smem_u32(stage_buf), col_mem0 + c0, m0 + quarter * 32);
            bulk_commit_group();
          }
          store_pending = true;
          continue;
        }
        // ---- staging -> global: lane handles 16-byte piece (lane % 8) of row i8*4 + lane/8.
        //      All old-value loads of an accumulate are issued before the first store so eight
        //      requests per lane are in flight. ----
        const int piece = lane & 7;
        const int epp = 16 / esize;  // elements per piece
        const int pcol = c0 + piece * epp;
        const bool col_ok = pcol < n_limit;
        const bool col_full = pcol + epp <= n_limit;
        uint8_t* gps[8];
        bool vec_ok[8], any_ok[8];
        uint4 olds[8];
#pragma unroll
        for (int i8 = 0; i8 < 8; ++i8) {
          const int r = i8 * 4 + (lane >> 3);
          const long long ro = __shfl_sync(0xffffffffu, row_off, r);
          const long long e_idx = base + ro * p.ldc + pcol;
          uint8_t* obase = reinterpret_cast<uint8_t*>(p.out);
          if (kCanPush && p.push_n > 1) {
            int owner = 0;
#pragma unroll
            for (int r = 1; r < 8; ++r) owner += (r < p.push_n && e_idx >= p.push_bound[r]) ? 1 : 0;
            obase = reinterpret_cast<uint8_t*>(p.push_base[owner]);
          }
          gps[i8] = obase + e_idx * esize;
          any_ok[i8] = ro >= 0 && col_ok;
          vec_ok[i8] = FAST ? any_ok[i8]
                            : (any_ok[i8] && col_full &&
                               ((reinterpret_cast<uintptr_t>(gps[i8]) & 15) == 0));
        }
        if (p.accumulate) {
#pragma unroll
          for (int i8 = 0; i8 < 8; ++i8)
            if (vec_ok[i8]) olds[i8] = *reinterpret_cast<const uint4*>(gps[i8]);
        }
#pragma unroll
        for (int i8 = 0; i8 < 8; ++i8) {
          if (!any_ok[i8]) continue;
          const int r = i8 * 4 + (lane >> 3);
          uint4 val = *reinterpret_cast<const uint4*>(stage_buf + r * kStageRowBytes + piece * 16);
          if (vec_ok[i8]) {
            if (p.accumulate) {
              if (p.out_fp32) {
                float4* v4 = reinterpret_cast<float4*>(&val);
                const float4* o4 = reinterpret_cast<const float4*>(&olds[i8]);
                v4->x += o4->x; v4->y += o4->y; v4->z += o4->z; v4->w += o4->w;
              } else {
                __half2* a2 = reinterpret_cast<__half2*>(&val);
                const __half2* o2 = reinterpret_cast<const __half2*>(&olds[i8]);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float2 x = __half22float2(a2[j]), y = __half22float2(o2[j]);
                  a2[j] = __floats2half2_rn(x.x + y.x, x.y + y.y);
                }
              }
            }
            *reinterpret_cast<uint4*>(gps[i8]) = val;
          } else if (FAST) {
            // unreachable: FAST launches only see full aligned pieces
          } else if (p.out_fp32) {
            const float* e = reinterpret_cast<const float*>(&val);
            float* go = reinterpret_cast<float*>(gps[i8]);
            for (int j = 0; j < 4; ++j)
              if (pcol + j < n_limit) go[j] = p.accumulate ? go[j] + e[j] : e[j];
          } else {
            const __half* e = reinterpret_cast<const __half*>(&val);
            __half* go = reinterpret_cast<__half*>(gps[i8]);
            for (int j = 0; j < 8; ++j)
              if (pcol + j < n_limit)
                go[j] = p.accumulate ? __float2half_rn(__half2float(go[j]) + __half2float(e[j]))
                                     : e[j];
          }
        }
        __syncwarp();
      }
    }
    if (do_stats && st_nblk >= 0) flush_stats(st_nblk);
    if (kTmaStoreMode && tma_store && store_pending && lane == 0) bulk_wait_group_all();
    // remote (NVLink) stores of a fused reduce-scatter must be performed system-wide before
    // this kernel counts as complete for the ranks that wait on the stream-ordered READY flag
    if (kCanPush && p.push_n > 1) __threadfence_system();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---------------------------------------------------------------------------
// Deterministic split-K reduction: out[i] = (accumulate ? out[i] : 0) + sum_z ws[z][i]
// ---------------------------------------------------------------------------
// Block = (32 float4 columns, kSplitLanes split lanes): lane y sums slices y, y+L, ... so the
// loads of a thin output (a 1x1 conv's few thousand weights) still fill the memory pipeline;
// the lane sums are combined in lane order, so the result does not depend on the launch.
constexpr int kSplitLanes = 8;
// launch shape: as many split lanes as there are splits (up to 8), 256 threads per block
static inline void splitk_launch_shape(long long n4, int splits, dim3* grid, dim3* block) {
  int lanes = 1;
  while (lanes < kSplitLanes && lanes < splits) lanes <<= 1;
  const int bx = 256 / lanes;
  *block = dim3(bx, lanes);
  *grid = dim3((unsigned)((n4 + bx - 1) / bx));
}
__global__ void __launch_bounds__(256)
    splitk_reduce_kernel(const float* __restrict__ ws, float* __restrict__ out, long long n,
                         int splits, int accumulate, float scale) {
  __shared__ float4 sh[256];  // [blockDim.y lanes][blockDim.x columns], 256 threads in all
  const long long i4 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  const bool full = i4 + 4 <= n;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i4 < n) {
    for (int z = threadIdx.y; z < splits; z += blockDim.y) {
      const float* p = ws + static_cast<long long>(z) * n + i4;
      if (full) {
        const float4 v = *reinterpret_cast<const float4*>(p);
        acc.x += v.x;
        acc.y += v.y;
        acc.z += v.z;
        acc.w += v.w;
      } else {
        acc.x += p[0];
        if (i4 + 1 < n) acc.y += p[1];
        if (i4 + 2 < n) acc.z += p[2];
      }
    }
  }
  sh[threadIdx.y * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y != 0 || i4 >= n) return;
  for (int l = 1; l < static_cast<int>(blockDim.y); ++l) {
    const float4 v = sh[l * blockDim.x + threadIdx.x];
    acc.x += v.x;
    acc.y += v.y;
    acc.z += v.z;
    acc.w += v.w;
  }
  acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
  if (full) {
    float4* o = reinterpret_cast<float4*>(out + i4);
    if (accumulate) {
      const float4 old = *o;
      acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
    }
    *o = acc;
  } else {
    const float e[3] = {acc.x, acc.y, acc.z};
    for (int j = 0; i4 + j < n; ++j) out[i4 + j] = accumulate ? out[i4 + j] + e[j] : e[j];
  }
}

// Split-K reduction whose result is scattered to the owning ranks' staging slots (the second
// stage of a fused wgrad + reduce-scatter when the GEMM needed more than one K split).
struct PushTarget {
  int n;
  long long bound[8];
  float* base[8];
};

__global__ void __launch_bounds__(256)
    splitk_reduce_push_kernel(const float* __restrict__ ws, PushTarget t, long long n, int splits,
                              float scale) {
  __shared__ float4 sh[256];  // [blockDim.y lanes][blockDim.x columns], 256 threads in all
  const long long i4 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  const bool full = i4 + 4 <= n;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i4 < n) {
    for (int z = threadIdx.y; z < splits; z += blockDim.y) {
      const float* p = ws + static_cast<long long>(z) * n + i4;
      if (full) {
        const float4 v = *reinterpret_cast<const float4*>(p);
        acc.x += v.x;
        acc.y += v.y;
        acc.z += v.z;
        acc.w += v.w;
      } else {
        acc.x += p[0];
        if (i4 + 1 < n) acc.y += p[1];
        if (i4 + 2 < n) acc.z += p[2];
      }
    }
  }
  sh[threadIdx.y * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y != 0 || i4 >= n) return;
  for (int l = 1; l < static_cast<int>(blockDim.y); ++l) {
    const float4 v = sh[l * blockDim.x + threadIdx.x];
    acc.x += v.x;
    acc.y += v.y;
    acc.z += v.z;
    acc.w += v.w;
  }
  acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
  int owner = 0;
#pragma unroll
  for (int r = 1; r < 8; ++r) owner += (r < t.n && i4 >= t.bound[r]) ? 1 : 0;
  float* out = t.base[owner];
  if (full) {
    *reinterpret_cast<float4*>(out + i4) = acc;
  } else {
    const float e[3] = {acc.x, acc.y, acc.z};
    for (int j = 0; i4 + j < n; ++j) out[i4 + j] = e[j];
  }
  __threadfence_system();
}

// ---------------------------------------------------------------------------
// Host: tensor-map encoding through the driver entry point (no -lcuda needed)
// ---------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*PFN_encodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                     const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                     cuuint32_t, cuuint32_t, const cuuint32_t*,
                                     CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled g_encodeTiled = nullptr;
static PFN_encodeIm2col g_encodeIm2col = nullptr;

static int load_driver_fns() {
  if (g_encodeTiled && g_encodeIm2col) return 0;
  cudaDriverEntryPointQueryResult q;
  void* fn = nullptr;
  KFB_CHECK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  KFB_REQUIRE(fn != nullptr && q == cudaDriverEntryPointSuccess,
              "cuTensorMapEncodeTiled not available from the driver");
  g_encodeTiled = reinterpret_cast<PFN_encodeTiled>(fn);
  fn = nullptr;
  KFB_CHECK_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &q));
  KFB_REQUIRE(fn != nullptr && q == cudaDriverEntryPointSuccess,
              "cuTensorMapEncodeIm2col not available from the driver");
  g_encodeIm2col = reinterpret_cast<PFN_encodeIm2col>(fn);
  return 0;
}

// Tensor maps are pure functions of (pointer, geometry): a training step launches the same few
// hundred GEMMs on the same static buffers every step, so the encoded maps are kept in a small
// host-side table and a launch outside a CUDA graph costs one hash lookup per operand instead of
// a driver call (KFB_TMAP_CACHE=0 encodes every time).
struct TmapKey {
  uint64_t v[20];
  bool operator==(const TmapKey& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    uint64_t h = 1469598103934665603ull;
    for (uint64_t x : k.v) h = (h ^ x) * 1099511628211ull;
    return (size_t)h;
  }
};
static std::mutex g_tmap_mu;
static std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> g_tmap_cache;
static bool tmap_cache_on() {
  static const bool on = [] {
    const char* e = getenv("KFB_TMAP_CACHE");
    return !(e != nullptr && *e != 0 && atoi(e) == 0);
  }();
  return on;
}
static bool tmap_lookup(const TmapKey& k, CUtensorMap* m) {
  if (!tmap_cache_on()) return false;
  std::lock_guard<std::mutex> lock(g_tmap_mu);
  auto it = g_tmap_cache.find(k);
  if (it == g_tmap_cache.end()) return false;
  *m = it->second;
  return true;
}
static void tmap_store(const TmapKey& k, const CUtensorMap& m) {
  if (!tmap_cache_on()) return;
  std::lock_guard<std::mutex> lock(g_tmap_mu);
  if (g_tmap_cache.size() > 8192) g_tmap_cache.clear();
  g_tmap_cache.emplace(k, m);
}

// fp16 tensor, rank 2 or 3 (dims innermost first), 128B swizzle, zero OOB fill.
static int encode_tiled(CUtensorMap* m, const void* ptr, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes /* rank-1 */, const uint32_t* box) {
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.v[0] = 1;
  key.v[1] = (uint64_t)(uintptr_t)ptr;
  key.v[2] = (uint64_t)rank;
  for (int i = 0; i < rank; ++i) {
    key.v[3 + i] = dims[i];
    key.v[8 + i] = box[i];
  }
  for (int i = 0; i + 1 < rank; ++i) key.v[13 + i] = strides_bytes[i];
  if (tmap_lookup(key, m)) return 0;
  CUresult r = g_encodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, rank, const_cast<void*>(ptr),
                             gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  KFB_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeTiled failed (%d): rank %d dims [%llu,%llu,%llu] box [%u,%u,%u] "
              "stride0 %llu ptr %p",
              (int)r, rank, (unsigned long long)dims[0], (unsigned long long)dims[1],
              (unsigned long long)(rank > 2 ? dims[2] : 0), box[0], box[1], rank > 2 ? box[2] : 0,
              (unsigned long long)strides_bytes[0], ptr);
  tmap_store(key, *m);
  return 0;
}

// NHWC fp16 activation viewed as {C, W, H, N}; im2col traversal.
static int encode_im2col(CUtensorMap* m, const void* ptr, int N, int H, int W, int C, int lower_w,
                         int lower_h, int upper_w, int upper_h, int channels_per_pixel,
                         int pixels_per_column, int stride_w, int stride_h, int pix_stride = 0,
                         int row_stride = 0) {
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  const cuuint64_t ps = pix_stride > 0 ? (cuuint64_t)pix_stride : (cuuint64_t)C;
  const cuuint64_t rs = row_stride > 0 ? (cuuint64_t)row_stride : (cuuint64_t)W * ps;
  cuuint64_t gstr[3] = {ps * 2, rs * 2, (cuuint64_t)H * rs * 2};
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t es[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)stride_h, 1};
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.v[0] = 2;
  key.v[1] = (uint64_t)(uintptr_t)ptr;
  key.v[2] = ((uint64_t)(uint32_t)N << 32) | (uint32_t)H;
  key.v[3] = ((uint64_t)(uint32_t)W << 32) | (uint32_t)C;
  key.v[4] = ((uint64_t)(uint32_t)lower_w << 32) | (uint32_t)lower_h;
  key.v[5] = ((uint64_t)(uint32_t)upper_w << 32) | (uint32_t)upper_h;
  key.v[6] = ((uint64_t)(uint32_t)channels_per_pixel << 32) | (uint32_t)pixels_per_column;
  key.v[7] = ((uint64_t)(uint32_t)stride_w << 32) | (uint32_t)stride_h;
  key.v[8] = ((uint64_t)(uint32_t)pix_stride << 32) | (uint32_t)row_stride;
  if (tmap_lookup(key, m)) return 0;
  CUresult r = g_encodeIm2col(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(ptr), gdim,
                              gstr, lower, upper, (cuuint32_t)channels_per_pixel,
                              (cuuint32_t)pixels_per_column, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  KFB_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeIm2col failed (%d): NHWC [%d,%d,%d,%d] lower [%d,%d] upper "
              "[%d,%d] cpp %d ppc %d stride [%d,%d]",
              (int)r, N, H, W, C, lower_w, lower_h, upper_w, upper_h, channels_per_pixel,
              pixels_per_column, stride_w, stride_h);
  tmap_store(key, *m);
  return 0;
}

static int g_sm_count = 0;
static int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

// n / d == (n * mul) >> sh for 0 <= n < 2^31: mul = floor(2^(32+s) / d) + 1, s = floor(log2 d)
static void magic_div(unsigned d, unsigned long long* mul, int* sh) {
  int s = 0;
  while (s < 30 && (2u << s) <= d) ++s;
  *sh = 32 + s;
  *mul = ((1ull << (32 + s)) / d) + 1;
}
static void fill_scatter_magic(GemmParams* p) {
  if (p->sc_P <= 0) return;
  magic_div((unsigned)(p->sc_P * p->sc_Q), &p->sc_pq_mul, &p->sc_pq_sh);
  magic_div((unsigned)p->sc_Q, &p->sc_q_mul, &p->sc_q_sh);
}

// KFB_BRES=0 switches the resident-weights variants off
static int bres_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("KFB_BRES");
    v = (e != nullptr && *e != 0 && atoi(e) == 0) ? 0 : 1;
  }
  return v;
}

template <int BN, int MODE>
static int launch_cfg(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams p, dim3 tiles,
                      cudaStream_t stream) {
  using Cfg = TileCfg<BN>;
  static bool attr_set[16] = {false};  // per device
  int dev = 0;
  KFB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 16 && !attr_set[dev]) {
    KFB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, MODE, false>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::smem_bytes(Cfg::max_stages())));
    KFB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, MODE, true>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::smem_bytes(Cfg::max_stages())));
    if constexpr (MODE == GEMM_TN || MODE == FPROP || MODE == DGRAD) {
      KFB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, MODE, true, true>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          Cfg::smem_bytes(Cfg::max_stages())));
    }
    if constexpr (MODE == DGRAD) {
      KFB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, MODE, true, true, true>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          Cfg::smem_bytes_acc(Cfg::max_stages_acc())));
    }
    if constexpr (BN == 64 && (MODE == FPROP || MODE == DGRAD)) {
      KFB_CHECK_CUDA(cudaFuncSetAttribute(
          gemm_tcgen05_kernel<BN, MODE, true, true, false, false, true>,
          cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget));
    }
    attr_set[dev] = true;
  }
  // every 16-byte piece of the output is full and aligned?
  const int esz = p.out_fp32 ? 4 : 2;
  const int epp = 16 / esz;
  const int n_valid = (MODE == WGRAD && !p.wg_flat) ? p.Cn : p.N;
  const bool fast = (n_valid % epp == 0) && (p.ldc % epp == 0) && (p.split_stride % epp == 0) &&
                    ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0) &&
                    (MODE != WGRAD || p.wg_flat || (p.Cn % epp == 0));
  // TMA-store epilogue: the default where it is eligible (measured +0.9 % on the headline step,
  // +2.3 % together with the fused BN statistics, profiles/README.md); KFB_TMA_STORE=0 disables
  static int tma_store_env = -1;
  if (tma_store_env < 0) {
    const char* e = getenv("KFB_TMA_STORE");
    tma_store_env = (e != nullptr && *e != 0 && atoi(e) == 0) ? 0 : 1;
  }
  CUtensorMap tmC = tmA;  // placeholder, only dereferenced when p.tma_store is set
  p.tma_store = 0;
  static int tma_acc_env = -1;   // KFB_TMA_ACC=0: accumulating outputs keep the per-lane epilogue
  if (tma_acc_env < 0) {
    const char* e = getenv("KFB_TMA_ACC");
    tma_acc_env = (e != nullptr && *e != 0 && atoi(e) == 0) ? 0 : 1;
  }
  if (p.acc_src != nullptr) p.accumulate = 1;  // the masked source takes the accumulate path
  const bool acc_ok = !p.accumulate || ((tma_acc_env == 1 || p.acc_src != nullptr) && MODE == DGRAD);
  if (tma_store_env == 1 && fast && (MODE == GEMM_TN || MODE == FPROP || MODE == DGRAD) &&
      !p.out_fp32 && acc_ok && p.sc_P == 0 && tiles.z == 1 && p.push_n <= 1 &&
      p.N >= 64 && p.M >= 32 && (p.ldc % 8) == 0) {
    const uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.M};
    const uint64_t strides[1] = {(uint64_t)p.ldc * 2};
    const uint32_t box[2] = {64, 32};
    if (encode_tiled(&tmC, p.out, 2, dims, strides, box)) return 1;
    p.tma_store = 1;
  }
  CUtensorMap tmD = tmC;  // source of the values an accumulating epilogue adds (default: out)
  if (p.acc_src != nullptr) {
    KFB_REQUIRE(p.tma_store == 1,
                "fused residual gradient: this shape does not take the TMA-store epilogue");
    const uint64_t dims[2] = {(uint64_t)p.N, (uint64_t)p.M};
    const uint64_t strides[1] = {(uint64_t)p.acc_src_ld * 2};
    const uint32_t box[2] = {64, 32};
    if (encode_tiled(&tmD, p.acc_src, 2, dims, strides, box)) return 1;
  }
  p.stages = Cfg::max_stages();
  fill_scatter_magic(&p);
  p.m_tiles = (int)tiles.x;
  p.n_tiles = (int)tiles.y;
  p.splits = (int)tiles.z;
  const long long total = (long long)tiles.x * tiles.y * tiles.z;
  KFB_REQUIRE(total > 0 && total < (1ll << 31), "bad tile count %lld", total);
  const int ctas = total < sm_count() ? (int)total : sm_count();
  p.stats_rpn = (int)tiles.x < ctas ? (int)tiles.x : ctas;
  if (fast && p.tma_store && p.accumulate) {
    if constexpr (MODE == DGRAD) {
      p.stages = Cfg::max_stages_acc();
      gemm_tcgen05_kernel<BN, MODE, true, true, true>
          <<<ctas, Cfg::kThreads, Cfg::smem_bytes_acc(p.stages), stream>>>(tmA, tmB, tmC, tmD, p);
    }
  } else if (BN == 64 && (MODE == FPROP || MODE == DGRAD) && fast && p.tma_store &&
             bres_enabled() && tiles.y == 1 && tiles.z == 1 &&
             p.num_kb * Cfg::kTileBBytes <= Cfg::kBresMax) {
    if constexpr (BN == 64 && (MODE == FPROP || MODE == DGRAD)) {
      p.bres_bytes = p.num_kb * Cfg::kTileBBytes;
      p.stages = Cfg::stages_bres(kTileABytes, p.bres_bytes);
      gemm_tcgen05_kernel<BN, MODE, true, true, false, false, true>
          <<<ctas, Cfg::kThreads, Cfg::smem_bytes_bres(kTileABytes, p.bres_bytes, p.stages),
             stream>>>(tmA, tmB, tmC, tmD, p);
    }
  } else if (fast && p.tma_store) {
    if constexpr (MODE == GEMM_TN || MODE == FPROP || MODE == DGRAD) {
      gemm_tcgen05_kernel<BN, MODE, true, true>
          <<<ctas, Cfg::kThreads, Cfg::smem_bytes(p.stages), stream>>>(tmA, tmB, tmC, tmD, p);
    }
  } else if (fast)
    gemm_tcgen05_kernel<BN, MODE, true><<<ctas, Cfg::kThreads, Cfg::smem_bytes(p.stages), stream>>>(
        tmA, tmB, tmC, tmD, p);
  else
    gemm_tcgen05_kernel<BN, MODE, false><<<ctas, Cfg::kThreads, Cfg::smem_bytes(p.stages), stream>>>(
        tmA, tmB, tmC, tmD, p);
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// SV3 launch (FPROP / DGRAD, BN <= 128): per-lane epilogue (the virtual rows scatter to real
// pixels), the deepest pipeline its 17 KiB + 3 B-tile stages allow.
template <int BN, int MODE>
static int launch_sv3_cfg(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams p, dim3 tiles,
                          cudaStream_t stream) {
  using Cfg = TileCfg<BN>;
  static bool attr_set[16] = {false};  // per device
  int dev = 0;
  KFB_CHECK_CUDA(cudaGetDevice(&dev));
  if (dev < 16 && !attr_set[dev]) {
    KFB_CHECK_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, MODE, true, false, false, true>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::smem_bytes_sv3(Cfg::max_stages_sv3())));
    if constexpr (BN == 64) {
      KFB_CHECK_CUDA(cudaFuncSetAttribute(
          gemm_tcgen05_kernel<BN, MODE, true, false, false, true, true>,
          cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget));
    }
    attr_set[dev] = true;
  }
  p.tma_store = 0;
  fill_scatter_magic(&p);
  p.stages = Cfg::max_stages_sv3();
  p.m_tiles = (int)tiles.x;
  p.n_tiles = (int)tiles.y;
  p.splits = 1;
  const long long total = (long long)tiles.x * tiles.y;
  KFB_REQUIRE(total > 0 && total < (1ll << 31), "bad tile count %lld", total);
  const int ctas = total < sm_count() ? (int)total : sm_count();
  p.stats_rpn = (int)tiles.x < ctas ? (int)tiles.x : ctas;
  if (BN == 64 && bres_enabled() && tiles.y == 1 &&
      p.num_kb * 3 * Cfg::kTileBBytes <= Cfg::kBresMax) {
    if constexpr (BN == 64) {
      p.bres_bytes = p.num_kb * 3 * Cfg::kTileBBytes;
      p.stages = Cfg::stages_bres(Cfg::kSv3ABytes, p.bres_bytes);
      gemm_tcgen05_kernel<BN, MODE, true, false, false, true, true>
          <<<ctas, Cfg::kThreads, Cfg::smem_bytes_bres(Cfg::kSv3ABytes, p.bres_bytes, p.stages),
             stream>>>(tmA, tmB, tmA, tmA, p);
    }
  } else {
    gemm_tcgen05_kernel<BN, MODE, true, false, false, true>
        <<<ctas, Cfg::kThreads, Cfg::smem_bytes_sv3(p.stages), stream>>>(tmA, tmB, tmA, tmA, p);
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int MODE>
static int launch_sv3(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                      dim3 grid, cudaStream_t stream) {
  if (bn == 64) return launch_sv3_cfg<64, MODE>(tmA, tmB, p, grid, stream);
  if (bn == 128) return launch_sv3_cfg<128, MODE>(tmA, tmB, p, grid, stream);
  set_error("SV3: unsupported BN %d", bn);
  return 2;
}

template <int MODE>
static int launch_mode(int bn, const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p,
                       dim3 grid, cudaStream_t stream) {
  switch (bn) {
    case 64: return launch_cfg<64, MODE>(tmA, tmB, p, grid, stream);
    case 128: return launch_cfg<128, MODE>(tmA, tmB, p, grid, stream);
    case 256: return launch_cfg<256, MODE>(tmA, tmB, p, grid, stream);
  }
  set_error("unsupported BN %d", bn);
  return 2;
}

// Tile width: the widest tile that covers N (wider tiles amortise the A-operand smem reads);
// short-K, wide-N problems keep 128-wide tiles so more tiles exist to balance across SMs.
// Policy knobs for on-device sweeps without a rebuild (read once; defaults = measured setup):
// KFB_GEMM_MAX_BN (64 | 128 | 256, default 256) caps the tile width, KFB_SPLIT_ROUNDS (default 1;
// 2-4 measured 1-4 % slower on the headline step) is the number of persistent-grid rounds the
// split-K count aims for.
static int env_knob(const char* name, int dflt, int lo, int hi) {
  const char* e = getenv(name);
  if (e == nullptr || *e == 0) return dflt;
  const int v = atoi(e);
  return v < lo ? lo : (v > hi ? hi : v);
}
static int max_bn() {
  static const int v = env_knob("KFB_GEMM_MAX_BN", 256, 64, 256);
  return v >= 256 ? 256 : (v >= 128 ? 128 : 64);
}
static int split_rounds() {
  static const int v = env_knob("KFB_SPLIT_ROUNDS", 1, 1, 8);
  return v;
}
static int pick_bn(int n) {
  const int bn = n <= 64 ? 64 : (n <= 128 ? 128 : 256);
  return bn < max_bn() ? bn : max_bn();
}
static int pick_bn_for(int n, int num_kb) {
  int bn = pick_bn(n);
  if (num_kb <= 2 && bn == 256) bn = 128;
  return bn;
}

// Split count such that tiles * splits fills k whole rounds of the persistent grid (k = 2 when
// possible): a partial last round would leave most SMs idle for a full tile time.
static int balanced_splits(int tiles, int num_kb, int sms) {
  int splits = (split_rounds() * sms) / tiles;
  if (splits < 1) splits = 1;
  if (splits > num_kb) splits = num_kb;
  return splits < 1 ? 1 : splits;
}
static void init_params(GemmParams* p) { memset(p, 0, sizeof(*p)); }

// SV3 eligibility (see the kernel): three taps per filter row at unit stride with one padding
// column on each side, whole 64-channel K blocks, 16-byte output pieces, tiles of at most 128
// columns (three B tiles per stage must fit). KFB_SV3=0 switches the variant off.
static int sv3_enabled() {
  static const int v = env_knob("KFB_SV3", 1, 0, 1);
  return v;
}
static bool sv3_fprop_ok(const kfb_conv_desc* d, int bn, long long ldy, const void* y) {
  return sv3_enabled() && d->s == 3 && d->stride_h == 1 && d->stride_w == 1 &&
         d->pad_w_lo == 1 && d->pad_w_hi == 1 && d->q == d->w && d->c % 64 == 0 && bn <= 128 &&
         d->k % 8 == 0 && ldy % 8 == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0;
}
static bool sv3_dgrad_ok(const kfb_conv_desc* d, int bn, const void* dx) {
  return sv3_enabled() && d->s == 3 && d->stride_h == 1 && d->stride_w == 1 &&
         d->pad_w_lo == 1 && d->pad_w_hi == 1 && d->q == d->w && d->k % 64 == 0 && bn <= 128 &&
         d->c % 8 == 0 && (reinterpret_cast<uintptr_t>(dx) & 15) == 0;
}

}  // namespace kfb

using namespace kfb;

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" const char* kfb_last_error(void) { return kfb::g_err; }

// Layout of the fused statistics a GEMM with M rows, N columns and num_kb 64-wide K blocks
// writes: `rows` partial rows of [2][bn] floats per bn-wide column block.
extern "C" int kfb_gemm_stats_layout(int M, int N, int num_kb, int* rows, int* bn) {
  KFB_REQUIRE(M > 0 && N > 0 && num_kb > 0 && rows != nullptr && bn != nullptr,
              "gemm_stats_layout: bad arguments");
  const int b = pick_bn_for(N, num_kb);
  const long long m_tiles = (M + 127) / 128, n_tiles = (N + b - 1) / b;
  const long long total = m_tiles * n_tiles;
  const long long ctas = total < sm_count() ? total : sm_count();
  // 64-wide tiles: the two epilogue warps of a lane quarter take alternate tiles, a row each
  *rows = (int)(m_tiles < ctas ? m_tiles : ctas) * (b == 64 ? 8 : 4);
  *bn = b;
  return 0;
}

// The same for a convolution's forward GEMM (which may traverse virtual rows, SV3).
extern "C" int kfb_conv_stats_layout(const kfb_conv_desc* d, long long ldy, int* rows, int* bn) {
  KFB_REQUIRE(d != nullptr && rows != nullptr && bn != nullptr, "conv_stats_layout: bad arguments");
  const int num_kb = d->r * d->s * ((d->c + 63) / 64);
  const int b = pick_bn_for(d->k, num_kb);
  const int M = sv3_fprop_ok(d, b, ldy, nullptr) ? d->n * d->p * (d->q + 1) : d->n * d->p * d->q;
  return kfb_gemm_stats_layout(M, d->k, num_kb, rows, bn);
}

extern "C" int kfb_gemm_tn(const void* A, const void* B, void* C, int M, int N, int K,
                           long long lda, long long ldb, long long ldc, const float* bias,
                           int relu, int out_fp32, float* stats, void* stream) {
  KFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_tn: bad shape %d %d %d", M, N, K);
  KFB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_tn: lda/ldb must be multiples of 8 elements");
  if (load_driver_fns()) return 1;
  const int bn = pick_bn_for(N, (K + 63) / 64);
  CUtensorMap tmA, tmB;
  uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, sa[1] = {(uint64_t)lda * 2};
  uint32_t ba[2] = {64, 128};
  if (encode_tiled(&tmA, A, 2, da, sa, ba)) return 1;
  uint64_t db[2] = {(uint64_t)K, (uint64_t)N}, sb[1] = {(uint64_t)ldb * 2};
  uint32_t bb[2] = {64, (uint32_t)bn};
  if (encode_tiled(&tmB, B, 2, db, sb, bb)) return 1;
  GemmParams p;
  init_params(&p);
  p.mode = GEMM_TN;
  p.M = M;
  p.N = N;
  p.num_kb = (K + 63) / 64;
  p.kb_per_split = p.num_kb;
  p.out = C;
  p.ldc = ldc;
  p.bias = bias;
  p.relu = relu;
  p.out_fp32 = out_fp32;
  p.stats = stats;
  KFB_REQUIRE(stats == nullptr || !out_fp32, "gemm_tn: fused statistics need fp16 output");
  dim3 grid((M + 127) / 128, (N + bn - 1) / bn, 1);
  return launch_mode<GEMM_TN>(bn, tmA, tmB, p, grid, (cudaStream_t)stream);
}

// C[M,N] (+)= A[M,K] * B[K,N]; B is read in place as an MN-major operand.
extern "C" int kfb_gemm_nn(const void* A, const void* B, void* C, int M, int N, int K,
                           long long lda, long long ldb, long long ldc, int out_fp32,
                           int accumulate, void* stream) {
  KFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_nn: bad shape %d %d %d", M, N, K);
  KFB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_nn: lda/ldb must be multiples of 8 elements");
  if (load_driver_fns()) return 1;
  const int bn = pick_bn_for(N, (K + 63) / 64);
  CUtensorMap tmA, tmB;
  uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, sa[1] = {(uint64_t)lda * 2};
  uint32_t ba[2] = {64, 128};
  if (encode_tiled(&tmA, A, 2, da, sa, ba)) return 1;
  uint64_t db[2] = {(uint64_t)N, (uint64_t)K}, sb[1] = {(uint64_t)ldb * 2};
  uint32_t bb[2] = {64, 64};
  if (encode_tiled(&tmB, B, 2, db, sb, bb)) return 1;
  GemmParams p;
  init_params(&p);
  p.mode = GEMM_NN;
  p.M = M;
  p.N = N;
  p.num_kb = (K + 63) / 64;
  p.kb_per_split = p.num_kb;
  p.out = C;
  p.ldc = ldc;
  p.out_fp32 = out_fp32;
  p.accumulate = accumulate;
  dim3 grid((M + 127) / 128, (N + bn - 1) / bn, 1);
  return launch_mode<GEMM_NN>(bn, tmA, tmB, p, grid, (cudaStream_t)stream);
}


// Fills the kernel-side description of a fused reduce-scatter from the C ABI struct.
static int fill_push(const kfb_push_target* t, long long out_elems, int* n_out,
                     long long* bound, float** base) {
  *n_out = 0;
  if (t == nullptr || t->nranks <= 1) return 0;
  KFB_REQUIRE(t->nranks <= 8, "push: at most 8 ranks (%d)", t->nranks);
  KFB_REQUIRE(t->per > 0 && t->per % 4 == 0 && t->rel >= 0 && t->rel % 4 == 0,
              "push: shard length %lld / offset %lld must be positive multiples of 4", t->per,
              t->rel);
  for (int r = 0; r < 8; ++r) {
    bound[r] = r < t->nranks ? (long long)r * t->per - t->rel : (1ll << 62);
    base[r] = r < t->nranks ? (float*)t->dst[r] : nullptr;
    if (r < t->nranks) KFB_REQUIRE(base[r] != nullptr, "push: null destination for rank %d", r);
  }
  (void)out_elems;
  *n_out = t->nranks;
  return 0;
}

// Final stage shared by the weight-gradient entry points: either the deterministic split-K
// reduction into `out`, or (push) the same reduction scattered to the owners' staging slots.
static int finish_wgrad(float* workspace, float* out, long long slice, int splits, int accumulate,
                        float scale, const kfb_push_target* push, cudaStream_t st) {
  const long long n4 = (slice + 3) / 4;
  dim3 grid, block;
  splitk_launch_shape(n4, splits, &grid, &block);
  if (push != nullptr && push->nranks > 1) {
    PushTarget t;
    if (fill_push(push, slice, &t.n, t.bound, t.base)) return 2;
    splitk_reduce_push_kernel<<<grid, block, 0, st>>>(workspace, t, slice, splits, scale);
  } else {
    splitk_reduce_kernel<<<grid, block, 0, st>>>(workspace, out, slice, splits, accumulate,
                                                 scale);
  }
  KFB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// C[M,N] (fp32) = A[K,M]^T * B[K,N]; split over K into `workspace` then reduced. With a push
// target and a single K split the tcgen05 kernel's epilogue itself scatters the tiles to their
// owners (GEMM + reduce-scatter in one kernel); with several splits the reduction kernel does.
static int gemm_nt_impl(const void* A, const void* B, float* C, int M, int N, int K,
                        long long lda, long long ldb, long long ldc, float* workspace,
                        long long workspace_bytes, int accumulate, float scale,
                        const kfb_push_target* push, void* stream) {
  KFB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_nt: bad shape %d %d %d", M, N, K);
  KFB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_nt: lda/ldb must be multiples of 8 elements");
  KFB_REQUIRE(ldc == N, "gemm_nt: output must be dense (ldc == N)");
  if (load_driver_fns()) return 1;
  const int bn = pick_bn(N);
  CUtensorMap tmA, tmB;
  uint64_t da[2] = {(uint64_t)M, (uint64_t)K}, sa[1] = {(uint64_t)lda * 2};
  uint32_t bx[2] = {64, 64};
  if (encode_tiled(&tmA, A, 2, da, sa, bx)) return 1;
  uint64_t db[2] = {(uint64_t)N, (uint64_t)K}, sb[1] = {(uint64_t)ldb * 2};
  if (encode_tiled(&tmB, B, 2, db, sb, bx)) return 1;
  GemmParams p;
  init_params(&p);
  p.mode = GEMM_NT;
  p.M = M;
  p.N = N;
  p.num_kb = (K + 63) / 64;
  const int tiles = ((M + 127) / 128) * ((N + bn - 1) / bn);
  int splits = balanced_splits(tiles, p.num_kb, sm_count() < 148 ? sm_count() : 148);
  const long long slice = (long long)M * N;
  long long max_splits = workspace_bytes / (slice * 4);
  if (splits > max_splits) splits = (int)max_splits;
  KFB_REQUIRE(splits >= 1, "gemm_nt: workspace too small (%lld bytes for %lld-element slices)",
              workspace_bytes, slice);
  p.kb_per_split = (p.num_kb + splits - 1) / splits;
  splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.out = workspace;
  p.ldc = N;
  p.split_stride = slice;
  p.out_fp32 = 1;
  dim3 grid((M + 127) / 128, (N + bn - 1) / bn, splits);
  const bool pushing = push != nullptr && push->nranks > 1;
  KFB_REQUIRE(!(pushing && accumulate), "gemm_nt: a pushed gradient cannot accumulate");
  if (pushing && splits == 1 && scale == 1.0f) {
    if (fill_push(push, slice, &p.push_n, p.push_bound, p.push_base)) return 2;
    return launch_mode<GEMM_NT>(bn, tmA, tmB, p, grid, (cudaStream_t)stream);
  }
  if (launch_mode<GEMM_NT>(bn, tmA, tmB, p, grid, (cudaStream_t)stream)) return 1;
  return finish_wgrad(workspace, C, slice, splits, accumulate, scale, push, (cudaStream_t)stream);
}

extern "C" int kfb_gemm_nt(const void* A, const void* B, float* C, int M, int N, int K,
                           long long lda, long long ldb, long long ldc, float* workspace,
                           long long workspace_bytes, int accumulate, float scale, void* stream) {
  return gemm_nt_impl(A, B, C, M, N, K, lda, ldb, ldc, workspace, workspace_bytes, accumulate,
                      scale, nullptr, stream);
}

extern "C" int kfb_gemm_nt_rs(const void* A, const void* B, int M, int N, int K, long long lda,
                              long long ldb, float* workspace, long long workspace_bytes,
                              float scale, const kfb_push_target* push, void* stream) {
  KFB_REQUIRE(push != nullptr && push->nranks > 1, "gemm_nt_rs: needs a push target");
  return gemm_nt_impl(A, B, nullptr, M, N, K, lda, ldb, N, workspace, workspace_bytes, 0, scale,
                      push, stream);
}

static int check_conv(const kfb_conv_desc* d, const char* who) {
  KFB_REQUIRE(d != nullptr, "%s: null descriptor", who);
  KFB_REQUIRE(d->c % 8 == 0 && d->k % 8 == 0, "%s: C (%d) and Cout (%d) must be multiples of 8",
              who, d->c, d->k);
  KFB_REQUIRE(d->n > 0 && d->h > 0 && d->w > 0 && d->p > 0 && d->q > 0, "%s: bad geometry", who);
  KFB_REQUIRE(d->x_pix_stride >= 0 && d->x_row_stride >= 0 && d->x_pix_stride % 8 == 0 &&
                  d->x_row_stride % 8 == 0,
              "%s: input strides must be multiples of 8 elements", who);
  const int up_h = d->pad_h_hi - (d->r - 1), up_w = d->pad_w_hi - (d->s - 1);
  KFB_REQUIRE(d->pad_h_lo <= 127 && d->pad_w_lo <= 127 && up_h >= -128 && up_w >= -128,
              "%s: padding/filter exceed the TMA im2col corner range", who);
  KFB_REQUIRE((d->h + d->pad_h_lo + d->pad_h_hi - d->r) / d->stride_h + 1 == d->p &&
                  (d->w + d->pad_w_lo + d->pad_w_hi - d->s) / d->stride_w + 1 == d->q,
              "%s: output size (%d,%d) inconsistent with geometry", who, d->p, d->q);
  return 0;
}

// y[N,P,Q,ldy>=K] = conv(x[N,H,W,C], w[K,R,S,C]) (+bias) (ReLU)
extern "C" int kfb_conv_fprop(const kfb_conv_desc* d, const void* x, const void* w, void* y,
                              long long ldy, const float* bias, int relu, float* stats,
                              void* stream) {
  if (check_conv(d, "conv_fprop")) return 2;
  if (load_driver_fns()) return 1;
  const int M = d->n * d->p * d->q;
  const int bn = pick_bn_for(d->k, d->r * d->s * ((d->c + 63) / 64));
  CUtensorMap tmA, tmB;
  const bool sv3 = sv3_fprop_ok(d, bn, ldy, y);
  // the statistics layout was sized by kfb_conv_stats_layout(d, K) without the pointer
  KFB_REQUIRE(stats == nullptr || sv3 == sv3_fprop_ok(d, bn, d->k, nullptr),
              "conv_fprop: fused statistics need a 16-byte aligned output with ldy % 8 == 0");
  const int Mv = d->n * d->p * (d->q + 1);  // SV3: rows of Q + 1 virtual pixels
  if (encode_im2col(&tmA, x, d->n, d->h, d->w, d->c, -d->pad_w_lo, -d->pad_h_lo,
                    d->pad_w_hi - (d->s - 1) + (sv3 ? 1 : 0), d->pad_h_hi - (d->r - 1), 64,
                    sv3 ? 130 : 128, d->stride_w, d->stride_h, d->x_pix_stride, d->x_row_stride))
    return 1;
  const uint64_t Kt = (uint64_t)d->r * d->s * d->c;
  uint64_t db[2] = {Kt, (uint64_t)d->k}, sb[1] = {Kt * 2};
  uint32_t bb[2] = {64, (uint32_t)bn};
  if (encode_tiled(&tmB, w, 2, db, sb, bb)) return 1;
  GemmParams p;
  init_params(&p);
  p.mode = FPROP;
  p.M = M;
  p.N = d->k;
  p.R = d->r;
  p.S = d->s;
  p.CB = (d->c + 63) / 64;
  p.Ck = d->c;
  p.num_kb = d->r * d->s * p.CB;
  p.kb_per_split = p.num_kb;
  p.P = d->p;
  p.Q = d->q;
  p.lower_h = -d->pad_h_lo;
  p.lower_w = -d->pad_w_lo;
  p.stride_h = d->stride_h;
  p.stride_w = d->stride_w;
  p.out = y;
  p.ldc = ldy;
  p.bias = bias;
  p.relu = relu;
  p.stats = stats;
  if (sv3) {
    p.M = Mv;
    p.Q = d->q + 1;
    p.num_kb = d->r * p.CB;
    p.kb_per_split = p.num_kb;
    p.sc_P = d->p;
    p.sc_Q = d->q + 1;
    p.sc_H = d->p;
    p.sc_W = d->q;
    p.sc_sh = p.sc_sw = 1;
    p.sv3 = 1;
    dim3 grid_v((Mv + 127) / 128, (d->k + bn - 1) / bn, 1);
    return launch_sv3<FPROP>(bn, tmA, tmB, p, grid_v, (cudaStream_t)stream);
  }
  dim3 grid((M + 127) / 128, (d->k + bn - 1) / bn, 1);
  return launch_mode<FPROP>(bn, tmA, tmB, p, grid, (cudaStream_t)stream);
}

// dx[N,H,W,C] (+)= conv^T(dy[N,P,Q,K], w[K,R,S,C]).
// The transposed convolution is decomposed by output phase (h mod stride_h, w mod stride_w):
// every phase is a stride-1 correlation of dy with the sub-filter of taps r = r0 + stride*t,
// gathered by TMA im2col over dy and scattered to the phase's pixels of dx. Stride 1 is the
// single-phase case. Phases without any tap (e.g. 1x1 filters with stride 2) stay zero: without
// `accumulate` the caller zero-fills dx first (kfb_conv_dgrad_needs_zero_fill tells).
static int dgrad_phase_taps(int R, int stride, int pad_lo, int phase, int* r0, int* nR,
                            int* base) {
  *r0 = (phase + pad_lo) % stride;
  *nR = *r0 < R ? (R - *r0 + stride - 1) / stride : 0;
  *base = (phase + pad_lo - *r0) / stride;
  return 0;
}

extern "C" int kfb_conv_dgrad_needs_zero_fill(const kfb_conv_desc* d) {
  for (int ph = 0; ph < d->stride_h; ++ph)
    for (int pw = 0; pw < d->stride_w; ++pw) {
      int r0, nR, bh, s0, nS, bw;
      dgrad_phase_taps(d->r, d->stride_h, d->pad_h_lo, ph, &r0, &nR, &bh);
      dgrad_phase_taps(d->s, d->stride_w, d->pad_w_lo, pw, &s0, &nS, &bw);
      if (nR == 0 || nS == 0) return 1;
      // pixels of dx beyond the last window also receive nothing but are covered by OOB zeros
    }
  return 0;
}

static int conv_dgrad_impl(const kfb_conv_desc* d, const void* dy, const void* w, void* dx,
                           int accumulate, const void* res_src, long long res_ld,
                           const unsigned char* res_mask, long long res_mask_ld, void* stream);

extern "C" int kfb_conv_dgrad(const kfb_conv_desc* d, const void* dy, const void* w, void* dx,
                              int accumulate, void* stream) {
  return conv_dgrad_impl(d, dy, w, dx, accumulate, nullptr, 0, nullptr, 0, stream);
}

// dx = conv^T(dy, w) + res_src * mask: the data gradient of the first convolution of a residual
// block together with the gradient that reaches the block input through the identity shortcut
// (res_src = gradient of the block output, mask = packed ReLU mask of the block output, one bit
// per element, as kfb_bn_forward_train_fused writes it). The separate "shortcut gradient" tensor
// is then never written or read. Unit-stride convolutions whose output takes the TMA-store
// epilogue only (kfb_conv_dgrad_residual_ok).
extern "C" int kfb_conv_dgrad_residual_ok(const kfb_conv_desc* d) {
  return d->stride_h == 1 && d->stride_w == 1 && d->c >= 64 && d->c % 8 == 0 &&
         (long long)d->n * d->h * d->w >= 32;
}

extern "C" int kfb_conv_dgrad_residual(const kfb_conv_desc* d, const void* dy, const void* w,
                                       void* dx, const void* res_src, long long res_ld,
                                       const unsigned char* res_mask, long long res_mask_ld,
                                       void* stream) {
  KFB_REQUIRE(res_src != nullptr && kfb_conv_dgrad_residual_ok(d),
              "conv_dgrad_residual: needs a unit-stride convolution with C >= 64");
  KFB_REQUIRE(res_ld % 8 == 0, "conv_dgrad_residual: source row stride must be a multiple of 8");
  return conv_dgrad_impl(d, dy, w, dx, 0, res_src, res_ld, res_mask, res_mask_ld, stream);
}

static int conv_dgrad_impl(const kfb_conv_desc* d, const void* dy, const void* w, void* dx,
                           int accumulate, const void* res_src, long long res_ld,
                           const unsigned char* res_mask, long long res_mask_ld, void* stream) {
  if (check_conv(d, "conv_dgrad")) return 2;
  KFB_REQUIRE(d->x_pix_stride == 0 && d->x_row_stride == 0,
              "conv_dgrad: a strided (overlapping) input view has no data gradient");
  if (load_driver_fns()) return 1;
  CUtensorMap tmB;
  uint64_t db[3] = {(uint64_t)d->c, (uint64_t)d->r * d->s, (uint64_t)d->k};
  uint64_t sb[2] = {(uint64_t)d->c * 2, (uint64_t)d->r * d->s * d->c * 2};
  uint32_t bb[3] = {64, 1, 64};
  if (encode_tiled(&tmB, w, 3, db, sb, bb)) return 1;
  const bool strided = d->stride_h != 1 || d->stride_w != 1;
  for (int ph = 0; ph < d->stride_h; ++ph) {
    for (int pw = 0; pw < d->stride_w; ++pw) {
      int r0, nR, base_h, s0, nS, base_w;
      dgrad_phase_taps(d->r, d->stride_h, d->pad_h_lo, ph, &r0, &nR, &base_h);
      dgrad_phase_taps(d->s, d->stride_w, d->pad_w_lo, pw, &s0, &nS, &base_w);
      const int Hp = (d->h - ph + d->stride_h - 1) / d->stride_h;  // dx rows of this phase
      const int Wp = (d->w - pw + d->stride_w - 1) / d->stride_w;
      if (nR == 0 || nS == 0 || Hp <= 0 || Wp <= 0) continue;
      // dy row read for dx row index i and sub-tap t' (flipped): i + lower_h + t'
      const int lower_h = base_h - (nR - 1), lower_w = base_w - (nS - 1);
      const int upper_h = Hp - d->p + lower_h, upper_w = Wp - d->q + lower_w;
      KFB_REQUIRE(lower_h >= -128 && lower_w >= -128 && upper_h <= 127 && upper_w <= 127 &&
                      lower_h <= 127 && lower_w <= 127 && upper_h >= -128 && upper_w >= -128,
                  "conv_dgrad: geometry exceeds the TMA im2col corner range");
      const int num_kb_taps = nR * nS * ((d->k + 63) / 64);
      const bool sv3 = !strided && res_src == nullptr && nS == 3 && lower_w == -1 &&
                       upper_w == -1 && sv3_dgrad_ok(d, pick_bn_for(d->c, num_kb_taps), dx);
      CUtensorMap tmA;
      if (encode_im2col(&tmA, dy, d->n, d->p, d->q, d->k, lower_w, lower_h,
                        upper_w + (sv3 ? 1 : 0), upper_h, 64, sv3 ? 130 : 128, 1, 1))
        return 1;
      GemmParams p;
      init_params(&p);
      p.mode = DGRAD;
      p.M = d->n * Hp * Wp;
      p.N = d->c;
      p.R = nR;
      p.S = nS;
      p.CB = (d->k + 63) / 64;
      p.Ck = d->k;
      p.num_kb = nR * nS * p.CB;
      p.kb_per_split = p.num_kb;
      const int bn = pick_bn_for(d->c, p.num_kb);
      p.P = Hp;
      p.Q = Wp;
      p.lower_h = lower_h;
      p.lower_w = lower_w;
      p.stride_h = 1;
      p.stride_w = 1;
      p.wt_r0 = r0;
      p.wt_rstep = d->stride_h;
      p.wt_s0 = s0;
      p.wt_sstep = d->stride_w;
      p.wt_S = d->s;
      p.out = dx;
      p.ldc = d->c;
      p.accumulate = accumulate;
      p.acc_src = res_src;
      p.acc_src_ld = res_ld;
      p.acc_mask = res_mask;
      p.acc_mask_ld = res_mask_ld;
      if (strided) {
        p.sc_P = Hp;
        p.sc_Q = Wp;
        p.sc_H = d->h;
        p.sc_W = d->w;
        p.sc_sh = d->stride_h;
        p.sc_sw = d->stride_w;
        p.sc_h0 = ph;
        p.sc_w0 = pw;
      }
      if (sv3) {
        p.M = d->n * Hp * (Wp + 1);
        p.Q = Wp + 1;
        p.num_kb = nR * p.CB;
        p.kb_per_split = p.num_kb;
        p.sc_P = Hp;
        p.sc_Q = Wp + 1;
        p.sc_H = d->h;
        p.sc_W = d->w;
        p.sc_sh = p.sc_sw = 1;
        p.sv3 = 1;
        dim3 grid_v((p.M + 127) / 128, (d->c + bn - 1) / bn, 1);
        if (launch_sv3<DGRAD>(bn, tmA, tmB, p, grid_v, (cudaStream_t)stream)) return 1;
        continue;
      }
      dim3 grid((p.M + 127) / 128, (d->c + bn - 1) / bn, 1);
      if (launch_mode<DGRAD>(bn, tmA, tmB, p, grid, (cudaStream_t)stream)) return 1;
    }
  }
  return 0;
}

// Tile plan of the weight-gradient GEMM. With C % 64 == 0 the N axis is the flattened
// (tap, channel) axis, so one CTA covers several taps and re-uses its dy tile for all of them;
// otherwise every n-tile belongs to one tap.
static void wgrad_plan(const kfb_conv_desc* d, int* bn, int* n_tiles, int* flat, int* ncb) {
  const int ntot = d->r * d->s * d->c;
  if (d->c % 64 == 0) {
    *flat = 1;
    *bn = pick_bn(ntot);
    *n_tiles = (ntot + *bn - 1) / *bn;
    *ncb = 0;
  } else {
    *flat = 0;
    *bn = pick_bn(d->c);
    *ncb = (d->c + *bn - 1) / *bn;
    *n_tiles = d->r * d->s * *ncb;
  }
}

static int wgrad_splits(const kfb_conv_desc* d, int tiles, int num_kb, int sms) {
  (void)d;
  return balanced_splits(tiles, num_kb, sms);
}

extern "C" long long kfb_conv_wgrad_workspace_bytes(const kfb_conv_desc* d) {
  const long long slice = (long long)d->k * d->r * d->s * d->c * 4;
  int bn, n_tiles, flat, ncb;
  wgrad_plan(d, &bn, &n_tiles, &flat, &ncb);
  const int tiles = ((d->k + 127) / 128) * n_tiles;
  const int num_kb = (d->n * d->p * d->q + 63) / 64;
  return slice * wgrad_splits(d, tiles, num_kb, 148);  // never fewer SMs assumed than present
}

// dw[K,R,S,C] (fp32) (+)= scale * sum_pixels dy[pix,K]^T * im2col(x)[pix,(r,s,C)]
static int conv_wgrad_impl(const kfb_conv_desc* d, const void* x, const void* dy, long long lddy,
                           float* dw, float* workspace, long long workspace_bytes, int accumulate,
                           float scale, const kfb_push_target* push, void* stream) {
  if (check_conv(d, "conv_wgrad")) return 2;
  if (load_driver_fns()) return 1;
  const int pixels = d->n * d->p * d->q;
  int bn, n_tiles, flat, ncb;
  wgrad_plan(d, &bn, &n_tiles, &flat, &ncb);
  CUtensorMap tmA, tmB;
  uint64_t da[2] = {(uint64_t)d->k, (uint64_t)pixels}, sa[1] = {(uint64_t)lddy * 2};
  uint32_t bx[2] = {64, 64};
  if (encode_tiled(&tmA, dy, 2, da, sa, bx)) return 1;
  if (encode_im2col(&tmB, x, d->n, d->h, d->w, d->c, -d->pad_w_lo, -d->pad_h_lo,
                    d->pad_w_hi - (d->s - 1), d->pad_h_hi - (d->r - 1), 64, 64, d->stride_w,
                    d->stride_h, d->x_pix_stride, d->x_row_stride))
    return 1;
  GemmParams p;
  init_params(&p);
  p.mode = WGRAD;
  p.M = d->k;
  p.N = d->r * d->s * d->c;
  p.R = d->r;
  p.S = d->s;
  p.Cn = d->c;
  p.NCB = ncb;
  p.wg_flat = flat;
  p.num_kb = (pixels + 63) / 64;
  p.P = d->p;
  p.Q = d->q;
  p.lower_h = -d->pad_h_lo;
  p.lower_w = -d->pad_w_lo;
  p.stride_h = d->stride_h;
  p.stride_w = d->stride_w;
  const int m_tiles = (d->k + 127) / 128;
  const int tiles = m_tiles * n_tiles;
  int splits = wgrad_splits(d, tiles, p.num_kb, sm_count() < 148 ? sm_count() : 148);
  const long long slice = (long long)d->k * d->r * d->s * d->c;
  const long long max_splits = workspace_bytes / (slice * 4);
  if (splits > max_splits) splits = (int)max_splits;
  KFB_REQUIRE(splits >= 1, "conv_wgrad: workspace too small (%lld bytes, slice %lld elements)",
              workspace_bytes, slice);
  p.kb_per_split = (p.num_kb + splits - 1) / splits;
  splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
  p.out = workspace;
  p.ldc = (long long)d->r * d->s * d->c;
  p.split_stride = slice;
  p.out_fp32 = 1;
  dim3 grid(m_tiles, n_tiles, splits);
  const bool pushing = push != nullptr && push->nranks > 1;
  KFB_REQUIRE(!(pushing && accumulate), "conv_wgrad: a pushed gradient cannot accumulate");
  if (pushing && splits == 1 && scale == 1.0f) {
    if (fill_push(push, slice, &p.push_n, p.push_bound, p.push_base)) return 2;
    return launch_mode<WGRAD>(bn, tmA, tmB, p, grid, (cudaStream_t)stream);
  }
  if (launch_mode<WGRAD>(bn, tmA, tmB, p, grid, (cudaStream_t)stream)) return 1;
  return finish_wgrad(workspace, dw, slice, splits, accumulate, scale, push, (cudaStream_t)stream);
}

extern "C" int kfb_conv_wgrad(const kfb_conv_desc* d, const void* x, const void* dy,
                              long long lddy, float* dw, float* workspace,
                              long long workspace_bytes, int accumulate, float scale,
                              void* stream) {
  return conv_wgrad_impl(d, x, dy, lddy, dw, workspace, workspace_bytes, accumulate, scale,
                         nullptr, stream);
}

extern "C" int kfb_conv_wgrad_rs(const kfb_conv_desc* d, const void* x, const void* dy,
                                 long long lddy, float* workspace, long long workspace_bytes,
                                 float scale, const kfb_push_target* push, void* stream) {
  KFB_REQUIRE(push != nullptr && push->nranks > 1, "conv_wgrad_rs: needs a push target");
  return conv_wgrad_impl(d, x, dy, lddy, nullptr, workspace, workspace_bytes, 0, scale, push,
                         stream);
}
