// K1: persistent, warp-specialised bf16 GEMM for sm_100a.
//   TMA (cp.async.bulk.tensor, 128B swizzle) -> shared-memory ring -> tcgen05.mma (one issuing thread,
//   fp32 accumulators in TMEM, double-buffered) -> tcgen05.ld epilogue warps (bias / GELU / gamma /
//   fp32 residual / bf16 or fp32 store).
//   C[b][m][n] = epi(alpha * sum_k A[b][m][k] * B[b][n][k]);  A, B K-major (nn.Linear's [out,in] layout
//   is exactly a K-major B operand, so weights are used as stored).
//
// Covers (SURVEY.md §8a') every dense contraction of the path: trunk qkv/proj/MLP linears
// (reference hieradet.py:56-81, sam2_utils.py:131-136), patch-embed as im2col GEMM (backbones/utils.py:89-93),
// FpnNeck 1x1 convs + conv_s0/s1 (image_encoder.py:102-134, sam2_base_official.py:571-577), and the
// memory-attention / memory-encoder linears and score/PV products.
#pragma once
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

// epilogue specialisations (compile-time): everything not listed is off / alpha = 1 / bias along N
// MODE_F32_BIAS_RES_LN: MODE_F32_BIAS_RES that also emits what the NEXT LayerNorm needs -- a bf16 copy of its fp32 output and
// per-row partial sums (sum x, sum x^2 per 32-column-chunk parity of every N tile) -- so that LayerNorm never runs as a kernel:
// the consuming GEMM multiplies the raw bf16 rows with gamma-scaled, row-centred weights and finishes the normalisation in
// its epilogue: LN(x) W^T + b = rstd * (x W''^T) + (b + W beta) with W' = W diag(gamma), W''[n][k] = W'[n][k] - mean_k W'[n][:]
// (sum_k x_k W''[n][k] = sum_k (x_k - mu) W'[n][k]: the mean subtraction lives in the weights, only rstd is per row).
enum GemmMode { MODE_GENERIC = 0, MODE_BF16_BIAS = 1, MODE_BF16_BIAS_GELU = 2, MODE_F32_BIAS_RES = 3, MODE_F32_BIAS_RES_LN = 4 };

int make_tmap(CUtensorMap* m, const bf16* ptr, long rows, long K, long ld, long batch, long batch_stride, int box_rows);
int make_tmap_out(CUtensorMap* m, const void* ptr, int is_bf16, long rows, long cols, long ld, long batch, long batch_stride);

static constexpr int BM = 128;      // UMMA M (cta_group::1)
static constexpr int BK = 64;       // 64 bf16 = 128 B = one swizzle row
static constexpr int UMMA_K = 16;
// accumulator stages in TMEM: 2 x 256 columns, or 4 x 128 columns when BLOCK_N <= 128 (GemmCfg::NACC / ACC_STRIDE)
static constexpr int TMEM_COLS = 512;
// Epilogue warps EW = 8 or 16: 2 or 4 warps per TMEM lane quarter, interleaved over the 32-column chunks of a tile.
// warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps 4.. epilogue.  EW = 16 (640 threads) doubles the warps that
// hide each other's latency in the epilogue-bound small-K / GELU GEMMs (no prefetch registers there: 94 per thread).
static constexpr int gemm_threads(int ew) { return 128 + 32 * ew; }
static constexpr int SMEM_TOTAL = 227 * 1024;
static constexpr int BAR_BYTES = 512;

struct GemmKernelParams {
  int M, N, K, batch;
  int tiles_m, tiles_n, num_kb;
  uint32_t div_n_mul, div_n_shr, div_m_mul, div_m_shr;   // exact division of the tile index by tiles_n / tiles_m (mulhi + shift)
  int b_batched;
  float alpha;
  const float* bias; int bias_along_m;
  int act;
  const float* gamma;
  int has_res; int res_batched; int res_row_mod;
  // RoPE epilogue (MODE_GENERIC, bf16 output): the "residual" boxes are (cos, sin) pairs of the rotation table
  // [res_row_mod positions][256], column pair (2i, 2i + 1) of a row is rotated by its angle; rows (within a batch) at or past
  // rope_rows (rope_rows_dev[b] when given) pass through un-rotated
  int rope; int rope_rows; const int* rope_rows_dev;
  int out_is_bf16;
  // LayerNorm fold, consumer side: rows of A are raw (un-normalised); ln_stats [M][ln_parts][2] partial (sum, sum of squares)
  const float* ln_stats; int ln_parts; float ln_inv_c; float ln_eps;
  // producer side (MODE_F32_BIAS_RES_LN): partial sums of this GEMM's output rows, [M][2 * tiles_n][2]
  float* stats_out;
  int dbg;   // timing experiments (SAM2B_GEMM_DBG): 1 = no TMA store, 2 = no proxy fence, 4 = no staging writes
};

// NCTA = 2: a CTA pair (cluster of 2, `tcgen05.mma.cta_group::2`) works on a 256 x BN tile.  Each CTA holds 128 rows of A and
// of the accumulator and HALF of the B tile (BN/2 rows) in its own shared memory, so a CTA moves (128 + BN/2) x 64 operands per
// k-block instead of (128 + BN) x 64: one third less L2 -> SM traffic at BN = 256 and a deeper ring in the same smem.
template <int BN, int NCTA = 1, int MODE = 0, int EW = 8>
struct GemmCfg {
  static constexpr int NACC = BN <= 128 ? 4 : 2;
  static constexpr int ACC_STRIDE = 512 / NACC;
  static constexpr int STAGE_A = BM * BK * 2;
  static constexpr int B_ROWS = BN / NCTA;            // rows of the B tile this CTA loads
  static constexpr int STAGE_B = B_ROWS * BK * 2;
  static constexpr int STAGE = STAGE_A + STAGE_B;
  // Epilogue staging per warp: 32 rows x 128 B (32 fp32 or, for bf16, 32 rows x 64 B), the box of one TMA store.
  // BN <= 192 leaves room for two 4 KB buffers per warp (store of chunk i overlaps chunk i+1, residual prefetch);
  // BN > 192 gets one 4 KB buffer (two 2 KB buffers for bf16 outputs).
  // MODE_F32_BIAS_RES_LN adds a 2 KB bf16 staging box per warp behind the fp32 buffers.
  static constexpr int EPI_F32_BYTES = (BN > 192 || EW > 8) ? 4096 : 8192;
  static constexpr int EPI_WARP_BYTES = EPI_F32_BYTES + (MODE == MODE_F32_BIAS_RES_LN ? 2048 : 0);
  static constexpr int EPI_SMEM = EW * EPI_WARP_BYTES;
  static constexpr int AVAIL = SMEM_TOTAL - EPI_SMEM - 1024 /*align slack*/ - BAR_BYTES;
  static constexpr int STAGES = (AVAIL / STAGE) > 8 ? 8 : (AVAIL / STAGE);
  static constexpr int SMEM = STAGES * STAGE + EPI_SMEM + 1024 + BAR_BYTES;
};

// q = t / d for 0 <= t < 2^31 with (mul, shr) from fast_div_setup(d): q = umulhi(t, mul) >> shr
SAM2B_DEVINL uint32_t fast_div(uint32_t t, uint32_t mul, uint32_t shr) { return mul ? (__umulhi(t, mul) >> shr) : t; }
// tile index -> (n block, m block, batch), n fastest
SAM2B_DEVINL void decode_tile(const GemmKernelParams& p, int t, int& n_blk, int& m_blk, int& b) {
  const uint32_t r = fast_div((uint32_t)t, p.div_n_mul, p.div_n_shr);
  n_blk = t - (int)r * p.tiles_n;
  b = (int)fast_div(r, p.div_m_mul, p.div_m_shr);
  m_blk = (int)r - b * p.tiles_m;
}

template <int BN, int MODE, int NCTA, int EW>
__global__ void __launch_bounds__(gemm_threads(EW), 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_r,
                         const __grid_constant__ CUtensorMap tma_c2, const GemmKernelParams p) {
  using Cfg = GemmCfg<BN, NCTA, MODE, EW>;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int TILE_M = BM * NCTA;
  constexpr int NSUB = EW / 4;                        // epilogue warps per TMEM lane quarter
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * Cfg::STAGE_A;
  uint8_t* smem_epi = smem + STAGES * Cfg::STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + Cfg::EPI_SMEM);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = bars + 2 * STAGES + 4;
  uint64_t* res_bar = bars + 2 * STAGES + 8;          // [EW][2]: residual chunk landed in the staging buffer
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(res_bar + 2 * EW);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n * p.batch;
  const int cta_rank = NCTA == 2 ? (int)cluster_ctarank() : 0;      // 0 = leader (issues the MMAs of the pair)
  const int tile0 = NCTA == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = NCTA == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_c);
    if (p.has_res) tma_prefetch_desc(&tma_r);
    if (MODE == MODE_F32_BIAS_RES_LN) tma_prefetch_desc(&tma_c2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < Cfg::NACC; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], EW * NCTA); }
    for (int i = 0; i < 2 * EW; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    if (NCTA == 2) { tmem_alloc_2cta(tmem_ptr, TMEM_COLS); tmem_relinquish_2cta(); }
    else { tmem_alloc(tmem_ptr, TMEM_COLS); tmem_relinquish(); }
  }
  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();     // the pair's barriers are initialised before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = tile0; t < num_tiles; t += tile_step) {
        int n_blk, m_blk, b;
        decode_tile(p, t, n_blk, m_blk, b);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (NCTA == 2) {
            // both CTAs' loads complete on the LEADER's full barrier, which expects the bytes of the pair
            if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE);
            tma_load_3d_2cta(smem_a + stage * Cfg::STAGE_A, &tma_a, &full_bar[stage], kb * BK, m_blk * TILE_M + cta_rank * BM, b);
            tma_load_3d_2cta(smem_b + stage * Cfg::STAGE_B, &tma_b, &full_bar[stage], kb * BK, n_blk * BN + cta_rank * Cfg::B_ROWS,
                             p.b_batched ? b : 0);
          } else {
            mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE);
            tma_load_3d(smem_a + stage * Cfg::STAGE_A, &tma_a, &full_bar[stage], kb * BK, m_blk * BM, b);
            tma_load_3d(smem_b + stage * Cfg::STAGE_B, &tma_b, &full_bar[stage], kb * BK, n_blk * BN, p.b_batched ? b : 0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(TILE_M, BN);
      int stage = 0; uint32_t phase = 0;
      int local = 0;
      for (int t = tile0; t < num_tiles; t += tile_step, ++local) {
        const int acc = local % Cfg::NACC;
        const uint32_t acc_phase = (local / Cfg::NACC) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * Cfg::ACC_STRIDE;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint64_t da = make_smem_desc_sw128(smem_u32(smem_a + stage * Cfg::STAGE_A));
          const uint64_t db = make_smem_desc_sw128(smem_u32(smem_b + stage * Cfg::STAGE_B));
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advance 32 B (16 bf16) along K inside the 128 B swizzle row: +2 in the >>4 address field
            if (NCTA == 2) umma_bf16_2cta(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            else umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
          if (NCTA == 2) umma_commit_2cta(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        // accumulator complete -> epilogue warps (of both CTAs)
        if (NCTA == 2) umma_commit_2cta(&tmem_full[acc]); else umma_commit(&tmem_full[acc]);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> swizzled smem row -> TMA store =====================
    // Warp (4+we) may read TMEM lanes 32*(we&3)..+31 (hardware rule: warp id mod 4); the two warps of a lane quarter
    // take alternate 32-column chunks.  tcgen05.ld 32x32b gives thread i row i of the chunk, which it finishes in
    // registers (alpha, bias, GELU/ReLU, gamma, + fp32 residual that TMA dropped into the same staging buffer) and
    // writes as one swizzled row of the staging box; one lane then issues cp.async.bulk.tensor (smem -> global), which
    // also clips ragged M / N edges.  No per-thread global loads/stores, no transposition pass.
    const int we = warp - 4;
    const int q = we & 3;
    const int half = we >> 2;                 // which of the NSUB warps of this lane quarter: chunks half, half + NSUB, ...
    constexpr int NCHUNK = BN / 32;
    constexpr int WB = Cfg::EPI_WARP_BYTES;
    constexpr bool GEN = MODE == MODE_GENERIC;
    constexpr bool LNP = MODE == MODE_F32_BIAS_RES_LN;          // producer of LayerNorm inputs
    const bool bf = GEN ? (p.out_is_bf16 != 0) : (MODE == MODE_BF16_BIAS || MODE == MODE_BF16_BIAS_GELU);
    const bool has_res = GEN ? (p.has_res != 0) : (MODE == MODE_F32_BIAS_RES || LNP);
    const bool ln_fold = !LNP && MODE != MODE_F32_BIAS_RES && p.ln_stats != nullptr;   // consumer of LayerNorm inputs
    const bool bias_n = GEN ? (p.bias != nullptr && !p.bias_along_m) : true;
    const int act = GEN ? p.act : (MODE == MODE_BF16_BIAS_GELU ? (int)ACT_GELU : (int)ACT_NONE);
    const bool rope = GEN && p.rope != 0;                       // fp32 table boxes in, bf16 boxes out of the same staging buffer
    const uint32_t buf_bytes = (bf && !rope) ? 2048u : 4096u;
    const bool two_buf = (bf && !rope) || Cfg::EPI_F32_BYTES >= 8192;
    const uint32_t buf16 = smem_u32(smem_epi + we * WB) + Cfg::EPI_F32_BYTES;     // LNP: bf16 copy staging
    const uint32_t buf_base = smem_u32(smem_epi + we * WB);
    uint64_t* my_res_bar = res_bar + 2 * we;
    const uint32_t row_f32 = (uint32_t)lane * 128, xor_f32 = (uint32_t)(lane & 7) << 4;
    const uint32_t row_b16 = (uint32_t)lane * 64, xor_b16 = (uint32_t)((lane >> 1) & 3) << 4;
    uint32_t i = 0;                       // live chunks processed by this warp: buffer = i & 1 (two_buf), res phase
    int local = 0;
    for (int t = tile0; t < num_tiles; t += tile_step, ++local) {
      int n_blk, m_blk, b;
      decode_tile(p, t, n_blk, m_blk, b);
      const int acc = local % Cfg::NACC;
      const uint32_t acc_phase = (local / Cfg::NACC) & 1;
      const int row_g = m_blk * TILE_M + cta_rank * BM + q * 32;   // first row of this warp's 32-row slab
      const int n0 = n_blk * BN;
      const bool rows_live = row_g < p.M;
      const bool full_n = n0 + BN <= p.N;                // no column guards needed anywhere in this tile
      const int res_row = p.res_row_mod ? ((m_blk * TILE_M + cta_rank * BM) % p.res_row_mod) + q * 32 : row_g;
      const int res_b = p.res_batched ? b : 0;
      auto issue_res = [&](uint32_t ii, int c) {          // lane 0: residual chunk c -> staging buffer of chunk ii
        const uint32_t bsel = two_buf ? (ii & 1) : 0;
        mbar_arrive_expect_tx(&my_res_bar[bsel], 4096);
        tma_load_3d_addr(buf_base + bsel * buf_bytes, &tma_r, &my_res_bar[bsel], rope ? ((n0 + c * 32) & 255) : (n0 + c * 32), res_row, res_b);
      };
      const int rope_limit = rope ? (p.rope_rows_dev ? __ldg(p.rope_rows_dev + b) : p.rope_rows) : 0;
      auto load_vec8 = [&](const float* base, int c, float fill, float4 (&o)[8]) {
        // 32 consecutive per-column values (bias / gamma), same addresses for every lane (L1 broadcast)
        const int col0 = n0 + c * 32;
        if (base != nullptr && col0 + 32 <= p.N) {             // whole chunk inside N: 8 unguarded vector loads
          const float4* src = reinterpret_cast<const float4*>(base + col0);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = __ldg(src + j);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int col = col0 + 4 * j;
            o[j] = make_float4(fill, fill, fill, fill);
            if (base != nullptr) {
              if (col < p.N) o[j].x = __ldg(base + col);
              if (col + 1 < p.N) o[j].y = __ldg(base + col + 1);
              if (col + 2 < p.N) o[j].z = __ldg(base + col + 2);
              if (col + 3 < p.N) o[j].w = __ldg(base + col + 3);
            }
          }
        }
      };
      const bool first_live = rows_live && half < NCHUNK && (n0 + half * 32 < p.N);
      if (has_res && first_live && lane == 0) {
        if (two_buf) bulk_wait_read<1>(); else bulk_wait_read<0>();
        issue_res(i, half);
      }
      float4 bn[8];                                      // bias of the chunk in flight (EW = 8: prefetched one chunk ahead)
      if (EW != 16) load_vec8(bias_n ? p.bias : nullptr, half, 0.f, bn);
      float bias_m = 0.f;
      if (GEN) bias_m = (p.bias && p.bias_along_m && row_g + lane < p.M) ? __ldg(p.bias + row_g + lane) : 0.f;
      float ln_a = 1.f;                                  // LayerNorm fold: v = rstd(row) * acc + bias[n]
      if (ln_fold) {
        float s1 = 0.f, s2 = 0.f;
        if (row_g + lane < p.M) {
          const float2* st = reinterpret_cast<const float2*>(p.ln_stats) + (long)(row_g + lane) * p.ln_parts;
          for (int k = 0; k < p.ln_parts; ++k) { const float2 t = __ldg(st + k); s1 += t.x; s2 += t.y; }
        }
        const float mu = s1 * p.ln_inv_c;
        const float var = fmaxf(fmaf(s2, p.ln_inv_c, -mu * mu), 0.f);
        ln_a = rsqrtf(var + p.ln_eps);
      }
      float st1 = 0.f, st2 = 0.f;                         // LNP: this warp's partial row sums over its chunks of the tile
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * Cfg::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
      uint32_t r[32];
      if (EW != 16 && half < NCHUNK) tmem_ld_x32(taddr + half * 32, r);     // EW = 8: TMEM chunks are prefetched one ahead
#pragma unroll(EW == 16 ? 1 : 8)
      for (int c = half; c < NCHUNK; c += NSUB) {
        if (EW == 16) {     // 112 registers per thread: no prefetch, the other three warps of the SM sub-partition hide the latency
          tmem_ld_x32(taddr + c * 32, r);
          load_vec8(bias_n ? p.bias : nullptr, c, 0.f, bn);
        }
        tmem_ld_wait();
        if (c + NSUB >= NCHUNK) {
          // last chunk of this warp has left TMEM: hand the accumulator stage back to the MMA warp now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) { if (NCTA == 2) mbar_arrive_cluster(&tmem_empty[acc], 0); else mbar_arrive(&tmem_empty[acc]); }
        }
        const bool live = rows_live && (n0 + c * 32 < p.N);
        const bool next_live = rows_live && (c + NSUB < NCHUNK) && (n0 + (c + NSUB) * 32 < p.N);
        if (live) {
          const uint32_t bsel = two_buf ? (i & 1) : 0;
          const uint32_t buf = buf_base + bsel * buf_bytes;
          if (lane == 0) {
            if (has_res) {
              if (two_buf) { if (next_live) { bulk_wait_read<0>(); issue_res(i + 1, c + NSUB); } }
              else if (c != half) { bulk_wait_read<0>(); issue_res(i, c); }
            } else {
              if (two_buf) bulk_wait_read<1>(); else bulk_wait_read<0>();     // staging buffer `bsel` is free again
            }
          }
          __syncwarp();
          float4 v[8];
          if (GEN) {
            const float al = p.alpha * ln_a;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[j].x = fmaf(__uint_as_float(r[4 * j]), al, bn[j].x + bias_m);
              v[j].y = fmaf(__uint_as_float(r[4 * j + 1]), al, bn[j].y + bias_m);
              v[j].z = fmaf(__uint_as_float(r[4 * j + 2]), al, bn[j].z + bias_m);
              v[j].w = fmaf(__uint_as_float(r[4 * j + 3]), al, bn[j].w + bias_m);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[j].x = fmaf(__uint_as_float(r[4 * j]), ln_a, bn[j].x);
              v[j].y = fmaf(__uint_as_float(r[4 * j + 1]), ln_a, bn[j].y);
              v[j].z = fmaf(__uint_as_float(r[4 * j + 2]), ln_a, bn[j].z);
              v[j].w = fmaf(__uint_as_float(r[4 * j + 3]), ln_a, bn[j].w);
            }
          }
          if (EW != 16 && c + NSUB < NCHUNK) {
            tmem_ld_x32(taddr + (c + NSUB) * 32, r);   // next chunk streams out of TMEM (and its bias out of L1) meanwhile
            load_vec8(bias_n ? p.bias : nullptr, c + NSUB, 0.f, bn);
          }
          if (act == ACT_GELU) {
            if (MODE == MODE_BF16_BIAS_GELU) {
              // packed pairs (FFMA2): the polynomial of two columns per instruction - this epilogue is bound by instruction issue
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                gelu_fast_bf16x2(v[j].x, v[j].y, v[j].x, v[j].y);
                gelu_fast_bf16x2(v[j].z, v[j].w, v[j].z, v[j].w);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                v[j].x = gelu_fast(v[j].x); v[j].y = gelu_fast(v[j].y); v[j].z = gelu_fast(v[j].z); v[j].w = gelu_fast(v[j].w);
              }
            }
          } else if (act == ACT_RELU) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[j].x = fmaxf(v[j].x, 0.f); v[j].y = fmaxf(v[j].y, 0.f); v[j].z = fmaxf(v[j].z, 0.f); v[j].w = fmaxf(v[j].w, 0.f);
            }
          }
          if (GEN && p.gamma) {
            float4 gm[8];
            load_vec8(p.gamma, c, 1.f, gm);
#pragma unroll
            for (int j = 0; j < 8; ++j) { v[j].x *= gm[j].x; v[j].y *= gm[j].y; v[j].z *= gm[j].z; v[j].w *= gm[j].w; }
          }
          if (has_res) {
            mbar_wait(&my_res_bar[bsel], (two_buf ? (i >> 1) : i) & 1);
            if (rope) {
              const bool rot = row_g + lane < rope_limit;
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 cs = ld_shared_f4(buf + row_f32 + (((uint32_t)j << 4) ^ xor_f32));     // (cos, sin) of two column pairs
                if (!rot) cs = make_float4(1.f, 0.f, 1.f, 0.f);
                const float4 x = v[j];
                v[j].x = x.x * cs.x - x.y * cs.y;
                v[j].y = x.x * cs.y + x.y * cs.x;
                v[j].z = x.z * cs.z - x.w * cs.w;
                v[j].w = x.z * cs.w + x.w * cs.z;
              }
              __syncwarp();          // every lane has read its table row before the (narrower) bf16 rows overwrite the box
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float4 rv = ld_shared_f4(buf + row_f32 + (((uint32_t)j << 4) ^ xor_f32));
                v[j].x += rv.x; v[j].y += rv.y; v[j].z += rv.z; v[j].w += rv.w;
              }
            }
          }
          if (LNP) {
            // columns >= N hold exact zeros here (zero-filled B rows, guarded bias, zero-filled residual box)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              st1 += (v[j].x + v[j].y) + (v[j].z + v[j].w);
              st2 = fmaf(v[j].x, v[j].x, st2); st2 = fmaf(v[j].y, v[j].y, st2); st2 = fmaf(v[j].z, v[j].z, st2); st2 = fmaf(v[j].w, v[j].w, st2);
            }
            if (lane == 0) bulk_wait_read<0>();          // the previous chunk's stores have left the bf16 staging box
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_u4(buf16 + row_b16 + (((uint32_t)j << 4) ^ xor_b16), pack_bf16x2(v[2 * j].x, v[2 * j].y),
                           pack_bf16x2(v[2 * j].z, v[2 * j].w), pack_bf16x2(v[2 * j + 1].x, v[2 * j + 1].y),
                           pack_bf16x2(v[2 * j + 1].z, v[2 * j + 1].w));
          }
          if (p.dbg & 4) {
            if (v[0].x + v[3].y + v[5].z + v[7].w == 123.456f) st_shared_u4(buf, 0, 0, 0, 0);
          } else if (bf) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_shared_u4(buf + row_b16 + (((uint32_t)j << 4) ^ xor_b16), pack_bf16x2(v[2 * j].x, v[2 * j].y),
                           pack_bf16x2(v[2 * j].z, v[2 * j].w), pack_bf16x2(v[2 * j + 1].x, v[2 * j + 1].y),
                           pack_bf16x2(v[2 * j + 1].z, v[2 * j + 1].w));
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              st_shared_u4(buf + row_f32 + (((uint32_t)j << 4) ^ xor_f32), __float_as_uint(v[j].x), __float_as_uint(v[j].y),
                           __float_as_uint(v[j].z), __float_as_uint(v[j].w));
          }
          if (!(p.dbg & 2)) fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0 && !(p.dbg & 1)) {
            tma_store_3d(&tma_c, buf, n0 + c * 32, row_g, b);
            if (LNP) tma_store_3d(&tma_c2, buf16, n0 + c * 32, row_g, b);     // same bulk group as the fp32 box
            bulk_commit();
          }
          ++i;
        } else {
          if (EW != 16 && c + NSUB < NCHUNK) {
            load_vec8(bias_n ? p.bias : nullptr, c + NSUB, 0.f, bn);   // (dead chunks: values unused)
            tmem_ld_x32(taddr + (c + NSUB) * 32, r);
          }
        }
      }
      if (LNP && p.stats_out != nullptr && row_g + lane < p.M) {
        float2* so = reinterpret_cast<float2*>(p.stats_out) + ((long)b * p.M + row_g + lane) * (NSUB * p.tiles_n) + NSUB * n_blk + half;
        *so = make_float2(st1, st2);
      }
      if (half >= NCHUNK) {   // more warps per quarter than chunks: a warp without a chunk still owes its arrival
        if (lane == 0) { if (NCTA == 2) mbar_arrive_cluster(&tmem_empty[acc], 0); else mbar_arrive(&tmem_empty[acc]); }
      }
    }
    if (lane == 0) bulk_wait_read<0>();    // staging smem must outlive the last TMA store's read
  }

  tc_fence_before();
  if (NCTA == 2) cluster_sync_all(); else __syncthreads();     // the peer's smem / TMEM / barriers stay alive until both are done
  if (warp == 2) {
    tc_fence_after();
    if (NCTA == 2) tmem_dealloc_2cta(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// host side: launch of one instantiation
// ------------------------------------------------------------------------------------------
// (mul, shr) with umulhi(t, mul) >> shr == t / d for every 0 <= t < 2^31 (round-up method); d == 1 -> mul = 0 (identity)
static inline void fast_div_setup(uint32_t d, uint32_t* mul, uint32_t* shr) {
  if (d <= 1) { *mul = 0; *shr = 0; return; }
  uint32_t l = 0;
  while ((1ull << l) < d) ++l;                     // l = ceil(log2 d)
  const unsigned long long m = ((1ull << (31 + l)) + d - 1) / d;   // < 2^32 because 2^(l-1) < d
  *mul = (uint32_t)m; *shr = l - 1;
}

template <int BN, int MODE, int NCTA, int EW>
static int launch_gemm(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = GemmCfg<BN, NCTA, MODE, EW>;
  constexpr int NUM_THREADS = gemm_threads(EW);
  static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "BLOCK_N must be a multiple of 32");
  static_assert(Cfg::STAGES >= 2, "pipeline too shallow");
  if (Cfg::STAGES < 3) { set_error("gemm<%d,%d,%d,%d>: only %d smem stages fit; not a supported combination", BN, MODE, NCTA, EW, Cfg::STAGES); return 1; }
  static_assert(NCTA == 1 || (BN % 32 == 0 && (BN / 2) % 8 == 0), "a CTA pair splits the B tile in two halves of whole swizzle atoms");
  static int max_pairs = 0;       // resident CTAs / CTA pairs of this instantiation (identical GPUs assumed within a process)
  static PerDeviceFlag attr;      // attributes set on the current device
  auto kern = gemm_bf16_tcgen05_kernel<BN, MODE, NCTA, EW>;
  if (!attr.is_set()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(gemm<%d,%d>): %s", BN, NCTA, cudaGetErrorString(e)); return 1; }
    int mp = num_sms() / NCTA;
    if (NCTA == 2) {
      cudaLaunchConfig_t qc = {};
      qc.gridDim = dim3(num_sms() / 2 * 2); qc.blockDim = dim3(NUM_THREADS); qc.dynamicSmemBytes = Cfg::SMEM;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension; qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
      qc.attrs = qa; qc.numAttrs = 1;
      int n_clusters = 0;
      e = cudaOccupancyMaxActiveClusters(&n_clusters, kern, &qc);
      if (e != cudaSuccess || n_clusters <= 0) { set_error("cudaOccupancyMaxActiveClusters(gemm<%d,2>): %s (%d)", BN, cudaGetErrorString(e), n_clusters); return 1; }
      if (n_clusters < mp) mp = n_clusters;
    }
    max_pairs = mp;
    attr.set();
  }
  CUtensorMap ta, tb, tc, tr, tc2;
  if (make_tmap(&ta, a.A, a.M, a.K, a.lda, a.batch, a.a_batch_stride, BM)) return 1;
  const bool b_batched = a.batch > 1 && a.b_batch_stride != 0;
  if (make_tmap(&tb, a.B, a.N, a.K, a.ldb, b_batched ? a.batch : 1, a.b_batch_stride, Cfg::B_ROWS)) return 1;
  const int out_bf16 = a.out_bf16 != nullptr;
  if (make_tmap_out(&tc, out_bf16 ? (const void*)a.out_bf16 : (const void*)a.out_f32, out_bf16, a.M, a.N, a.ldc, a.batch,
                    a.c_batch_stride)) return 1;
  const bool res_batched = a.residual && a.batch > 1 && a.res_batch_stride != 0;
  if (a.rope_table) {
    if (make_tmap_out(&tr, a.rope_table, 0, a.rope_positions, 256, 256, 1, 0)) return 1;
  } else if (a.residual) {
    const long res_rows = a.res_row_mod ? a.res_row_mod : a.M;
    if (make_tmap_out(&tr, a.residual, 0, res_rows, a.N, a.ld_res, res_batched ? a.batch : 1, a.res_batch_stride)) return 1;
  } else {
    tr = tc;
  }
  if (MODE == MODE_F32_BIAS_RES_LN) {
    if (make_tmap_out(&tc2, a.out2_bf16, 1, a.M, a.N, a.ldc2, a.batch, a.c2_batch_stride)) return 1;
  } else {
    tc2 = tc;
  }
  GemmKernelParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.batch = a.batch;
  p.tiles_m = (a.M + BM * NCTA - 1) / (BM * NCTA); p.tiles_n = (a.N + BN - 1) / BN; p.num_kb = (a.K + BK - 1) / BK;
  fast_div_setup((uint32_t)p.tiles_n, &p.div_n_mul, &p.div_n_shr);
  fast_div_setup((uint32_t)p.tiles_m, &p.div_m_mul, &p.div_m_shr);
  p.b_batched = b_batched ? 1 : 0;
  p.alpha = a.alpha; p.bias = a.bias; p.bias_along_m = a.bias_along_m; p.act = a.act; p.gamma = a.gamma;
  p.has_res = a.residual ? 1 : 0; p.res_batched = res_batched ? 1 : 0; p.res_row_mod = a.res_row_mod;
  p.rope = 0; p.rope_rows = 0; p.rope_rows_dev = nullptr;
  if (a.rope_table) {
    p.has_res = 1; p.res_batched = 0; p.res_row_mod = a.rope_positions;
    p.rope = 1; p.rope_rows = a.rope_rows; p.rope_rows_dev = a.rope_rows_dev;
  }
  p.out_is_bf16 = out_bf16;
  p.ln_stats = a.ln_stats; p.ln_parts = a.ln_parts; p.ln_inv_c = a.ln_stats ? 1.0f / (float)a.K : 0.f; p.ln_eps = a.ln_eps;
  p.stats_out = a.stats_out;
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("SAM2B_GEMM_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
  const long tiles = (long)p.tiles_m * p.tiles_n * p.batch;
  const int grid = (int)(tiles < max_pairs ? tiles : max_pairs) * NCTA;
  char nm[96];
  if (g_prof_on) snprintf(nm, sizeof(nm), "gemm m%d n%d k%d b%d bn%d%s%s md%d act%d%s%s%s", a.M, a.N, a.K, a.batch, BN, NCTA == 2 ? "x2" : "", EW == 16 ? "w16" : "", MODE, a.act, a.residual ? " res" : (a.rope_table ? " rope" : ""), a.out_bf16 ? " o16" : " o32", a.ln_stats ? " ln" : "");
  const double esz = a.out_bf16 ? 2.0 : 4.0;
  ProfRange pr(stream, nm, 2.0 * a.M * a.N * (double)a.K * a.batch,
               ((double)a.M * a.K * 2 + (b_batched ? 1.0 : 1.0 / a.batch) * (double)a.N * a.K * 2 + (double)a.M * a.N * (esz + (a.residual ? 4.0 : 0.0))) * a.batch);
  if (NCTA == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(NUM_THREADS); cfg.dynamicSmemBytes = Cfg::SMEM; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ta, tb, tc, tr, tc2, p);
    if (e != cudaSuccess) { set_error("cudaLaunchKernelEx(gemm<%d,2>): %s", BN, cudaGetErrorString(e)); return 1; }
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
  } else {
    kern<<<grid, NUM_THREADS, Cfg::SMEM, stream>>>(ta, tb, tc, tr, tc2, p);
    SAM2B_CHECK_LAUNCH("gemm_bf16_tcgen05");
  }
  return 0;
}


// One translation unit per BLOCK_N (parallel compilation): the specialised modes for the large tiles, generic otherwise.
// `cta2`: run as CTA pairs (256 x BN tiles); `ew16`: 16 epilogue warps instead of 8.
#define SAM2B_GEMM_DISPATCH_MODES(BN_, NCTA_, EW_)                                                   \
  switch (mode) {                                                                                   \
    case MODE_BF16_BIAS: return launch_gemm<BN_, MODE_BF16_BIAS, NCTA_, EW_>(a, stream);            \
    case MODE_BF16_BIAS_GELU: return launch_gemm<BN_, MODE_BF16_BIAS_GELU, NCTA_, EW_>(a, stream);  \
    case MODE_F32_BIAS_RES: return launch_gemm<BN_, MODE_F32_BIAS_RES, NCTA_, EW_>(a, stream);      \
    case MODE_F32_BIAS_RES_LN: return launch_gemm<BN_, MODE_F32_BIAS_RES_LN, NCTA_, EW_>(a, stream); \
    default: return launch_gemm<BN_, MODE_GENERIC, NCTA_, EW_>(a, stream);                          \
  }
#define SAM2B_GEMM_INSTANTIATE_FULL(BN_)                                                            \
  int launch_gemm_bn##BN_(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream) {    \
    if (cta2 && ew16) { SAM2B_GEMM_DISPATCH_MODES(BN_, 2, 16) }                                     \
    if (cta2) { SAM2B_GEMM_DISPATCH_MODES(BN_, 2, 8) }                                              \
    if (ew16) { SAM2B_GEMM_DISPATCH_MODES(BN_, 1, 16) }                                             \
    SAM2B_GEMM_DISPATCH_MODES(BN_, 1, 8)                                                            \
  }
#define SAM2B_GEMM_INSTANTIATE_GENERIC(BN_)                                                         \
  int launch_gemm_bn##BN_(const GemmArgs& a, int, int, int, cudaStream_t stream) { return launch_gemm<BN_, MODE_GENERIC, 1, 8>(a, stream); }

int launch_gemm_bn256(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn224(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn192(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn160(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn128(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn96(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn64(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);
int launch_gemm_bn32(const GemmArgs& a, int mode, int cta2, int ew16, cudaStream_t stream);

}  // namespace sam2b
