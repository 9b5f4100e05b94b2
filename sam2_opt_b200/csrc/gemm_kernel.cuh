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
enum GemmMode { MODE_GENERIC = 0, MODE_BF16_BIAS = 1, MODE_BF16_BIAS_GELU = 2, MODE_F32_BIAS_RES = 3 };

int make_tmap(CUtensorMap* m, const bf16* ptr, long rows, long K, long ld, long batch, long batch_stride, int box_rows);
int make_tmap_out(CUtensorMap* m, const void* ptr, int is_bf16, long rows, long cols, long ld, long batch, long batch_stride);

static constexpr int BM = 128;      // UMMA M (cta_group::1)
static constexpr int BK = 64;       // 64 bf16 = 128 B = one swizzle row
static constexpr int UMMA_K = 16;
// accumulator stages in TMEM: 2 x 256 columns, or 4 x 128 columns when BLOCK_N <= 128 (GemmCfg::NACC / ACC_STRIDE)
static constexpr int TMEM_COLS = 512;
static constexpr int NUM_EPI_WARPS = 8;  // two warps per TMEM lane quarter, interleaved over 32-column chunks
static constexpr int NUM_THREADS = 128 + 32 * NUM_EPI_WARPS;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-11 epilogue
static constexpr int SMEM_TOTAL = 227 * 1024;
static constexpr int BAR_BYTES = 512;

struct GemmKernelParams {
  int M, N, K, batch;
  int tiles_m, tiles_n, num_kb;
  int b_batched;
  float alpha;
  const float* bias; int bias_along_m;
  int act;
  const float* gamma;
  int has_res; int res_batched; int res_row_mod;
  int out_is_bf16;
  int dbg;   // timing experiments (SAM2B_GEMM_DBG): 1 = no TMA store, 2 = no proxy fence, 4 = no staging writes
};

template <int BN>
struct GemmCfg {
  static constexpr int NACC = BN <= 128 ? 4 : 2;
  static constexpr int ACC_STRIDE = 512 / NACC;
  static constexpr int STAGE_A = BM * BK * 2;
  static constexpr int STAGE_B = BN * BK * 2;
  static constexpr int STAGE = STAGE_A + STAGE_B;
  // Epilogue staging per warp: 32 rows x 128 B (32 fp32 or, for bf16, 32 rows x 64 B), the box of one TMA store.
  // BN <= 192 leaves room for two 4 KB buffers per warp (store of chunk i overlaps chunk i+1, residual prefetch);
  // BN > 192 gets one 4 KB buffer (two 2 KB buffers for bf16 outputs).
  static constexpr int EPI_WARP_BYTES = BN > 192 ? 4096 : 8192;
  static constexpr int EPI_SMEM = NUM_EPI_WARPS * EPI_WARP_BYTES;
  static constexpr int AVAIL = SMEM_TOTAL - EPI_SMEM - 1024 /*align slack*/ - BAR_BYTES;
  static constexpr int STAGES = (AVAIL / STAGE) > 8 ? 8 : (AVAIL / STAGE);
  static constexpr int SMEM = STAGES * STAGE + EPI_SMEM + 1024 + BAR_BYTES;
};

// Exact-erf GELU, x * Phi(x), evaluated as relu(x) - |x| * h(|x|) with h(t) = Phi(-t) = 2^p(t); p is a degree-6
// fit of log2(0.5 erfc(t/sqrt 2)) on [0,6] (max |error| of the GELU value 2e-7 in fp32) -> 6 FMA + 1 MUFU.EX2.
SAM2B_DEVINL float gelu_fast(float x) {
  const float t = fminf(fabsf(x), 6.0f);
  float p = 3.242628146753161e-05f;
  p = fmaf(p, t, -0.0007628580443300231f);
  p = fmaf(p, t, 0.008056423025940528f);
  p = fmaf(p, t, -0.05336353987295394f);
  p = fmaf(p, t, -0.4588251982307528f);
  p = fmaf(p, t, -1.1511703007880216f);
  p = fmaf(p, t, -1.0f);
  float h;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(h) : "f"(p));
  return fmaf(-fabsf(x), h, fmaxf(x, 0.0f));
}

template <int BN, int MODE>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_r,
                         const GemmKernelParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
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
  uint64_t* res_bar = bars + 2 * STAGES + 8;          // [NUM_EPI_WARPS][2]: residual chunk landed in the staging buffer
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(res_bar + 2 * NUM_EPI_WARPS);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n * p.batch;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_c);
    if (p.has_res) tma_prefetch_desc(&tma_r);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < Cfg::NACC; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], NUM_EPI_WARPS); }
    for (int i = 0; i < 2 * NUM_EPI_WARPS; ++i) mbar_init(&res_bar[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int n_blk = t % p.tiles_n;
        const int m_blk = (t / p.tiles_n) % p.tiles_m;
        const int b = t / (p.tiles_n * p.tiles_m);
        for (int kb = 0; kb < p.num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE);
          tma_load_3d(smem_a + stage * Cfg::STAGE_A, &tma_a, &full_bar[stage], kb * BK, m_blk * BM, b);
          tma_load_3d(smem_b + stage * Cfg::STAGE_B, &tma_b, &full_bar[stage], kb * BK, n_blk * BN, p.b_batched ? b : 0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN);
      int stage = 0; uint32_t phase = 0;
      int local = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++local) {
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
            umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);   // frees the smem slot once these MMAs have read it
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[acc]);       // accumulator complete -> epilogue
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
    const int half = we >> 2;
    constexpr int NCHUNK = BN / 32;
    constexpr int WB = Cfg::EPI_WARP_BYTES;
    constexpr bool GEN = MODE == MODE_GENERIC;
    const bool bf = GEN ? (p.out_is_bf16 != 0) : (MODE == MODE_BF16_BIAS || MODE == MODE_BF16_BIAS_GELU);
    const bool has_res = GEN ? (p.has_res != 0) : (MODE == MODE_F32_BIAS_RES);
    const bool bias_n = GEN ? (p.bias != nullptr && !p.bias_along_m) : true;
    const int act = GEN ? p.act : (MODE == MODE_BF16_BIAS_GELU ? (int)ACT_GELU : (int)ACT_NONE);
    const uint32_t buf_bytes = bf ? 2048u : 4096u;
    const bool two_buf = bf || WB >= 8192;
    const uint32_t buf_base = smem_u32(smem_epi + we * WB);
    uint64_t* my_res_bar = res_bar + 2 * we;
    const uint32_t row_f32 = (uint32_t)lane * 128, xor_f32 = (uint32_t)(lane & 7) << 4;
    const uint32_t row_b16 = (uint32_t)lane * 64, xor_b16 = (uint32_t)((lane >> 1) & 3) << 4;
    uint32_t i = 0;                       // live chunks processed by this warp: buffer = i & 1 (two_buf), res phase
    int local = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++local) {
      const int n_blk = t % p.tiles_n;
      const int m_blk = (t / p.tiles_n) % p.tiles_m;
      const int b = t / (p.tiles_n * p.tiles_m);
      const int acc = local % Cfg::NACC;
      const uint32_t acc_phase = (local / Cfg::NACC) & 1;
      const int row_g = m_blk * BM + q * 32;             // first row of this warp's 32-row slab
      const int n0 = n_blk * BN;
      const bool rows_live = row_g < p.M;
      const bool full_n = n0 + BN <= p.N;                // no column guards needed anywhere in this tile
      const int res_row = p.res_row_mod ? ((m_blk * BM) % p.res_row_mod) + q * 32 : row_g;
      const int res_b = p.res_batched ? b : 0;
      auto issue_res = [&](uint32_t ii, int c) {          // lane 0: residual chunk c -> staging buffer of chunk ii
        const uint32_t bsel = two_buf ? (ii & 1) : 0;
        mbar_arrive_expect_tx(&my_res_bar[bsel], 4096);
        tma_load_3d_addr(buf_base + bsel * buf_bytes, &tma_r, &my_res_bar[bsel], n0 + c * 32, res_row, res_b);
      };
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
      float4 bn[8];                                      // bias of the chunk in flight (reloaded right after use)
      load_vec8(bias_n ? p.bias : nullptr, half, 0.f, bn);
      float bias_m = 0.f;
      if (GEN) bias_m = (p.bias && p.bias_along_m && row_g + lane < p.M) ? __ldg(p.bias + row_g + lane) : 0.f;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * Cfg::ACC_STRIDE + ((uint32_t)(q * 32) << 16);
      uint32_t r[32];
      if (half < NCHUNK) tmem_ld_x32(taddr + half * 32, r);
#pragma unroll 1
      for (int c = half; c < NCHUNK; c += 2) {
        tmem_ld_wait();
        if (c + 2 >= NCHUNK) {
          // last chunk of this warp has left TMEM: hand the accumulator stage back to the MMA warp now
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        const bool live = rows_live && (n0 + c * 32 < p.N);
        const bool next_live = rows_live && (c + 2 < NCHUNK) && (n0 + (c + 2) * 32 < p.N);
        if (live) {
          const uint32_t bsel = two_buf ? (i & 1) : 0;
          const uint32_t buf = buf_base + bsel * buf_bytes;
          if (lane == 0) {
            if (has_res) {
              if (two_buf) { if (next_live) { bulk_wait_read<0>(); issue_res(i + 1, c + 2); } }
              else if (c != half) { bulk_wait_read<0>(); issue_res(i, c); }
            } else {
              if (two_buf) bulk_wait_read<1>(); else bulk_wait_read<0>();     // staging buffer `bsel` is free again
            }
          }
          __syncwarp();
          float4 v[8];
          if (GEN) {
            const float al = p.alpha;
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
              v[j].x = __uint_as_float(r[4 * j]) + bn[j].x;
              v[j].y = __uint_as_float(r[4 * j + 1]) + bn[j].y;
              v[j].z = __uint_as_float(r[4 * j + 2]) + bn[j].z;
              v[j].w = __uint_as_float(r[4 * j + 3]) + bn[j].w;
            }
          }
          if (c + 2 < NCHUNK) {
            tmem_ld_x32(taddr + (c + 2) * 32, r);   // next chunk streams out of TMEM (and its bias out of L1) meanwhile
            load_vec8(bias_n ? p.bias : nullptr, c + 2, 0.f, bn);
          }
          if (act == ACT_GELU) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[j].x = gelu_fast(v[j].x); v[j].y = gelu_fast(v[j].y); v[j].z = gelu_fast(v[j].z); v[j].w = gelu_fast(v[j].w);
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
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 rv = ld_shared_f4(buf + row_f32 + (((uint32_t)j << 4) ^ xor_f32));
              v[j].x += rv.x; v[j].y += rv.y; v[j].z += rv.z; v[j].w += rv.w;
            }
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
            bulk_commit();
          }
          ++i;
        } else {
          if (c + 2 < NCHUNK) {
            load_vec8(bias_n ? p.bias : nullptr, c + 2, 0.f, bn);
            tmem_ld_x32(taddr + (c + 2) * 32, r);
          }
        }
      }
      if (half >= NCHUNK) {   // BN = 32: the second warp of each quarter has no chunk but still owes its arrival
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      }
    }
    if (lane == 0) bulk_wait_read<0>();    // staging smem must outlive the last TMA store's read
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// host side: launch of one instantiation
// ------------------------------------------------------------------------------------------
template <int BN, int MODE>
static int launch_gemm(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "BLOCK_N must be a multiple of 32");
  static_assert(Cfg::STAGES >= 3, "pipeline too shallow");
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(gemm<%d>): %s", BN, cudaGetErrorString(e)); return 1; }
    attr_set = true;
  }
  CUtensorMap ta, tb, tc, tr;
  if (make_tmap(&ta, a.A, a.M, a.K, a.lda, a.batch, a.a_batch_stride, BM)) return 1;
  const bool b_batched = a.batch > 1 && a.b_batch_stride != 0;
  if (make_tmap(&tb, a.B, a.N, a.K, a.ldb, b_batched ? a.batch : 1, a.b_batch_stride, BN)) return 1;
  const int out_bf16 = a.out_bf16 != nullptr;
  if (make_tmap_out(&tc, out_bf16 ? (const void*)a.out_bf16 : (const void*)a.out_f32, out_bf16, a.M, a.N, a.ldc, a.batch,
                    a.c_batch_stride)) return 1;
  const bool res_batched = a.residual && a.batch > 1 && a.res_batch_stride != 0;
  if (a.residual) {
    const long res_rows = a.res_row_mod ? a.res_row_mod : a.M;
    if (make_tmap_out(&tr, a.residual, 0, res_rows, a.N, a.ld_res, res_batched ? a.batch : 1, a.res_batch_stride)) return 1;
  } else {
    tr = tc;
  }
  GemmKernelParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.batch = a.batch;
  p.tiles_m = (a.M + BM - 1) / BM; p.tiles_n = (a.N + BN - 1) / BN; p.num_kb = (a.K + BK - 1) / BK;
  p.b_batched = b_batched ? 1 : 0;
  p.alpha = a.alpha; p.bias = a.bias; p.bias_along_m = a.bias_along_m; p.act = a.act; p.gamma = a.gamma;
  p.has_res = a.residual ? 1 : 0; p.res_batched = res_batched ? 1 : 0; p.res_row_mod = a.res_row_mod;
  p.out_is_bf16 = out_bf16;
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("SAM2B_GEMM_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
  const long tiles = (long)p.tiles_m * p.tiles_n * p.batch;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  char nm[96];
  if (g_prof_on) snprintf(nm, sizeof(nm), "gemm m%d n%d k%d b%d bn%d md%d act%d%s%s", a.M, a.N, a.K, a.batch, BN, MODE, a.act, a.residual ? " res" : "", a.out_bf16 ? " o16" : " o32");
  const double esz = a.out_bf16 ? 2.0 : 4.0;
  ProfRange pr(stream, nm, 2.0 * a.M * a.N * (double)a.K * a.batch,
               ((double)a.M * a.K * 2 + (b_batched ? 1.0 : 1.0 / a.batch) * (double)a.N * a.K * 2 + (double)a.M * a.N * (esz + (a.residual ? 4.0 : 0.0))) * a.batch);
  gemm_bf16_tcgen05_kernel<BN, MODE><<<grid, NUM_THREADS, Cfg::SMEM, stream>>>(ta, tb, tc, tr, p);
  SAM2B_CHECK_LAUNCH("gemm_bf16_tcgen05");
  return 0;
}


// One translation unit per BLOCK_N (parallel compilation): the specialised modes for the large tiles, generic otherwise.
#define SAM2B_GEMM_INSTANTIATE_FULL(BN_)                                                     \
  int launch_gemm_bn##BN_(const GemmArgs& a, int mode, cudaStream_t stream) {                 \
    switch (mode) {                                                                          \
      case MODE_BF16_BIAS: return launch_gemm<BN_, MODE_BF16_BIAS>(a, stream);               \
      case MODE_BF16_BIAS_GELU: return launch_gemm<BN_, MODE_BF16_BIAS_GELU>(a, stream);     \
      case MODE_F32_BIAS_RES: return launch_gemm<BN_, MODE_F32_BIAS_RES>(a, stream);         \
      default: return launch_gemm<BN_, MODE_GENERIC>(a, stream);                             \
    }                                                                                        \
  }
#define SAM2B_GEMM_INSTANTIATE_GENERIC(BN_)                                                  \
  int launch_gemm_bn##BN_(const GemmArgs& a, int, cudaStream_t stream) { return launch_gemm<BN_, MODE_GENERIC>(a, stream); }

int launch_gemm_bn256(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn224(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn192(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn160(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn128(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn96(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn64(const GemmArgs& a, int mode, cudaStream_t stream);
int launch_gemm_bn32(const GemmArgs& a, int mode, cudaStream_t stream);

}  // namespace sam2b
