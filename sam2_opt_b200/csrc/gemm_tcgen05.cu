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
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

static constexpr int BM = 128;      // UMMA M (cta_group::1)
static constexpr int BK = 64;       // 64 bf16 = 128 B = one swizzle row
static constexpr int UMMA_K = 16;
static constexpr int ACC_STRIDE = 256;   // TMEM columns between the two accumulator stages
static constexpr int TMEM_COLS = 512;
static constexpr int NUM_EPI_WARPS = 8;  // two warps per TMEM lane quarter, interleaved over 32-column chunks
static constexpr int NUM_THREADS = 128 + 32 * NUM_EPI_WARPS;  // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-11 epilogue
static constexpr int EPI_STAGE_BYTES = 32 * 32 * 4;           // one 32x32 fp32 transpose buffer per epilogue warp (XOR swizzled)
static constexpr int EPI_SMEM = NUM_EPI_WARPS * EPI_STAGE_BYTES;
static constexpr int SMEM_TOTAL = 227 * 1024;

struct GemmKernelParams {
  int M, N, K, batch;
  int tiles_m, tiles_n, num_kb;
  int b_batched;
  float alpha;
  const float* bias; int bias_along_m;
  int act;
  const float* gamma;
  const float* residual; long ld_res; long res_batch_stride; int res_row_mod;
  float* out_f32; bf16* out_bf16; long ldc; long c_batch_stride;
  int dbg;   // timing experiments only (SAM2B_GEMM_DBG): 1 = skip global stores, 2 = skip the smem transpose
};

template <int BN>
struct GemmCfg {
  static constexpr int STAGE_A = BM * BK * 2;
  static constexpr int STAGE_B = BN * BK * 2;
  static constexpr int STAGE = STAGE_A + STAGE_B;
  static constexpr int AVAIL = SMEM_TOTAL - EPI_SMEM - 1024 /*align slack*/ - 256 /*barriers*/;
  static constexpr int STAGES = (AVAIL / STAGE) > 8 ? 8 : (AVAIL / STAGE);
  static constexpr int SMEM = STAGES * STAGE + EPI_SMEM + 1024 + 256;
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

template <int BN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const GemmKernelParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * Cfg::STAGE_A;
  uint8_t* smem_epi = smem + STAGES * Cfg::STAGE;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + EPI_SMEM);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n * p.batch;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], NUM_EPI_WARPS); }
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
        const int acc = local & 1;
        const uint32_t acc_phase = (local >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * ACC_STRIDE;
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
    // ===================== epilogue: TMEM -> registers -> smem transpose -> coalesced global =====================
    // Warp (4+we) may read TMEM lanes 32*(we&3)..+31 (hardware rule: warp id mod 4).  The two warps of a lane quarter
    // take alternate 32-column chunks.  tcgen05.ld gives thread i row i of the chunk; the XOR-swizzled staging buffer
    // turns that into "8 lanes x float4 = one 128 B row segment", so every global load/store instruction of the warp
    // covers 4 full 128 B lines (fp32) instead of 32 scattered 16 B pieces.  Every epilogue option is a warp-uniform
    // branch taken once per 32x32 chunk (not per element), and addresses advance by precomputed strides.
    const int we = warp - 4;
    const int q = we & 3;
    const int half = we >> 2;
    const uint32_t stg_base = smem_u32(smem_epi + we * EPI_STAGE_BYTES);
    const int lr = lane >> 3;            // row inside a group of 4 rows (transposed layout)
    const int lc = lane & 7;             // float4 column inside the 32-column chunk
    const uint32_t sts_base = stg_base + lane * 128;                       // row `lane`, slot j ^ (lane & 7)
    const uint32_t sts_xor = (uint32_t)(lane & 7) << 4;
    // transposed read: row rl = 4k + lr, slot lc ^ (rl & 7); (4k + lr) & 7 = lr + 4 (k & 1)
    const uint32_t lds_even = stg_base + lr * 128 + ((uint32_t)(lc ^ lr) << 4);          // k even
    const uint32_t lds_odd = stg_base + (lr + 4) * 128 + ((uint32_t)(lc ^ (lr + 4)) << 4);  // k odd
    constexpr int NCHUNK = (BN + 31) / 32;
    const bool has_res = p.residual != nullptr;
    const bool bias_n = p.bias != nullptr && !p.bias_along_m;
    const bool bias_m_on = p.bias != nullptr && p.bias_along_m;
    const long row_step_c = 4 * p.ldc;       // elements between this lane's consecutive rows
    const long row_step_r = 4 * p.ld_res;
    int local = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++local) {
      const int n_blk = t % p.tiles_n;
      const int m_blk = (t / p.tiles_n) % p.tiles_m;
      const int b = t / (p.tiles_n * p.tiles_m);
      const int acc = local & 1;
      const uint32_t acc_phase = (local >> 1) & 1;
      const int row0 = m_blk * BM + q * 32 + lr;           // this lane's rows: row0 + 4k
      const int ncol0 = n_blk * BN + lc * 4;               // this lane's columns in chunk c: ncol0 + 32c .. +3
      uint32_t row_mask = 0;                               // bit k: row0 + 4k < M
#pragma unroll
      for (int k = 0; k < 8; ++k) row_mask |= (row0 + 4 * k < p.M) ? (1u << k) : 0u;
      const long out_off0 = (long)b * p.c_batch_stride + (long)row0 * p.ldc + ncol0;
      const float* res_ptr0 = nullptr;
      if (has_res) {
        // res_row_mod (broadcast of a [mod, N] table over row blocks) must be a multiple of BM: one modulo per tile
        const int rbase = p.res_row_mod ? ((m_blk * BM) % p.res_row_mod) + q * 32 + lr : row0;
        res_ptr0 = p.residual + (long)b * p.res_batch_stride + (long)rbase * p.ld_res + ncol0;
      }
      float4 res[8];
      auto load_res = [&](int c) {
        const bool col_ok = (c * 32 + lc * 4 < BN) && (ncol0 + c * 32 + 4 <= p.N);
        const float* rp = res_ptr0 + c * 32;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          res[k] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (col_ok && ((row_mask >> k) & 1)) res[k] = ld_global_f4(rp);
          rp += row_step_r;
        }
      };
      if (has_res && half < NCHUNK) load_res(half);       // overlaps the wait for the accumulator
      float bias_m[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (bias_m_on) {
#pragma unroll
        for (int k = 0; k < 8; ++k) bias_m[k] = ((row_mask >> k) & 1) ? __ldg(p.bias + row0 + 4 * k) : 0.f;
      }
      // bias / gamma of this lane's 4 columns: fetched one chunk ahead so the load latency never sits on the critical path
      auto load_vec = [&](const float* base, int c, float fill) {
        const int col = ncol0 + c * 32;
        float4 v = make_float4(fill, fill, fill, fill);
        if (base && (c * 32 + lc * 4 < BN) && (col + 4 <= p.N)) v = __ldg(reinterpret_cast<const float4*>(base + col));
        return v;
      };
      float4 bn_next = load_vec(bias_n ? p.bias : nullptr, half, 0.f);
      float4 gm_next = load_vec(p.gamma, half, 1.f);
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + acc * ACC_STRIDE + ((uint32_t)(q * 32) << 16);
      uint32_t r[32];
      if (half < NCHUNK) tmem_ld_x32(taddr + half * 32, r);   // columns past BN (last chunk of BN=144) are never stored
#pragma unroll 1
      for (int c = half; c < NCHUNK; c += 2) {
        tmem_ld_wait();
        if (!(p.dbg & 2)) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          st_shared_u4(sts_base + (((uint32_t)j << 4) ^ sts_xor), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        }
        float4 v[8];
        if (p.dbg & 2) {
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = make_float4(__uint_as_float(r[4 * k]), __uint_as_float(r[4 * k + 1]), __uint_as_float(r[4 * k + 2]), __uint_as_float(r[4 * k + 3]));
        }
        __syncwarp();
        if (c + 2 < NCHUNK) {
          tmem_ld_x32(taddr + (c + 2) * 32, r);     // next chunk streams out of TMEM while this one is processed
        } else {
          // last chunk of this warp has left TMEM: hand the accumulator stage back to the MMA warp now
          tc_fence_before();
          if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        }
        const float4 bn = bn_next, gm = gm_next;
        if (c + 2 < NCHUNK) {
          bn_next = load_vec(bias_n ? p.bias : nullptr, c + 2, 0.f);
          gm_next = load_vec(p.gamma, c + 2, 1.f);
        }
        if (!(p.dbg & 2)) {
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = ld_shared_f4(((k & 1) ? lds_odd : lds_even) + (k >> 1) * 1024);
        }
        __syncwarp();
        const int col = ncol0 + c * 32;
        const bool in_tile = c * 32 + lc * 4 < BN;
        if (in_tile && col + 4 <= p.N) {
          {
            const float al = p.alpha;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              v[k].x = fmaf(v[k].x, al, bn.x + bias_m[k]); v[k].y = fmaf(v[k].y, al, bn.y + bias_m[k]);
              v[k].z = fmaf(v[k].z, al, bn.z + bias_m[k]); v[k].w = fmaf(v[k].w, al, bn.w + bias_m[k]);
            }
          }
          if (p.act == ACT_GELU) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              v[k].x = gelu_fast(v[k].x); v[k].y = gelu_fast(v[k].y); v[k].z = gelu_fast(v[k].z); v[k].w = gelu_fast(v[k].w);
            }
          } else if (p.act == ACT_RELU) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              v[k].x = fmaxf(v[k].x, 0.f); v[k].y = fmaxf(v[k].y, 0.f); v[k].z = fmaxf(v[k].z, 0.f); v[k].w = fmaxf(v[k].w, 0.f);
            }
          }
          if (p.gamma) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { v[k].x *= gm.x; v[k].y *= gm.y; v[k].z *= gm.z; v[k].w *= gm.w; }
          }
          if (has_res) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { v[k].x += res[k].x; v[k].y += res[k].y; v[k].z += res[k].z; v[k].w += res[k].w; }
            if (c + 2 < NCHUNK) load_res(c + 2);   // next chunk's residual in flight during the stores
          }
          if (p.dbg & 1) {
            if (v[0].x + v[1].y + v[2].z + v[3].w + v[4].x + v[5].y + v[6].z + v[7].w == 123.456f) p.out_f32[0] = 0.f;
          } else if (p.out_f32) {
            float* op = p.out_f32 + out_off0 + c * 32;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if ((row_mask >> k) & 1) st_global_f4(op, v[k]);
              op += row_step_c;
            }
          }
          if (!(p.dbg & 1) && p.out_bf16) {
            bf16* op = p.out_bf16 + out_off0 + c * 32;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              if ((row_mask >> k) & 1) st_global_u2(op, pack_bf16x2(v[k].x, v[k].y), pack_bf16x2(v[k].z, v[k].w));
              op += row_step_c;
            }
          }
        } else {
          if (in_tile && col < p.N) {
            // ragged N tail (N % 4 != 0 or the tile edge cuts this float4): scalar path
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const int row = row0 + 4 * k;
              if (row < p.M) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  if (col + e < p.N) {
                    float o = (e == 0 ? v[k].x : e == 1 ? v[k].y : e == 2 ? v[k].z : v[k].w) * p.alpha;
                    if (p.bias) o += p.bias_along_m ? __ldg(p.bias + row) : __ldg(p.bias + col + e);
                    if (p.act == ACT_GELU) o = gelu_fast(o); else if (p.act == ACT_RELU) o = fmaxf(o, 0.f);
                    if (p.gamma) o *= __ldg(p.gamma + col + e);
                    if (has_res) o += res_ptr0[c * 32 + (long)k * row_step_r + e];
                    const long off = out_off0 + c * 32 + (long)k * row_step_c + e;
                    if (p.out_f32) p.out_f32[off] = o;
                    if (p.out_bf16) p.out_bf16[off] = __float2bfloat16(o);
                  }
                }
              }
            }
          }
          if (has_res && c + 2 < NCHUNK) load_res(c + 2);
        }
      }
      if (half >= NCHUNK) {   // BN = 32: the second warp of each quarter has no chunk but still owes its arrival
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  return fn;
}

// 3-D bf16 tensor map over [batch][rows][K] (K contiguous), box = {64, box_rows, 1}, 128B swizzle, zero OOB fill.
static int make_tmap(CUtensorMap* m, const bf16* ptr, long rows, long K, long ld, long batch, long batch_stride,
                     int box_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * 2) % 16 || (batch > 1 && (batch_stride * 2) % 16)) {
    set_error("gemm operand not 16B aligned (ptr=%p ld=%ld batch_stride=%ld)", (const void*)ptr, ld, batch_stride);
    return 1;
  }
  cuuint64_t dims[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)(batch > 0 ? batch : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)(batch > 1 ? batch_stride : ld * rows) * 2};
  cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<bf16*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) rows=%ld K=%ld ld=%ld batch=%ld bs=%ld box_rows=%d", (int)r, rows, K,
              ld, batch, batch_stride, box_rows);
    return 1;
  }
  return 0;
}

template <int BN>
static int launch_gemm(const GemmArgs& a, cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(gemm_bf16_tcgen05_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(gemm<%d>): %s", BN, cudaGetErrorString(e)); return 1; }
    attr_set = true;
  }
  CUtensorMap ta, tb;
  if (make_tmap(&ta, a.A, a.M, a.K, a.lda, a.batch, a.a_batch_stride, BM)) return 1;
  const bool b_batched = a.batch > 1 && a.b_batch_stride != 0;
  if (make_tmap(&tb, a.B, a.N, a.K, a.ldb, b_batched ? a.batch : 1, a.b_batch_stride, BN)) return 1;
  GemmKernelParams p;
  p.M = a.M; p.N = a.N; p.K = a.K; p.batch = a.batch;
  p.tiles_m = (a.M + BM - 1) / BM; p.tiles_n = (a.N + BN - 1) / BN; p.num_kb = (a.K + BK - 1) / BK;
  p.b_batched = b_batched ? 1 : 0;
  p.alpha = a.alpha; p.bias = a.bias; p.bias_along_m = a.bias_along_m; p.act = a.act; p.gamma = a.gamma;
  p.residual = a.residual; p.ld_res = a.ld_res; p.res_batch_stride = a.res_batch_stride; p.res_row_mod = a.res_row_mod;
  p.out_f32 = a.out_f32; p.out_bf16 = a.out_bf16; p.ldc = a.ldc; p.c_batch_stride = a.c_batch_stride;
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("SAM2B_GEMM_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
  const long tiles = (long)p.tiles_m * p.tiles_n * p.batch;
  const int grid = (int)(tiles < num_sms() ? tiles : num_sms());
  char nm[96];
  if (g_prof_on) snprintf(nm, sizeof(nm), "gemm m%d n%d k%d b%d bn%d act%d%s%s", a.M, a.N, a.K, a.batch, BN, a.act, a.residual ? " res" : "", a.out_bf16 ? " o16" : " o32");
  const double esz = a.out_bf16 ? 2.0 : 4.0;
  ProfRange pr(stream, nm, 2.0 * a.M * a.N * (double)a.K * a.batch,
               ((double)a.M * a.K * 2 + (b_batched ? 1.0 : 1.0 / a.batch) * (double)a.N * a.K * 2 + (double)a.M * a.N * (esz + (a.residual ? 4.0 : 0.0))) * a.batch);
  gemm_bf16_tcgen05_kernel<BN><<<grid, NUM_THREADS, Cfg::SMEM, stream>>>(ta, tb, p);
  SAM2B_CHECK_LAUNCH("gemm_bf16_tcgen05");
  return 0;
}

static int pick_bn(const GemmArgs& a) {
  static const int cand[] = {256, 192, 144, 128, 96, 64, 32};
  const long tiles_m = (a.M + BM - 1) / BM;
  double best = 1e30; int best_bn = 128;
  for (int bn : cand) {
    const long tiles = tiles_m * ((a.N + bn - 1) / bn) * a.batch;
    const long waves = (tiles + num_sms() - 1) / num_sms();
    // per-tile cost ~ MMA time (prop. to BN) + fixed per-tile overhead (epilogue drain, barriers)
    const double cost = (double)waves * (bn + 64.0);
    if (cost < best - 1e-9) { best = cost; best_bn = bn; }
  }
  return best_bn;
}

int gemm_bf16(const GemmArgs& a, cudaStream_t stream) {
  if (a.M <= 0 || a.N <= 0 || a.K <= 0 || a.batch <= 0) { set_error("gemm: empty problem M=%d N=%d K=%d batch=%d", a.M, a.N, a.K, a.batch); return 1; }
  if (!a.out_f32 && !a.out_bf16) { set_error("gemm: no output"); return 1; }
  if ((a.out_f32 && (a.ldc % 4 || (reinterpret_cast<uintptr_t>(a.out_f32) & 15))) ||
      (a.out_bf16 && (a.ldc % 8 || (reinterpret_cast<uintptr_t>(a.out_bf16) & 15))) ||
      (a.residual && (a.ld_res % 4 || (reinterpret_cast<uintptr_t>(a.residual) & 15)))) {
    set_error("gemm: output/residual leading dimension or pointer not vector aligned (ldc=%ld ld_res=%ld)", a.ldc, a.ld_res);
    return 1;
  }
  if (a.residual && a.res_row_mod && (a.res_row_mod % BM)) {
    set_error("gemm: res_row_mod=%d must be a multiple of %d", a.res_row_mod, BM);
    return 1;
  }
  const int bn = a.force_bn ? a.force_bn : pick_bn(a);
  switch (bn) {
    case 256: return launch_gemm<256>(a, stream);
    case 192: return launch_gemm<192>(a, stream);
    case 144: return launch_gemm<144>(a, stream);
    case 128: return launch_gemm<128>(a, stream);
    case 96: return launch_gemm<96>(a, stream);
    case 64: return launch_gemm<64>(a, stream);
    case 32: return launch_gemm<32>(a, stream);
  }
  set_error("gemm: unsupported BLOCK_N %d", bn);
  return 1;
}

}  // namespace sam2b
