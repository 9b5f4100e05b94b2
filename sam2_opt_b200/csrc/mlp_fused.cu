// Fused trunk MLP for the narrow stages (C = 144, 288): out = x + fc2(GELU(fc1(LN(x)))) in one kernel per 128 rows.
// Reference: MLP.forward (sam2/modeling/sam2_utils.py:131-136) inside MultiScaleBlock.forward (backbones/hieradet.py:160-166),
// with norm2 folded into fc1 as everywhere else in the trunk (weights.py: row-centred W diag(gamma), b + W beta).
//
// Why: in stages 1 and 2 (K = 144 / 288) the two MLP GEMMs are not tensor bound - fc1's GELU epilogue is issue bound and
// the 4C-wide hidden tensor is written to and read back from HBM (1.2 GB / 0.6 GB per block per 16 images).  Here the
// hidden activations never leave the SM:
//   for every 96-wide chunk h of the hidden layer
//     S_b[128 x 96] = A[128 x C] W1[h]^T                  tcgen05.mma, A tile resident in shared memory, W1 streamed by TMA;
//                                                         b = h & 1: two S buffers, chunk h + 1 is computed while
//     P             = bf16(GELU(rstd * S_b + b1[h]))       the epilogue warps work on chunk h; P is written back over S_b's
//                                                         columns as packed pairs
//     O[128 x C]   += P W2[:, h]^T                         tcgen05.mma with the A operand (P) read from TENSOR MEMORY
//   x_out = O + b2 + x; bf16 copy and per-row (sum, sum of squares) for the next block's folded LayerNorm
// TMEM: S 2 x 96 columns + O C columns (<= 480).  Warps: 0 and 3 TMA producers (A + W1 / W2), 1 MMA issuer, 2 TMEM allocation, 4-11 epilogue:
// lane quarter = warp & 3; group (warp - 4) >> 2 takes the even / odd hidden chunks and one half of O's columns.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

namespace fm {
static constexpr int BM = 128;
static constexpr int BH = 96;                        // hidden chunk
static constexpr int THREADS = 384;
static constexpr int A_KB_BYTES = BM * 128;          // [128 rows][64 k] bf16, 128B swizzle
static constexpr int W1_BYTES = BH * 128;            // [96 hidden][64 k]
static constexpr int COL_S = 0;                      // S buffer b at COL_S + 96 b; P over its first 48 columns
static constexpr int COL_O = 192;
template <int C>
struct Cfg {
  static constexpr int KB = (C + 63) / 64;           // k-blocks of the A tile (the last one partly zero-filled by TMA)
  static constexpr int NH = 4 * C / BH;              // hidden chunks
  static constexpr int NSPLIT = C / 144;             // fc2 instructions per k-step (N = 144 each; tcgen05.mma N <= 256)
  static constexpr int W2_BYTES = C * 128;           // [C rows][64 hidden]
  static constexpr int NB2 = 4 * C / 64;             // 64-wide W2 blocks per tile (a block may straddle two hidden chunks)
  static constexpr int NS1 = C == 144 ? 4 : 3;       // W1 ring depth
  static constexpr int NS2 = 3;                      // W2 ring depth
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W1 = OFF_A + KB * A_KB_BYTES;
  static constexpr int OFF_W2 = OFF_W1 + NS1 * W1_BYTES;
  static constexpr int OFF_BAR = OFF_W2 + NS2 * W2_BYTES;
  static constexpr int SMEM = OFF_BAR + 256 + 1024;
  static_assert(4 * C % BH == 0 && C % 144 == 0 && NH % 2 == 0, "C must be 144 or 288");
  static_assert(COL_O + C <= 512, "TMEM columns");
};
}  // namespace fm

struct FusedMlpParams {
  const float* stats_in;     // [M][parts_in][2] partial (sum, sum of squares) of the input rows
  int parts_in;
  float eps, inv_c;
  const float* b1;           // [4C]  (b + W beta)
  const float* b2;           // [C]
  float* x;                  // [M][C] fp32 residual stream, updated in place
  bf16* xb_out;              // [M][C] bf16 copy of the new stream
  float* stats_out;          // [M][2][2]
  int M, num_tiles;
};

template <int C>
__global__ void __launch_bounds__(fm::THREADS, 1)
mlp_fused_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w1,
                 const __grid_constant__ CUtensorMap map_w2, const FusedMlpParams p) {
  using namespace fm;
  using K = Cfg<C>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + K::OFF_BAR);
  uint64_t* a_full = bars + 0;
  uint64_t* a_empty = bars + 1;
  uint64_t* w1_full = bars + 2;                 // [NS1]
  uint64_t* w1_empty = w1_full + K::NS1;
  uint64_t* w2_full = w1_empty + K::NS1;        // [NS2]
  uint64_t* w2_empty = w2_full + K::NS2;
  uint64_t* s_full = w2_empty + K::NS2;         // [2]
  uint64_t* p_full = s_full + 2;                // [2]
  uint64_t* o_full = p_full + 2;
  uint64_t* o_free = o_full + 1;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w1);
    tma_prefetch_desc(&map_w2);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(a_full, 1); mbar_init(a_empty, 1);
    for (int i = 0; i < K::NS1; ++i) { mbar_init(&w1_full[i], 1); mbar_init(&w1_empty[i], 1); }
    for (int i = 0; i < K::NS2; ++i) { mbar_init(&w2_full[i], 1); mbar_init(&w2_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4); }
    mbar_init(o_full, 1); mbar_init(o_free, 8);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_ptr, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer 1: the A tile and the W1 ring =====================
    if (lane == 0) {
      int i1 = 0, nt = 0;
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
        mbar_wait(a_empty, (nt & 1) ^ 1);
        mbar_arrive_expect_tx(a_full, K::KB * A_KB_BYTES);
#pragma unroll
        for (int kb = 0; kb < K::KB; ++kb) tma_load_3d(smem + K::OFF_A + kb * A_KB_BYTES, &map_a, a_full, kb * 64, t * BM, 0);
        for (int h = 0; h < K::NH; ++h) {
          for (int kb = 0; kb < K::KB; ++kb, ++i1) {
            const int s = i1 % K::NS1;
            mbar_wait(&w1_empty[s], ((i1 / K::NS1) & 1) ^ 1);
            mbar_arrive_expect_tx(&w1_full[s], W1_BYTES);
            tma_load_3d(smem + K::OFF_W1 + s * W1_BYTES, &map_w1, &w1_full[s], kb * 64, h * BH, 0);
          }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== TMA producer 2: the W2 ring (its own thread: a full W2 ring must not hold back W1) =====================
    if (lane == 0) {
      int i2 = 0;
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
        for (int blk = 0; blk < K::NB2; ++blk, ++i2) {
          const int s = i2 % K::NS2;
          mbar_wait(&w2_empty[s], ((i2 / K::NS2) & 1) ^ 1);
          mbar_arrive_expect_tx(&w2_full[s], K::W2_BYTES);
#pragma unroll
          for (int j = 0; j < K::NSPLIT; ++j)
            tma_load_3d(smem + K::OFF_W2 + s * K::W2_BYTES + j * 144 * 128, &map_w2, &w2_full[s], blk * 64, j * 144, 0);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc_fc1 = make_idesc_ex(128, BH, 0);
      constexpr uint32_t idesc_fc2 = make_idesc_ex(128, 144, 0);
      const uint32_t sbase = smem_u32(smem);
      int i1 = 0, i2 = 0, nt = 0;
      uint32_t n_s[2] = {0, 0};         // fc1 products issued on S buffer b
      uint32_t n_p[2] = {0, 0};         // hidden chunks handed over through buffer b (phase of p_full[b])
      auto issue_fc1 = [&](int h) {
        const int b = h & 1;
        for (int kb = 0; kb < K::KB; ++kb, ++i1) {
          const int s = i1 % K::NS1;
          mbar_wait(&w1_full[s], (i1 / K::NS1) & 1);
          tc_fence_after();
          const uint32_t a = sbase + K::OFF_A + kb * A_KB_BYTES, w = sbase + K::OFF_W1 + s * W1_BYTES;
          const int ksteps = (C - kb * 64) >= 64 ? 4 : (C - kb * 64) / 16;
          for (int k = 0; k < ksteps; ++k)
            umma_bf16(tmem_base + COL_S + BH * b, make_smem_desc(a + k * 32, 16, 1024, 2), make_smem_desc(w + k * 32, 16, 1024, 2),
                      idesc_fc1, (kb | k) != 0);
          umma_commit(&w1_empty[s]);
        }
        umma_commit(&s_full[b]);
        ++n_s[b];
      };
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
        mbar_wait(a_full, nt & 1);
        tc_fence_after();
        // S_b is reused by chunk h + 2: in order behind fc2 of chunk h, which read P(h) out of it after the epilogue warps had
        // read S(h) - no extra barrier needed
        issue_fc1(0);
        for (int h = 0; h < K::NH; ++h) {
          if (h + 1 < K::NH) issue_fc1(h + 1);                 // overlaps the GELU pass over chunk h
          if (h + 1 >= K::NH - 1 && h + 1 <= K::NH - 1) umma_commit(a_empty);   // last fc1 of the tile has been issued
          const int b = h & 1;
          mbar_wait(&p_full[b], n_p[b] & 1);
          ++n_p[b];
          if (h == 0) mbar_wait(o_free, (nt & 1) ^ 1);          // the previous tile's O has been read out
          tc_fence_after();
#pragma unroll 1
          for (int k = 0; k < BH / 16; ++k) {
            const int g = h * (BH / 16) + k;                    // k-step of this tile's fc2: 64-wide W2 block g / 4, step g % 4
            const int st = g & 3;
            const int s = (i2 + (g >> 2)) % K::NS2;
            if (st == 0 || k == 0) {
              if (st == 0) {                                      // first use of the block
                mbar_wait(&w2_full[s], ((i2 + (g >> 2)) / K::NS2) & 1);
                tc_fence_after();
              }
            }
            const uint32_t w = sbase + K::OFF_W2 + s * K::W2_BYTES;
            const uint32_t pa = tmem_base + COL_S + BH * b + k * 8;            // 16 hidden units = 8 columns of bf16 pairs
#pragma unroll
            for (int j = 0; j < K::NSPLIT; ++j)
              umma_bf16_ts(tmem_base + COL_O + j * 144, pa, make_smem_desc(w + j * 144 * 128 + st * 32, 16, 1024, 2), idesc_fc2,
                           (h | k) != 0);
            if (st == 3) umma_commit(&w2_empty[s]);
          }
        }
        i2 += K::NB2;
        umma_commit(o_full);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue warps =====================
    const int q = warp & 3, half = (warp - 4) >> 2;          // half: hidden chunks of this parity, and this half of O's columns
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    uint32_t n_s = 0;                                         // chunks this group has taken (phase of s_full[half])
    int nt = 0;
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
      const long row = (long)t * BM + q * 32 + lane;
      const bool live = row < p.M;
      float rstd = 1.f;
      if (live) {
        const float2* st = reinterpret_cast<const float2*>(p.stats_in) + row * p.parts_in;
        float s1 = 0.f, s2 = 0.f;
        for (int k = 0; k < p.parts_in; ++k) { const float2 v = __ldg(st + k); s1 += v.x; s2 += v.y; }
        const float mu = s1 * p.inv_c;
        rstd = rsqrtf(fmaxf(fmaf(s2, p.inv_c, -mu * mu), 0.f) + p.eps);
      }
      const uint32_t ts = tmem_base + COL_S + BH * half + lane_sel;
      for (int h = half; h < K::NH; h += 2, ++n_s) {
        mbar_wait(&s_full[half], n_s & 1);
        tc_fence_after();
        uint32_t pk[64];                                  // this row's 96 hidden units of the chunk as 48 bf16 pairs (+ slack)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          uint32_t r[32];
          tmem_ld_x32(ts + c * 32, r);
          tmem_ld_wait();
          const float4* bsrc = reinterpret_cast<const float4*>(p.b1 + h * BH + c * 32);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(bsrc + j);
            const float v0 = gelu_fast_bf16(fmaf(__uint_as_float(r[4 * j]), rstd, b.x));
            const float v1 = gelu_fast_bf16(fmaf(__uint_as_float(r[4 * j + 1]), rstd, b.y));
            const float v2 = gelu_fast_bf16(fmaf(__uint_as_float(r[4 * j + 2]), rstd, b.z));
            const float v3 = gelu_fast_bf16(fmaf(__uint_as_float(r[4 * j + 3]), rstd, b.w));
            pk[c * 16 + 2 * j] = pack_bf16x2(v0, v1);
            pk[c * 16 + 2 * j + 1] = pack_bf16x2(v2, v3);
          }
        }
        // P over the first 48 columns of the buffer this warp has just read (same lanes, same warp: no cross-warp hazard)
        tmem_st_x32(ts, *reinterpret_cast<uint32_t(*)[32]>(&pk[0]));
        tmem_st_x16(ts + 32, *reinterpret_cast<uint32_t(*)[32]>(&pk[32]));
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[half]);
      }
      // ---- tile done: x = O + b2 + x, bf16 copy, row statistics ----
      mbar_wait(o_full, nt & 1);
      tc_fence_after();
      constexpr int NC = C / 16, NC0 = (NC + 1) / 2;
      const int c0 = half == 0 ? 0 : NC0, c1 = half == 0 ? NC0 : NC;
      float* xr = p.x + row * C;
      bf16* xbr = p.xb_out + row * C;
      float s1 = 0.f, s2 = 0.f;
#pragma unroll 3
      for (int c = c0; c < c1; ++c) {
        uint32_t r[32];
        tmem_ld_x16(tmem_base + COL_O + c * 16 + lane_sel, r);
        tmem_ld_wait();
        if (c == c1 - 1) {                                // O has left tensor memory: the next tile's fc2 may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(o_free);
        }
        if (live) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.b2 + c * 16) + j);
            const float4 x = ld_global_f4(xr + c * 16 + 4 * j);
            v[4 * j] = __uint_as_float(r[4 * j]) + b.x + x.x;
            v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) + b.y + x.y;
            v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) + b.z + x.z;
            v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) + b.w + x.w;
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) { s1 += v[j]; s2 = fmaf(v[j], v[j], s2); }
#pragma unroll
          for (int j = 0; j < 4; ++j) st_global_f4(xr + c * 16 + 4 * j, make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]));
          uint4 lo, hi;
          lo.x = pack_bf16x2(v[0], v[1]); lo.y = pack_bf16x2(v[2], v[3]); lo.z = pack_bf16x2(v[4], v[5]); lo.w = pack_bf16x2(v[6], v[7]);
          hi.x = pack_bf16x2(v[8], v[9]); hi.y = pack_bf16x2(v[10], v[11]); hi.z = pack_bf16x2(v[12], v[13]); hi.w = pack_bf16x2(v[14], v[15]);
          *reinterpret_cast<uint4*>(xbr + c * 16) = lo;
          *reinterpret_cast<uint4*>(xbr + c * 16 + 8) = hi;
        }
      }
      if (live) reinterpret_cast<float2*>(p.stats_out)[row * 2 + half] = make_float2(s1, s2);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFnM)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFnM fm_encode_fn() {
  static EncodeTiledFnM fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFnM>(f);
  });
  return fn;
}
// [rows][cols] bf16 row-major (pitch ld), box {64, box_rows, 1}, 128B swizzle; out-of-range elements read as zero
static int fm_map(CUtensorMap* m, const bf16* ptr, long rows, long cols, long ld, int box_rows) {
  EncodeTiledFnM enc = fm_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * 2) % 16) { set_error("mlp_fused: operand not 16 B aligned"); return 1; }
  const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 1};
  const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * (cuuint64_t)rows};
  const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<bf16*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("mlp_fused: cuTensorMapEncodeTiled failed (%d)", (int)r); return 1; }
  return 0;
}

template <int C>
static int launch_mlp_fused(const CUtensorMap& ma, const CUtensorMap& m1, const CUtensorMap& m2, const FusedMlpParams& p, cudaStream_t stream) {
  using K = fm::Cfg<C>;
  static PerDeviceFlag attr;
  if (!attr.is_set()) {
    cudaError_t e = cudaFuncSetAttribute(mlp_fused_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(mlp_fused<%d>): %s", C, cudaGetErrorString(e)); return 1; }
    attr.set();
  }
  const int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  mlp_fused_kernel<C><<<grid, fm::THREADS, K::SMEM, stream>>>(ma, m1, m2, p);
  return 0;
}

bool mlp_fused_supported(int C) { return C == 144 || C == 288; }

// x[M][C] (fp32, in place) += fc2(GELU(rstd(row) * (xb W1''^T) + b1)) + b2; xb_out = bf16(x); stats_out[M][2][2]
int mlp_fused(const bf16* xb, const float* stats_in, int parts_in, float eps, const bf16* w1, const float* b1, const bf16* w2,
              const float* b2, float* x, bf16* xb_out, float* stats_out, long M, int C, cudaStream_t stream) {
  if (!mlp_fused_supported(C) || M <= 0) { set_error("mlp_fused: unsupported C=%d M=%ld", C, M); return 1; }
  CUtensorMap ma, m1, m2;
  if (fm_map(&ma, xb, M, C, C, fm::BM) || fm_map(&m1, w1, 4L * C, C, C, fm::BH) || fm_map(&m2, w2, C, 4L * C, 4L * C, 144)) return 1;
  FusedMlpParams p;
  p.stats_in = stats_in; p.parts_in = parts_in; p.eps = eps; p.inv_c = 1.0f / (float)C;
  p.b1 = b1; p.b2 = b2; p.x = x; p.xb_out = xb_out; p.stats_out = stats_out;
  p.M = (int)M; p.num_tiles = (int)((M + fm::BM - 1) / fm::BM);
  char nm[64];
  if (g_prof_on) snprintf(nm, sizeof(nm), "mlp_fused m%ld c%d", M, C);
  ProfRange pr(stream, nm, 16.0 * (double)M * C * C, (double)M * C * (2.0 + 4.0 + 4.0 + 2.0));
  const int rc = C == 144 ? launch_mlp_fused<144>(ma, m1, m2, p, stream) : launch_mlp_fused<288>(ma, m1, m2, p, stream);
  if (rc) return rc;
  SAM2B_CHECK_LAUNCH("mlp_fused");
  return 0;
}

}  // namespace sam2b
