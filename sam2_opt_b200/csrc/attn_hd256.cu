// K4: flash attention for the MemoryAttention layers (single head, head_dim 256) on tcgen05 / TMEM / TMA.
// Reference: RoPEAttention.forward / Attention.forward (sam2/modeling/sam/transformer.py:265-296,345-424):
// softmax(q k^T / 16) v with q [4096,256] and k, v [Lk,256], Lk = 4096 (self attention) or L*4096 + P (cross attention
// over the memory bank, up to 28 736 keys).  q / k arrive already projected and rotated (bf16), v projected (bf16);
// the score and probability matrices (470 MB per object and layer in the reference) never leave the SM.
//
// One persistent CTA works on units of (object, 128-query tile, KV split):
//   warp 0     TMA producer: Q tile (4 sub-tiles [128][64], 128B swizzle), then K and V tiles of 64 keys
//              (4 sub-tiles [64][64] each) through a 2-stage ring.  Rows past Lk are zero-filled by TMA.
//   warp 1     MMA issuer: S = Q K^T (M128 N64, 16 k-steps, double-buffered in TMEM), O += P V (V consumed as an
//              MN-major B operand: 4 x N64 per 16-key k-step), O [128 x 256] fp32 resident in TMEM for the whole unit.
//   warps 4-7  softmax warpgroup, thread = query row: fp32 running max / sum, P -> bf16 -> 128B-swizzled smem tile,
//              lazy rescale of the TMEM accumulator (only when the running max grows by more than 2^8).
// Every unit writes an un-normalised fp32 partial (O, m, l); attn256_combine_kernel merges the KV splits (one split for
// large batches, up to 8 for a single object so that all 148 SMs have work) and emits bf16 for the output projection.
#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

namespace fa256 {
static constexpr int D = 256;
static constexpr int TQ = 128;
static constexpr int TK = 64;
static constexpr int NSTAGE = 2;
static constexpr int Q_BYTES = TQ * D * 2;          // 64 KB: 4 sub-tiles of 16 KB
static constexpr int KV_BYTES = TK * D * 2;         // 32 KB: 4 sub-tiles of 8 KB
static constexpr int OFF_Q = 0;
static constexpr int OFF_K = OFF_Q + Q_BYTES;
static constexpr int OFF_V = OFF_K + NSTAGE * KV_BYTES;
static constexpr int OFF_BAR = OFF_V + NSTAGE * KV_BYTES;
static constexpr int SMEM = OFF_BAR + 256 + 1024;
static constexpr int THREADS = 256;                 // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-7 softmax
static constexpr int COL_S = 0;                     // S buffer b at COL_S + 64 b
static constexpr int COL_O = 128;                   // O: 256 columns
static constexpr int COL_Q = 384;                   // Q as the A operand of Q K^T: 128 rows x 256 bf16 = 128 columns
static constexpr int TMEM_COLS = 512;
static constexpr float RESCALE_THRESHOLD = 8.0f;    // log2 units
}  // namespace fa256

struct Fa256Params {
  float* o_part;      // [splits][n][Tq][256] un-normalised partial outputs
  float* ml_part;     // [splits][n][Tq][2]   (running max in log2 units, running sum)
  int n, Tq, Lk;
  int q_tiles;        // Tq / 128
  int kv_tiles;       // ceil(Lk / 64)
  int splits;
  int tiles_per_split;
  int num_units;      // n * q_tiles * splits
  float scale_log2;   // (1 / sqrt(256)) * log2(e)
};

struct Unit256 { int b, qt, sp, kv0, kv1; };
SAM2B_DEVINL Unit256 unit256(const Fa256Params& p, int u) {
  Unit256 c;
  c.sp = u % p.splits;
  const int r = u / p.splits;
  c.qt = r % p.q_tiles;
  c.b = r / p.q_tiles;
  c.kv0 = c.sp * p.tiles_per_split;
  c.kv1 = min(p.kv_tiles, c.kv0 + p.tiles_per_split);
  return c;
}

__global__ void __launch_bounds__(fa256::THREADS, 1)
flash_attn_hd256_tcgen05_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                                const __grid_constant__ CUtensorMap map_v, const Fa256Params p) {
  using namespace fa256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* q_full = bars + 0;
  uint64_t* q_empty = bars + 1;
  uint64_t* k_full = bars + 2;              // [NSTAGE]
  uint64_t* k_empty = k_full + NSTAGE;
  uint64_t* v_full = k_empty + NSTAGE;
  uint64_t* v_empty = v_full + NSTAGE;
  uint64_t* s_full = v_empty + NSTAGE;      // [2]
  uint64_t* s_free = s_full + 2;            // [2]
  uint64_t* p_full = s_free + 2;
  uint64_t* o_done = p_full + 1;            // one completion per P V product
  uint64_t* o_free = o_done + 1;            // softmax has read the final O of a unit
  uint64_t* qt_full = o_free + 1;           // the unit's Q tile has been copied shared -> tensor memory
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(qt_full + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_q);
    tma_prefetch_desc(&map_k);
    tma_prefetch_desc(&map_v);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1); mbar_init(q_empty, 4); mbar_init(qt_full, 4);
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&s_full[i], 1); mbar_init(&s_free[i], 4); }
    mbar_init(p_full, 4); mbar_init(o_done, 1); mbar_init(o_free, 4);
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
      int it = 0, nu = 0;
      for (int u = blockIdx.x; u < p.num_units; u += gridDim.x, ++nu) {
        const Unit256 c = unit256(p, u);
        mbar_wait(q_empty, (nu & 1) ^ 1);
        mbar_arrive_expect_tx(q_full, Q_BYTES);
#pragma unroll
        for (int s = 0; s < 4; ++s) tma_load_3d(smem + OFF_Q + s * 16384, &map_q, q_full, 64 * s, c.qt * TQ, c.b);
        for (int j = c.kv0; j < c.kv1; ++j, ++it) {
          const int st = it % NSTAGE;
          const uint32_t ph = (it / NSTAGE) & 1;
          uint8_t* k = smem + OFF_K + st * KV_BYTES;
          uint8_t* v = smem + OFF_V + st * KV_BYTES;
          mbar_wait(&k_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&k_full[st], KV_BYTES);
#pragma unroll
          for (int s = 0; s < 4; ++s) tma_load_3d(k + s * 8192, &map_k, &k_full[st], 64 * s, j * TK, c.b);
          mbar_wait(&v_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&v_full[st], KV_BYTES);
#pragma unroll
          for (int s = 0; s < 4; ++s) tma_load_3d(v + s * 8192, &map_v, &v_full[st], 64 * s, j * TK, c.b);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp, warp-uniform values; one elected lane issues: common.cuh umma_*_w) =====================
    {
      constexpr uint32_t idesc_qk = make_idesc_ex(128, 64, 0);
      constexpr uint32_t idesc_pv = make_idesc_ex(128, 256, 1);
      const uint32_t sbase = smem_u32(smem);
      uint32_t n_s = 0;       // Q K^T products issued (S buffer = n_s & 1)
      uint32_t n_o = 0;       // P V products issued
      int it = 0, nu = 0;
      auto issue_qk = [&](int st) {
        const uint32_t bsel = n_s & 1;
        mbar_wait(&s_free[bsel], ((n_s >> 1) & 1) ^ 1);
        tc_fence_after();
        // A = Q from tensor memory: with N = 64 a shared-memory A operand (4 KB re-read per instruction) would need 192 B/clk
        const uint32_t k = sbase + OFF_K + st * KV_BYTES;
        const uint32_t d = tmem_base + COL_S + 64 * bsel;
#pragma unroll
        for (int ks = 0; ks < 16; ++ks)
          umma_bf16_ts_w(d, tmem_base + COL_Q + 8 * ks, make_smem_desc(k + (ks >> 2) * 8192 + (ks & 3) * 32, 16, 1024, 2), idesc_qk,
                       ks != 0);
        umma_commit_w(&s_full[bsel]);
        ++n_s;
      };
      for (int u = blockIdx.x; u < p.num_units; u += gridDim.x, ++nu) {
        const Unit256 c = unit256(p, u);
        const int n_kv = c.kv1 - c.kv0;
        mbar_wait(qt_full, nu & 1);
        mbar_wait(&k_full[it % NSTAGE], (it / NSTAGE) & 1);
        tc_fence_after();
        issue_qk(it % NSTAGE);
        umma_commit_w(&k_empty[it % NSTAGE]);
        mbar_wait(o_free, (nu & 1) ^ 1);          // previous unit's accumulator has been read out
        for (int j = 0; j < n_kv; ++j, ++it) {
          const int st = it % NSTAGE, st1 = (it + 1) % NSTAGE;
          if (j + 1 < n_kv) {
            mbar_wait(&k_full[st1], ((it + 1) / NSTAGE) & 1);
            tc_fence_after();
            issue_qk(st1);                         // S(j+1) is computed while the softmax warps work on S(j)
            umma_commit_w(&k_empty[st1]);
          }
          mbar_wait(&v_full[st], (it / NSTAGE) & 1);
          mbar_wait(p_full, n_o & 1);
          tc_fence_after();
          const uint32_t v = sbase + OFF_V + st * KV_BYTES;
          const uint32_t pa = tmem_base + COL_S + 64 * (n_o & 1);      // P(j) lives in the first 32 columns of S(j)'s buffer
          // one N = 256 instruction per 16 keys: V is MN-major, its four 64-dim sub-tiles are 8 KB apart (LBO); P is read from
          // tensor memory, so the only shared-memory operand traffic is V (64 B/clk)
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)
            umma_bf16_ts_w(tmem_base + COL_O, pa + 8 * ks, make_smem_desc(v + ks * 2048, 8192, 1024, 2), idesc_pv, (j | ks) != 0);
          umma_commit_w(o_done);
          umma_commit_w(&v_empty[st]);
          ++n_o;
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax warpgroup (thread = query row) =====================
    const int wq = warp & 3;
    const int r = wq * 32 + lane;
    const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
    const uint32_t to = tmem_base + COL_O + lane_sel;
    uint32_t n = 0;          // KV tiles seen (S buffer = n & 1)
    uint32_t n_pv = 0;       // P V completions observed (always in order, never more than one behind)
    // Q tile: shared (TMA, 128B swizzle) -> registers -> tensor memory, row r -> lane r, two bf16 per column.  Called when no
    // Q K^T product that reads the previous Q is in flight (start of the kernel / after the unit's last P V has completed).
    const uint32_t q_row = smem_u32(smem + OFF_Q) + r * 128;
    const uint32_t q_xor = (uint32_t)(r & 7) << 4;
    auto stage_q = [&](uint32_t parity) {
      mbar_wait(q_full, parity);
#pragma unroll 1
      for (int sub = 0; sub < 4; ++sub) {
        uint32_t qv[32];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 f = ld_shared_f4(q_row + sub * 16384 + (((uint32_t)i << 4) ^ q_xor));
          qv[4 * i] = __float_as_uint(f.x); qv[4 * i + 1] = __float_as_uint(f.y);
          qv[4 * i + 2] = __float_as_uint(f.z); qv[4 * i + 3] = __float_as_uint(f.w);
        }
        tmem_st_x32(tmem_base + COL_Q + 32 * sub + lane_sel, qv);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { mbar_arrive(q_empty); mbar_arrive(qt_full); }
    };
    uint32_t nu = 0;
    if ((int)blockIdx.x < p.num_units) stage_q(0);
    for (int u = blockIdx.x; u < p.num_units; u += gridDim.x, ++nu) {
      const Unit256 c = unit256(p, u);
      const int n_kv = c.kv1 - c.kv0;
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j < n_kv; ++j, ++n) {
        const uint32_t bsel = n & 1;
        mbar_wait(&s_full[bsel], (n >> 1) & 1);
        tc_fence_after();
        uint32_t s[64];
        {
          uint32_t (&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
          uint32_t (&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
          const uint32_t ts = tmem_base + COL_S + 64 * bsel + lane_sel;
          tmem_ld_x32(ts, s0);
          tmem_ld_x32(ts + 32, s1);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[bsel]);      // S is in registers: the MMA warp may overwrite this buffer
        const int valid = p.Lk - (c.kv0 + j) * TK;       // keys of this tile that exist (>= 64 except on the last tile)
        if (valid < TK) {                                // ragged last tile only
#pragma unroll
          for (int i = 0; i < 64; ++i)
            if (i >= valid) s[i] = __float_as_uint(-INFINITY);
        }
        float mx = -INFINITY;
#pragma unroll
        for (int i = 0; i < 64; ++i) mx = fmaxf(mx, __uint_as_float(s[i]));
        mx *= p.scale_log2;
        // lazy rescale: keep the old reference max unless the new one exceeds it by more than 2^8
        float corr = 1.0f;
        const bool grow = mx > m_run + RESCALE_THRESHOLD;
        if (grow) { corr = exp2f(m_run - mx); m_run = mx; }     // first tile: m_run = -inf -> corr = 0 (nothing accumulated yet)
        // packed fp32 pairs: one FFMA2 (scale, minus max) and one FADD2 (row sum) per two keys
        const uint64_t sc2 = f2_pack(p.scale_log2, p.scale_log2), nm2 = f2_pack(-m_run, -m_run);
        uint64_t rs2a = f2_pack(0.f, 0.f), rs2b = rs2a;
        uint32_t pk[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float a0, a1, e0, e1;
          f2_unpack(f2_fma(f2_pack(__uint_as_float(s[2 * i]), __uint_as_float(s[2 * i + 1])), sc2, nm2), a0, a1);
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(a0));
          asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(a1));
          if (i & 1) rs2b = f2_add(rs2b, f2_pack(e0, e1)); else rs2a = f2_add(rs2a, f2_pack(e0, e1));
          pk[i] = pack_bf16x2(e0, e1);
        }
        float rs0, rs1;
        f2_unpack(f2_add(rs2a, rs2b), rs0, rs1);
        l_run = fmaf(l_run, corr, rs0 + rs1);
        if (j > 0) {
          // P V of the previous tile must be complete before P is overwritten / the accumulator is rescaled
          mbar_wait(o_done, n_pv & 1);
          ++n_pv;
          tc_fence_after();
          if (__any_sync(0xffffffffu, grow)) {
#pragma unroll 1
            for (int ch = 0; ch < 8; ++ch) {
              uint32_t o[32];
              tmem_ld_x32(to + 32 * ch, o);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
              tmem_st_x32(to + 32 * ch, o);
            }
            tmem_st_wait();
          }
        }
        tmem_st_x32(tmem_base + COL_S + 64 * bsel + lane_sel, pk);     // P(j) as the A operand of P V, over S(j)'s columns
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
      }
      // ---- unit done: un-normalised O, running max and sum -> partial buffers ----
      mbar_wait(o_done, n_pv & 1);
      ++n_pv;
      tc_fence_after();
      if (u + (int)gridDim.x < p.num_units) stage_q((nu + 1) & 1);   // next unit's first Q K^T overlaps this unit's read-out
      const long row_g = ((long)c.sp * p.n + c.b) * p.Tq + (long)c.qt * TQ + r;
      float* op = p.o_part + row_g * D;
#pragma unroll 1
      for (int ch = 0; ch < 8; ++ch) {
        uint32_t o[32];
        tmem_ld_x32(to + 32 * ch, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          st_global_f4(op + 32 * ch + 4 * i, make_float4(__uint_as_float(o[4 * i]), __uint_as_float(o[4 * i + 1]),
                                                          __uint_as_float(o[4 * i + 2]), __uint_as_float(o[4 * i + 3])));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);
      p.ml_part[row_g * 2] = m_run;
      p.ml_part[row_g * 2 + 1] = l_run;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, fa256::TMEM_COLS);
  }
}

// out[b][t][:] = sum_s O_s 2^(m_s - m) / sum_s l_s 2^(m_s - m)  (bf16), one warp per query row
__global__ void attn256_combine_kernel(const float* __restrict__ o_part, const float* __restrict__ ml_part, bf16* __restrict__ out,
                                       long rows, int splits) {
  const long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float m = -INFINITY;
  for (int s = 0; s < splits; ++s) m = fmaxf(m, ml_part[((long)s * rows + row) * 2]);
  float l = 0.f;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int s = 0; s < splits; ++s) {
    const float w = exp2f(ml_part[((long)s * rows + row) * 2] - m);
    l = fmaf(ml_part[((long)s * rows + row) * 2 + 1], w, l);
    const float4* src = reinterpret_cast<const float4*>(o_part + ((long)s * rows + row) * 256 + lane * 8);
    const float4 a = src[0], b = src[1];
    acc[0] = fmaf(a.x, w, acc[0]); acc[1] = fmaf(a.y, w, acc[1]); acc[2] = fmaf(a.z, w, acc[2]); acc[3] = fmaf(a.w, w, acc[3]);
    acc[4] = fmaf(b.x, w, acc[4]); acc[5] = fmaf(b.y, w, acc[5]); acc[6] = fmaf(b.z, w, acc[6]); acc[7] = fmaf(b.w, w, acc[7]);
  }
  const float inv = 1.0f / l;
  uint4 v;
  v.x = pack_bf16x2(acc[0] * inv, acc[1] * inv); v.y = pack_bf16x2(acc[2] * inv, acc[3] * inv);
  v.z = pack_bf16x2(acc[4] * inv, acc[5] * inv); v.w = pack_bf16x2(acc[6] * inv, acc[7] * inv);
  *reinterpret_cast<uint4*>(out + row * 256 + lane * 8) = v;
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn256)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                     CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn256 fa256_encode_fn() {
  static EncodeTiledFn256 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn256>(f);
  });
  return fn;
}

// [batch][rows][256] bf16, box {64, box_rows, 1}, 128B swizzle; rows past `rows` read as zero
static int fa256_map(CUtensorMap* m, const bf16* ptr, long rows, long batch, long batch_stride, int box_rows) {
  EncodeTiledFn256 enc = fa256_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (batch_stride * 2) % 16) { set_error("flash_attention_hd256: operand not 16B aligned"); return 1; }
  const cuuint64_t dims[3] = {256, (cuuint64_t)rows, (cuuint64_t)batch};
  const cuuint64_t strides[2] = {512, (cuuint64_t)batch_stride * 2};
  const cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<bf16*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("flash_attention_hd256: cuTensorMapEncodeTiled failed (%d)", (int)r); return 1; }
  return 0;
}

int flash_attention_hd256_splits(int n, int Tq, long Lk) {
  // units = q tiles x splits are dealt round-robin to one CTA per SM: pick the split count that minimises the longest CTA's
  // work, waves x (tiles per unit + per-unit prologue / read-out) plus the combine pass that grows with the split count
  const long q_units = (long)n * (Tq / fa256::TQ);
  const long kv_tiles = (Lk + fa256::TK - 1) / fa256::TK;
  const long sms = num_sms();
  if (const char* ev = getenv("SAM2B_FA256_SPLITS")) {          // experiments only
    long s = atol(ev);
    if (s >= 1 && s <= kv_tiles) {
      const long per = (kv_tiles + s - 1) / s;
      return (int)((kv_tiles + per - 1) / per);
    }
  }
  long best = 1;
  double best_cost = 1e30;
  for (long s = 1; s <= 16 && s <= kv_tiles; ++s) {
    const long per = (kv_tiles + s - 1) / s;
    const long eff = (kv_tiles + per - 1) / per;                // every split owns at least one tile
    if (eff != s) continue;
    const long waves = (q_units * s + sms - 1) / sms;
    const double cost = (double)waves * (per + 4.0) + 0.03 * s * (double)q_units;   // partials: 256 KB written + read per unit
    if (cost < best_cost - 1e-9) { best_cost = cost; best = s; }
  }
  return (int)best;
}

int flash_attention_hd256(const bf16* q, long q_batch_stride, const bf16* k, long k_batch_stride, const bf16* v,
                          long v_batch_stride, int n, int Tq, long Lk, float* o_part, float* ml_part, bf16* out,
                          cudaStream_t stream) {
  using namespace fa256;
  if (n <= 0 || Tq <= 0 || Tq % TQ || Lk <= 0) { set_error("flash_attention_hd256: bad shape n=%d Tq=%d Lk=%ld", n, Tq, Lk); return 1; }
  CUtensorMap mq, mk, mv;
  if (fa256_map(&mq, q, Tq, n, q_batch_stride, TQ) || fa256_map(&mk, k, Lk, n, k_batch_stride, TK) ||
      fa256_map(&mv, v, Lk, n, v_batch_stride, TK)) return 1;
  static PerDeviceFlag attr;
  if (!attr.is_set()) {
    cudaError_t e = cudaFuncSetAttribute(flash_attn_hd256_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(flash_attn_hd256): %s", cudaGetErrorString(e)); return 1; }
    attr.set();
  }
  Fa256Params p;
  p.o_part = o_part; p.ml_part = ml_part; p.n = n; p.Tq = Tq; p.Lk = (int)Lk;
  p.q_tiles = Tq / TQ;
  p.kv_tiles = (int)((Lk + TK - 1) / TK);
  p.splits = flash_attention_hd256_splits(n, Tq, Lk);
  p.tiles_per_split = (p.kv_tiles + p.splits - 1) / p.splits;
  p.num_units = n * p.q_tiles * p.splits;
  p.scale_log2 = 1.4426950408889634f / 16.0f;
  const int grid = p.num_units < num_sms() ? p.num_units : num_sms();
  {
    char nm[96];
    if (g_prof_on) snprintf(nm, sizeof(nm), "attn256 n%d q%d k%ld splits%d", n, Tq, Lk, p.splits);
    ProfRange pr(stream, nm, 4.0 * n * (double)Tq * (double)Lk * D, ((double)n * (Tq + 2.0 * Lk) * D) * 2.0);
    flash_attn_hd256_tcgen05_kernel<<<grid, THREADS, SMEM, stream>>>(mq, mk, mv, p);
    SAM2B_CHECK_LAUNCH("flash_attn_hd256_tcgen05");
  }
  {
    const long rows = (long)n * Tq;
    ProfRange pr(stream, "attn256_combine", 0.0, (double)rows * D * (4.0 * p.splits + 2.0));
    attn256_combine_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, stream>>>(o_part, ml_part, out, rows, p.splits);
    SAM2B_CHECK_LAUNCH("attn256_combine");
  }
  return 0;
}

}  // namespace sam2b
