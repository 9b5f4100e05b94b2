// K3: flash attention for Hiera's head_dim-72 MultiScaleAttention on the 5th-gen tensor cores (tcgen05 / TMEM / TMA).
// Covers the blocks that carry almost all of the trunk's attention FLOPs: the 16x16-window blocks of stage 3
// (256 queries x 256 keys per window) and the three global blocks 23/33/43 (4096 x 4096 per head).
// Reference: MultiScaleAttention.forward (hieradet.py:56-81) + window_partition/unpartition
// (backbones/utils.py:16-60), i.e. softmax(q k^T / sqrt(72)) v per (window, head).
//
// One persistent CTA works on units of 256 queries (one 16x16 window, or 4 rows of the 64x64 grid in global mode) of
// one (image, head), as two 128-row Q tiles that ping-pong on the tensor core:
//   warp 0      TMA producer: Q / K / V tiles are gathered straight out of the packed qkv tensor [B,H,W,3,heads,72]
//               by 5-D tensor maps (d, section*heads+head, x, y, b); window partition = box coordinates.  head_dim 72
//               is split 64 + 16: a 128 B-swizzled tile of d 0..63 and a 32 B-swizzled tile of d 64..79 whose last
//               8 columns are out of bounds of the d axis and therefore zero-filled by TMA.
//   warp 1      MMA issuer (one thread): S_t = Q_t K^T (M128 N128, 5 k-steps) into TMEM, then O_t = P_t V
//               (V consumed as an MN-major B operand, N = 64 + 16) into TMEM.
//   warps 4-7   softmax warpgroup of Q tile 0, warps 8-11 of Q tile 1: thread = query row (tcgen05.ld 32x32b),
//               fp32 online softmax in ONE pass over S (128 scores per thread held in registers), P written as bf16
//               into a 128B-swizzled K-major smem tile (A operand of P V); O_t accumulates in TMEM across the KV tiles
//               and is rescaled there only when the running max moves by more than 2^8 (lazy rescale).
#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

namespace fa {
static constexpr int HD = 72;
static constexpr int TQ = 128;          // rows per Q tile (UMMA M)
static constexpr int TK = 128;          // keys per KV tile
static constexpr int NSTAGE = 2;        // K and V ring depth
static constexpr int LO_BYTES = 128 * 128;   // [128 rows][64 d] bf16, SW128
static constexpr int HI_BYTES = 128 * 32;    // [128 rows][16 d] bf16, SW32
static constexpr int TILE_BYTES = LO_BYTES + HI_BYTES;
// V tiles are the MN-major B operand of P V.  A tcgen05.mma costs max(N / 2, ~50) clocks (tools/umma_rate.cu), so the 64 + 16
// split used for Q and K would pay 100 clocks per 16 keys; instead V is two 128B-swizzle atoms (d 0-63, d 64-127 with 72+ zero
// filled by TMA) 16 KB apart, which one N = 80 instruction walks through the descriptor's leading-dimension offset.
static constexpr int V_BYTES = 2 * LO_BYTES;
static constexpr int P_BYTES = 128 * 128 * 2;     // [128][128] bf16 as two SW128 sub-tiles
static constexpr int OFF_Q = 0;                                   // 2 Q tiles
static constexpr int OFF_K = OFF_Q + 2 * TILE_BYTES;
static constexpr int OFF_V = OFF_K + NSTAGE * TILE_BYTES;
static constexpr int OFF_P = OFF_V + NSTAGE * V_BYTES;            // 2 P tiles
static constexpr int OFF_BAR = OFF_P + 2 * P_BYTES;
static constexpr int SMEM = OFF_BAR + 256 + 1024;
static constexpr int THREADS = 384;
// TMEM columns
static constexpr int COL_S = 0;         // S_t at COL_S + 128 t
static constexpr int COL_O = 256;       // O_t at COL_O + 80 t
static constexpr int TMEM_COLS = 512;
static constexpr float RESCALE_THRESHOLD = 8.0f;   // log2 units
}  // namespace fa

struct FaParams {
  bf16* out;
  int B, H, W, heads, ws;
  int n_kv;            // KV tiles per unit (2 for 16x16 windows, H*W/128 in global mode)
  int blocks_per_img;  // 256-query blocks per image
  int nwx;             // windows per grid row (window mode)
  int rows_per_tile;   // grid rows covered by one 128-token tile = 128 / ws
  int num_units;
  float scale_log2;
  int pingpong;        // the two soft-max warpgroups take turns in their exponential phase (see the exp loop)
  long long* trace;    // development only (SAM2B_ATTN_TRACE): [role][256] pairs (event code, SM clock) written by CTA 0
};
#define FA_TR(role, code)                                                                                   \
  do {                                                                                                      \
    if (p.trace && blockIdx.x == 0 && lane == 0 && tr_n < 256) {                                            \
      p.trace[((role) * 256 + tr_n) * 2] = (code);                                                          \
      p.trace[((role) * 256 + tr_n) * 2 + 1] = clock64();                                                   \
      ++tr_n;                                                                                               \
    }                                                                                                       \
  } while (0)

struct UnitCoord { int b, head, x0, y0; };
SAM2B_DEVINL UnitCoord unit_coord(const FaParams& p, int u) {
  UnitCoord c;
  const int blk = u % p.blocks_per_img;
  const int bh = u / p.blocks_per_img;
  c.head = bh % p.heads;
  c.b = bh / p.heads;
  if (p.ws == p.W) { c.x0 = 0; c.y0 = blk * 2 * p.rows_per_tile; }          // global: 256 consecutive tokens
  else { c.x0 = (blk % p.nwx) * p.ws; c.y0 = (blk / p.nwx) * p.ws; }        // one window
  return c;
}

__global__ void __launch_bounds__(fa::THREADS, 1)
flash_attn_hd72_tcgen05_kernel(const __grid_constant__ CUtensorMap map_lo, const __grid_constant__ CUtensorMap map_hi,
                               const __grid_constant__ CUtensorMap map_out, const FaParams p) {
  using namespace fa;
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
  uint64_t* s_free = s_full + 2;
  uint64_t* p_full = s_free + 2;
  uint64_t* o_full = p_full + 2;
  uint64_t* o_free = o_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  int tr_n = 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_lo);
    tma_prefetch_desc(&map_hi);
    tma_prefetch_desc(&map_out);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1); mbar_init(q_empty, 1);
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1); mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
    for (int t = 0; t < 2; ++t) {
      mbar_init(&s_full[t], 1); mbar_init(&s_free[t], 4); mbar_init(&p_full[t], 4);
      mbar_init(&o_full[t], 1); mbar_init(&o_free[t], 4);
    }
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
        const UnitCoord c = unit_coord(p, u);
        FA_TR(0, 100 + nu);
        mbar_wait(q_empty, (nu & 1) ^ 1);
        FA_TR(0, 200 + nu);
        mbar_arrive_expect_tx(q_full, 2 * TILE_BYTES);
#pragma unroll
        for (int t = 0; t < 2; ++t) {
          uint8_t* q = smem + OFF_Q + t * TILE_BYTES;
          tma_load_5d(q, &map_lo, q_full, 0, c.head, c.x0, c.y0 + t * p.rows_per_tile, c.b);
          tma_load_5d(q + LO_BYTES, &map_hi, q_full, 64, c.head, c.x0, c.y0 + t * p.rows_per_tile, c.b);
        }
        const bool global = p.ws == p.W;
        for (int j = 0; j < p.n_kv; ++j, ++it) {
          const int st = it % NSTAGE;
          const uint32_t ph = (it / NSTAGE) & 1;
          const int ky = (global ? 0 : c.y0) + j * p.rows_per_tile;
          uint8_t* k = smem + OFF_K + st * TILE_BYTES;
          uint8_t* v = smem + OFF_V + st * V_BYTES;
          mbar_wait(&k_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&k_full[st], TILE_BYTES);
          tma_load_5d(k, &map_lo, &k_full[st], 0, p.heads + c.head, c.x0, ky, c.b);
          tma_load_5d(k + LO_BYTES, &map_hi, &k_full[st], 64, p.heads + c.head, c.x0, ky, c.b);
          mbar_wait(&v_empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&v_full[st], V_BYTES);
          tma_load_5d(v, &map_lo, &v_full[st], 0, 2 * p.heads + c.head, c.x0, ky, c.b);
          tma_load_5d(v + LO_BYTES, &map_lo, &v_full[st], 64, 2 * p.heads + c.head, c.x0, ky, c.b);
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (one thread) =====================
    if (lane == 0) {
      constexpr uint32_t idesc_qk = make_idesc_ex(128, 128, 0);
      constexpr uint32_t idesc_pv = make_idesc_ex(128, 80, 1);
      const uint32_t sbase = smem_u32(smem);
      uint32_t n_s[2] = {0, 0};     // QK^T products issued on S_t
      uint32_t n_o[2] = {0, 0};     // P V products issued on O_t
      int it = 0, nu = 0;
      auto issue_qk = [&](int t, int st) {
        mbar_wait(&s_free[t], (n_s[t] & 1) ^ 1);
        tc_fence_after();
        const uint32_t q = sbase + OFF_Q + t * TILE_BYTES, k = sbase + OFF_K + st * TILE_BYTES;
        const uint32_t d = tmem_base + COL_S + 128 * t;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          umma_bf16(d, make_smem_desc(q + 32 * ks, 16, 1024, 2), make_smem_desc(k + 32 * ks, 16, 1024, 2), idesc_qk, ks != 0);
        umma_bf16(d, make_smem_desc(q + LO_BYTES, 16, 256, 6), make_smem_desc(k + LO_BYTES, 16, 256, 6), idesc_qk, 1);
        umma_commit(&s_full[t]);
        ++n_s[t];
      };
      auto issue_pv = [&](int t, int st, int j) {
        mbar_wait(&p_full[t], n_o[t] & 1);                  // P_t(j) written; O_t rescaled if the running max moved
        if (j == 0) mbar_wait(&o_free[t], (nu & 1) ^ 1);    // the previous unit's O_t has been read out of TMEM
        tc_fence_after();
        const uint32_t pa = sbase + OFF_P + t * P_BYTES, v = sbase + OFF_V + st * V_BYTES;
        const uint32_t d = tmem_base + COL_O + 80 * t;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          // A: P[128][16 keys of k-step ks] (K-major, sub-tile ks/4); B: V rows 16ks..16ks+15 (MN-major: d contiguous)
          const uint64_t da = make_smem_desc(pa + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024, 2);
          umma_bf16(d, da, make_smem_desc(v + ks * 2048, LO_BYTES, 1024, 2), idesc_pv, (j | ks) != 0);
        }
        umma_commit(&o_full[t]);
        ++n_o[t];
      };
      for (int u = blockIdx.x; u < p.num_units; u += gridDim.x, ++nu) {
        FA_TR(1, 100 + nu);
        mbar_wait(q_full, nu & 1);
        tc_fence_after();
        FA_TR(1, 200 + nu);
        {
          const int st = it % NSTAGE;
          mbar_wait(&k_full[st], (it / NSTAGE) & 1);
          tc_fence_after();
          FA_TR(1, 300 + nu);
          issue_qk(0, st);
          issue_qk(1, st);
          FA_TR(1, 400 + nu);
          umma_commit(&k_empty[st]);
          if (p.n_kv == 1) umma_commit(q_empty);
        }
        for (int j = 0; j < p.n_kv; ++j, ++it) {
          const int st = it % NSTAGE;
          const int st1 = (it + 1) % NSTAGE;
          mbar_wait(&v_full[st], (it / NSTAGE) & 1);
          tc_fence_after();
          const bool more = j + 1 < p.n_kv;
          if (more) {
            // Q K^T of the NEXT tile first: S_t is free as soon as the softmax warps hold tile j's scores in registers (early
            // in their pass), so S_t(j+1) is being computed while they exponentiate tile j; P V of tile j follows.
            mbar_wait(&k_full[st1], ((it + 1) / NSTAGE) & 1);
            tc_fence_after();
            issue_qk(0, st1);
            issue_qk(1, st1);
            umma_commit(&k_empty[st1]);
            if (j + 2 == p.n_kv) umma_commit(q_empty);
          }
          FA_TR(1, 500 + 10 * nu + j);
          issue_pv(0, st, j);
          FA_TR(1, 600 + 10 * nu + j);
          issue_pv(1, st, j);
          FA_TR(1, 700 + 10 * nu + j);
          umma_commit(&v_empty[st]);
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax warpgroups (thread = query row) =====================
    const int t = (warp - 4) >> 2;            // Q tile of this warpgroup
    const int wq = warp & 3;                  // TMEM lane quarter
    const int r = wq * 32 + lane;             // row inside the Q tile
    const uint32_t lane_sel = (uint32_t)(wq * 32) << 16;
    const uint32_t ts = tmem_base + COL_S + 128 * t + lane_sel;
    const uint32_t to = tmem_base + COL_O + 80 * t + lane_sel;
    const uint32_t p_row = smem_u32(smem + OFF_P + t * P_BYTES) + r * 128;
    const uint32_t p_xor = (uint32_t)(r & 7) << 4;
    const int C = p.heads * HD;
    uint32_t n = 0;                           // KV tiles processed by this warpgroup (all units): phase of s_full[t]
    uint32_t n_pv = 0;                        // P V completions observed on o_full[t] (in order)
    bool stores_pending = false;
    if (p.pingpong && t == 1) named_bar_arrive(3, 256);       // warpgroup 0 takes the first turn
    for (int u = blockIdx.x; u < p.num_units; u += gridDim.x) {
      const UnitCoord c = unit_coord(p, u);
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = 0; j < p.n_kv; ++j, ++n) {
        if (wq == 0) FA_TR(2 + t, 100 + j);
        mbar_wait(&s_full[t], n & 1);
        tc_fence_after();
        if (wq == 0) FA_TR(2 + t, 200 + j);
        // the 128 scores of this row: one pass through TMEM, held in registers for max and exp
        uint32_t s[128];
        {
          uint32_t (&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
          uint32_t (&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
          uint32_t (&s2)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[64]);
          uint32_t (&s3)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[96]);
          tmem_ld_x32(ts, s0);
          tmem_ld_x32(ts + 32, s1);
          tmem_ld_x32(ts + 64, s2);
          tmem_ld_x32(ts + 96, s3);
          tmem_ld_wait();
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&s_free[t]);       // S_t is in registers: the next Q K^T may overwrite it
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          mx0 = fmaxf(mx0, __uint_as_float(s[i]));
          mx1 = fmaxf(mx1, __uint_as_float(s[32 + i]));
          mx2 = fmaxf(mx2, __uint_as_float(s[64 + i]));
          mx3 = fmaxf(mx3, __uint_as_float(s[96 + i]));
        }
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
        // lazy rescale: the reference max only moves when the new one exceeds it by more than 2^8 (exact: the factor
        // cancels in O / l; P stays below 2^8 and the fp32 accumulators have the headroom)
        float corr = 1.0f;
        const bool grow = mx > m_run + RESCALE_THRESHOLD;
        if (grow) { corr = ex2_approx(m_run - mx); m_run = mx; }      // first tile: m_run = -inf -> corr = 0
        if (j > 0) {
          // P V of the previous tile has completed (it was issued before this tile's Q K^T): P_t may be overwritten and
          // the accumulator rescaled
          if (wq == 0) FA_TR(2 + t, 300 + j);
          mbar_wait(&o_full[t], n_pv & 1);
          ++n_pv;
          tc_fence_after();
          if (wq == 0) FA_TR(2 + t, 400 + j);
          if (__any_sync(0xffffffffu, grow)) {
            uint32_t a[32];
#pragma unroll 1
            for (int ch = 0; ch < 2; ++ch) {
              tmem_ld_x32(to + 32 * ch, a);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) a[i] = __float_as_uint(__uint_as_float(a[i]) * corr);
              tmem_st_x32(to + 32 * ch, a);
            }
            tmem_ld_x16(to + 64, a);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = __float_as_uint(__uint_as_float(a[i]) * corr);
            tmem_st_x16(to + 64, a);
            tmem_st_wait();
          }
        }
        if (stores_pending) {
          // the previous unit's output rows were staged in this P tile: every warp's TMA store must have read them before any
          // warp writes P again (P rows are not warp-private).  The stores were issued a whole score tile ago, so this is free.
          if (lane == 0) bulk_wait_read<0>();
          __syncwarp();
          named_bar_sync(1 + t, 128);
          stores_pending = false;
        }
        // Turn-taking on the special-function unit: the exponentials of a tile keep MUFU busy for 1 024 clocks when one warpgroup
        // has it alone and for 2 048 when both are in this loop at once - and both arrive here together (their score tiles are
        // issued back to back), after which both do their MUFU-free work (S wait, tcgen05.ld, row maximum, rescale) together
        // too.  Strict alternation (warpgroup 0 first; same number of tiles on both sides, so it cannot deadlock) puts one
        // warpgroup's MUFU-free phase under the other one's exponentials.
        if (p.pingpong) named_bar_sync(3 + t, 256);
        // P = exp2(S * scale - m) -> bf16 -> swizzled K-major smem tile; fp32 row sum
        // packed fp32 pairs: one FFMA2 (scale, minus max) and one FADD2 (row sum) per two keys
        const uint64_t sc2 = f2_pack(p.scale_log2, p.scale_log2), nm2 = f2_pack(-m_run, -m_run);
        uint64_t rs2a = f2_pack(0.f, 0.f), rs2b = rs2a;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float a0, a1;
            f2_unpack(f2_fma(f2_pack(__uint_as_float(s[32 * ch + 2 * i]), __uint_as_float(s[32 * ch + 2 * i + 1])), sc2, nm2), a0, a1);
            const float e0 = ex2_approx(a0), e1 = ex2_approx(a1);
            if (i & 1) rs2b = f2_add(rs2b, f2_pack(e0, e1)); else rs2a = f2_add(rs2a, f2_pack(e0, e1));
            pk[i] = pack_bf16x2(e0, e1);
          }
          // 32 keys = 64 B = 16 B chunks (ch & 1) * 4 + i of sub-tile ch >> 1, XOR-swizzled by (row & 7)
          const uint32_t base = p_row + (ch >> 1) * 16384;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            st_shared_u4(base + (((uint32_t)((ch & 1) * 4 + i) << 4) ^ p_xor), pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
        }
        if (p.pingpong) named_bar_arrive(3 + (t ^ 1), 256);
        float rs0, rs1;
        f2_unpack(f2_add(rs2a, rs2b), rs0, rs1);
        l_run = fmaf(l_run, corr, rs0 + rs1);
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[t]);
        if (wq == 0) FA_TR(2 + t, 500 + j);
      }
      // ---- unit done: O_t / l -> bf16 -> out[token][head*72 ..] ----
      mbar_wait(&o_full[t], n_pv & 1);
      ++n_pv;
      tc_fence_after();
      if (wq == 0) FA_TR(2 + t, 600);
      uint32_t oa[32], ob[32], oc[32];
      tmem_ld_x32(to, oa);
      tmem_ld_x32(to + 32, ob);
      tmem_ld_x16(to + 64, oc);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&o_free[t]);
      // A thread owns one query row (144 B of output): written straight to global memory that is 32 separate 16 B pieces per
      // store instruction, nine times per warp - the load / store unit was busy with them for ~5 000 clocks per unit (traced).
      // The rows therefore go to shared memory (the P tile is free: its last P V has completed) and leave as ONE TMA store per
      // warp, 32 consecutive tokens of the tile each.
      const float inv = 1.0f / l_run;
      const uint32_t stg = smem_u32(smem + OFF_P + t * P_BYTES) + (uint32_t)r * 144u;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        st_shared_u4(stg + 16 * i, pack_bf16x2(__uint_as_float(oa[8 * i]) * inv, __uint_as_float(oa[8 * i + 1]) * inv),
                     pack_bf16x2(__uint_as_float(oa[8 * i + 2]) * inv, __uint_as_float(oa[8 * i + 3]) * inv),
                     pack_bf16x2(__uint_as_float(oa[8 * i + 4]) * inv, __uint_as_float(oa[8 * i + 5]) * inv),
                     pack_bf16x2(__uint_as_float(oa[8 * i + 6]) * inv, __uint_as_float(oa[8 * i + 7]) * inv));
#pragma unroll
      for (int i = 0; i < 4; ++i)
        st_shared_u4(stg + 64 + 16 * i, pack_bf16x2(__uint_as_float(ob[8 * i]) * inv, __uint_as_float(ob[8 * i + 1]) * inv),
                     pack_bf16x2(__uint_as_float(ob[8 * i + 2]) * inv, __uint_as_float(ob[8 * i + 3]) * inv),
                     pack_bf16x2(__uint_as_float(ob[8 * i + 4]) * inv, __uint_as_float(ob[8 * i + 5]) * inv),
                     pack_bf16x2(__uint_as_float(ob[8 * i + 6]) * inv, __uint_as_float(ob[8 * i + 7]) * inv));
      st_shared_u4(stg + 128, pack_bf16x2(__uint_as_float(oc[0]) * inv, __uint_as_float(oc[1]) * inv),
                   pack_bf16x2(__uint_as_float(oc[2]) * inv, __uint_as_float(oc[3]) * inv),
                   pack_bf16x2(__uint_as_float(oc[4]) * inv, __uint_as_float(oc[5]) * inv),
                   pack_bf16x2(__uint_as_float(oc[6]) * inv, __uint_as_float(oc[7]) * inv));
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        const int tok = t * TQ + wq * 32;                       // first token of this warp inside the 256-query unit
        tma_store_5d(&map_out, smem_u32(smem + OFF_P + t * P_BYTES) + (uint32_t)(wq * 32) * 144u, 0, c.head, c.x0 + tok % p.ws,
                     c.y0 + tok / p.ws, c.b);
        bulk_commit();
      }
      stores_pending = true;      // P_t doubles as the staging area: see the wait before the next unit's first P write
      if (wq == 0) FA_TR(2 + t, 700);
    }
    if (lane == 0) bulk_wait_read<0>();       // staging smem must outlive the last store's read
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, fa::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn5)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn5 fa_encode_fn() {
  static EncodeTiledFn5 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn5>(f);
  });
  return fn;
}

bool flash_attention_hd72_supported(int H, int W, int ws, int q_pool) {
  if (q_pool || H != W || (H * W) % 256) return false;
  if (ws == W) return W <= 128 && 128 % W == 0 && H % (256 / W) == 0;   // global: a tile = 128 / W full grid rows
  return ws == 16 && H % 16 == 0;
}

int flash_attention_hd72(const bf16* qkv, bf16* out, int B, int H, int W, int heads, int ws, cudaStream_t stream) {
  using namespace fa;
  if (!flash_attention_hd72_supported(H, W, ws, 0)) { set_error("flash_attention_hd72: unsupported H=%d W=%d ws=%d", H, W, ws); return 1; }
  EncodeTiledFn5 enc = fa_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  const long C = (long)heads * HD;
  CUtensorMap maps[2];
  const cuuint64_t dims[5] = {(cuuint64_t)HD, (cuuint64_t)(3 * heads), (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  const cuuint64_t strides[4] = {(cuuint64_t)HD * 2, (cuuint64_t)(3 * C * 2), (cuuint64_t)W * 3 * C * 2, (cuuint64_t)H * W * 3 * C * 2};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  for (int i = 0; i < 2; ++i) {
    const cuuint32_t box[5] = {(cuuint32_t)(i == 0 ? 64 : 16), 1, (cuuint32_t)ws, (cuuint32_t)(128 / ws), 1};
    CUresult r = enc(&maps[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<bf16*>(qkv), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, i == 0 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_32B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("flash_attention_hd72: cuTensorMapEncodeTiled failed (%d) map %d", (int)r, i); return 1; }
  }
  // output [B,H,W,heads,72] bf16: boxes of 32 consecutive tokens of a tile (one per softmax warp), 144 B per token
  CUtensorMap map_out;
  {
    const cuuint64_t odims[5] = {(cuuint64_t)HD, (cuuint64_t)heads, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    const cuuint64_t ostrides[4] = {(cuuint64_t)HD * 2, (cuuint64_t)(C * 2), (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    const cuuint32_t obox[5] = {(cuuint32_t)HD, 1, (cuuint32_t)(ws >= 32 ? 32 : ws), (cuuint32_t)(ws >= 32 ? 1 : 32 / ws), 1};
    CUresult r = enc(&map_out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, out, odims, ostrides, obox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("flash_attention_hd72: cuTensorMapEncodeTiled failed (%d) for the output", (int)r); return 1; }
  }
  static PerDeviceFlag attr;
  if (!attr.is_set()) {
    cudaError_t e = cudaFuncSetAttribute(flash_attn_hd72_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(flash_attn): %s", cudaGetErrorString(e)); return 1; }
    attr.set();
  }
  FaParams p;
  p.out = out; p.B = B; p.H = H; p.W = W; p.heads = heads; p.ws = ws;
  p.rows_per_tile = 128 / ws;
  p.n_kv = ws == W ? H * W / 128 : 2;
  p.blocks_per_img = H * W / 256;
  p.nwx = W / ws;
  p.num_units = B * heads * p.blocks_per_img;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  { static int pp = -1; if (pp < 0) { const char* e = getenv("SAM2B_ATTN_PINGPONG"); pp = e ? atoi(e) : 1; } p.pingpong = pp; }
  p.trace = nullptr;
  static long long* trace_dev = nullptr;
  const bool tracing = getenv("SAM2B_ATTN_TRACE") != nullptr;
  if (tracing) {
    if (!trace_dev) cudaMalloc(&trace_dev, 4 * 256 * 2 * sizeof(long long));
    cudaMemsetAsync(trace_dev, 0, 4 * 256 * 2 * sizeof(long long), stream);
    p.trace = trace_dev;
  }
  const int grid = p.num_units < num_sms() ? p.num_units : num_sms();
  char nm[96];
  if (g_prof_on) snprintf(nm, sizeof(nm), "attn h%d heads%d ws%d pool0 b%d tc", H, heads, ws, B);
  const double nq = (double)B * H * W * heads;
  ProfRange pr(stream, nm, 4.0 * nq * (ws == W ? H * W : 256) * HD, (double)B * H * W * 4 * C * 2.0);
  flash_attn_hd72_tcgen05_kernel<<<grid, THREADS, SMEM, stream>>>(maps[0], maps[1], map_out, p);
  SAM2B_CHECK_LAUNCH("flash_attn_hd72_tcgen05");
  if (tracing) {       // development: dump CTA 0's event trace of this launch
    static long long host[4 * 256 * 2];
    cudaStreamSynchronize(stream);
    cudaMemcpy(host, trace_dev, sizeof(host), cudaMemcpyDeviceToHost);
    static const char* roles[4] = {"tma", "mma", "softmax0", "softmax1"};
    long long t0 = 0;
    for (int r = 0; r < 4; ++r) for (int i = 0; i < 256; ++i) if (host[(r * 256 + i) * 2 + 1] && (!t0 || host[(r * 256 + i) * 2 + 1] < t0)) t0 = host[(r * 256 + i) * 2 + 1];
    for (int r = 0; r < 4; ++r) {
      fprintf(stderr, "ATTNTRACE ws=%d %s:", ws, roles[r]);
      for (int i = 0; i < 256 && host[(r * 256 + i) * 2 + 1]; ++i) fprintf(stderr, " %lld@%lld", host[(r * 256 + i) * 2], host[(r * 256 + i) * 2 + 1] - t0);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

}  // namespace sam2b
