// Fused trunk MLP for stage 1 (C = 144): out = x + fc2(GELU(fc1(LN(x)))) in one kernel per 128 rows.
// Reference: MLP.forward (sam2/modeling/sam2_utils.py:131-136) inside MultiScaleBlock.forward (backbones/hieradet.py:160-166),
// with norm2 folded into fc1 as everywhere else in the trunk (weights.py: row-centred W diag(gamma), b + W beta).
//
// Why: in stage 1 (K = 144) the two MLP GEMMs are HBM bound: the 4C-wide hidden tensor is written by fc1 and read back by
// fc2 (2.4 GB per block per 32 images, more than everything else in the block together).  Here the hidden activations never
// leave the SM:
//   for every 64-wide chunk h of the hidden layer (9 chunks)
//     S_b[128 x 64] = A[128 x C] W1[h]^T                  tcgen05.mma, A tile resident in shared memory, W1 streamed by TMA;
//                                                         b = h % 3: three S buffers, chunk h + 2 is computed while
//     P             = bf16(GELU(rstd * S_b + b1[h]))       the GELU warps work on chunk h; P is written back over S_b's
//                                                         first 32 columns as packed pairs
//     O[128 x C]   += P W2[:, h]^T                         tcgen05.mma with the A operand (P) read from TENSOR MEMORY
//   x_out = O + b2 + x; bf16 copy and per-row (sum, sum of squares) for the next block's folded LayerNorm
// Per 128-row tile the kernel must move 128 x 12 C bytes through HBM (x in / out in fp32, the bf16 copies): 9.5k clocks of the
// SM's share of the bus at C = 144, against 5.3k clocks of MMA issue and 5.5k of GELU arithmetic - so every role has its own
// warps and nothing but tensor memory and mbarriers couples them:
//   warp 0      TMA: A tile + W1 ring          warp 3      TMA: W2 ring
//   warp 1      MMA issuer (warp-uniform issue: elect.sync inside the asm keeps the descriptors in uniform registers)
//   warp 2      TMEM allocation
//   warps 4-11  GELU: group g = (warp - 4) / 4 takes the hidden chunks of parity g (lane quarter = warp & 3)
//   warps 12-15 read-out: O -> + b2 + residual -> x (fp32, TMA store), bf16 copy, row statistics.  The residual rows arrive by
//               TMA in [32 rows x 32 columns] boxes requested a whole tile ahead (a thread owns one ROW of tensor memory: plain
//               loads of its row would be 64 B pieces with a DRAM round trip each, and 8 warps of them jam the load / store
//               unit for the GELU warps' bias loads - measured, 37 % of all stall samples in the first version of this kernel)
// TMEM: S 3 x 64 columns + two O buffers of C columns (the read-out of tile t overlaps fc2 of tile t + 1): 480 columns.
// CTA pairs (CL = 2): both CTAs of a cluster walk the same chunk sequence, each loads HALF of every W1 / W2 block and
// multicasts it to both - the weight traffic out of L2 halves.
// C = 288 (stage 2) compiles in this design only with one O buffer and measured no faster than the GEMM pair (its hidden tensor is
// a quarter of stage 1's per block): it is not instantiated, mlp_fused_supported() says so.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

namespace fm {
static constexpr int BM = 128;
static constexpr int BH = 64;                        // hidden chunk = one 64-wide W2 block
static constexpr int NSB = 3;                        // S buffers: fc1 of chunk c + 2 never waits for the GELU pass over chunk c
static constexpr int THREADS = 512;
static constexpr int A_KB_BYTES = BM * 128;          // [128 rows][64 k] bf16, 128B swizzle
static constexpr int W1_BYTES = BH * 128;            // [64 hidden][64 k]
static constexpr int BOX_BYTES = 32 * 128;           // [32 rows][32 fp32 columns], 128B swizzle
static constexpr int COL_S = 0;                      // S buffer b at COL_S + 64 b; P over its first 32 columns
static constexpr int COL_O = NSB * BH;               // O buffer j at COL_O + j C
template <int C>
struct Cfg {
  static constexpr int KB = (C + 63) / 64;           // k-blocks of the A tile (the last one partly zero-filled by TMA)
  static constexpr int NH = 4 * C / BH;              // hidden chunks
  static constexpr int NSPLIT = C / 144;             // fc2 instructions per k-step (N = 144 each; tcgen05.mma N <= 256)
  static constexpr int W2_BYTES = C * 128;           // [C rows][64 hidden]
  static constexpr int NS1 = 3;                      // W1 ring depth in CHUNKS, one per S buffer (KB blocks of 8 KB each: one wait / commit per chunk)
  static constexpr int W1_STAGE = KB * W1_BYTES;
  static constexpr int NS2 = 2;                      // W2 ring depth
  static constexpr int NOB = C == 144 ? 2 : 1;       // O accumulators in tensor memory
  static constexpr int NCH = (C + 31) / 32;          // 32-column read-out chunks of a row (the last one may be half full)
  static constexpr int NBOX = 3;                     // residual boxes per read-out warp (a ring: NCH = 5 chunks per row)
  static constexpr int OFF_A = 0;
  static constexpr int OFF_W1 = OFF_A + KB * A_KB_BYTES;
  static constexpr int OFF_W2 = OFF_W1 + NS1 * W1_STAGE;
  static constexpr int OFF_BOX = OFF_W2 + NS2 * W2_BYTES;
  static constexpr int OFF_BAR = OFF_BOX + 4 * NBOX * BOX_BYTES;
  static constexpr int SMEM = OFF_BAR + 512 + 1024;
  static_assert(4 * C % BH == 0 && C % 144 == 0, "C must be 144 or 288");
  static_assert(COL_O + NOB * C <= 512, "TMEM columns");
  static_assert(OFF_BOX % 1024 == 0, "swizzled boxes need 1024 B alignment");
  static_assert(SMEM <= 227 * 1024, "shared memory");
};
}  // namespace fm

struct FusedMlpParams {
  const float* stats_in;     // [M][parts_in][2] partial (sum, sum of squares) of the input rows
  int parts_in;
  float eps, inv_c;
  const float* b1;           // [4C]  (b + W beta)
  const float* b2;           // [C]
  bf16* xb_out;              // [M][C] bf16 copy of the new stream
  float* stats_out;          // [M][1][2]
  int M, num_tiles;
  long long* trace;          // development only (SAM2B_MLP_TRACE): [role][256] pairs (event code, SM clock) written by CTA 0
};
#define FM_TR(role, code)                                                                                   \
  do {                                                                                                      \
    if (p.trace && blockIdx.x == 0 && lane == 0 && tr_n < 256) {                                            \
      p.trace[((role) * 256 + tr_n) * 2] = (code);                                                          \
      p.trace[((role) * 256 + tr_n) * 2 + 1] = clock64();                                                   \
      ++tr_n;                                                                                               \
    }                                                                                                       \
  } while (0)

// CL = 2: the kernel runs as clusters of two CTAs that work on neighbouring row tiles and SHARE the weight stream: each CTA
// loads half of every W1 / W2 block and multicasts it into both CTAs' rings, so a CTA pulls half the weight bytes through its
// own L2 -> shared-memory path and the ring covers twice the latency (the weights, 332 KB / 1.3 MB per tile at C = 144 / 288,
// are the dominant on-chip traffic of this kernel).  A ring slot is free again when BOTH CTAs have consumed it: the "empty"
// barriers count two arrivals, one multicast commit from each CTA's MMA warp.
template <int C, int CL>
__global__ void __launch_bounds__(fm::THREADS, 1)
mlp_fused_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w1,
                 const __grid_constant__ CUtensorMap map_w2, const __grid_constant__ CUtensorMap map_x, const FusedMlpParams p) {
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
  uint64_t* s_full = w2_empty + K::NS2;         // [NSB]
  uint64_t* p_full = s_full + NSB;              // [NSB]
  uint64_t* o_full = p_full + NSB;              // [NOB]
  uint64_t* o_free = o_full + K::NOB;           // [NOB]
  uint64_t* box_full = o_free + K::NOB;         // [4 read-out warps][NBOX]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(box_full + 4 * K::NBOX);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int tr_n = 0;
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w1);
    tma_prefetch_desc(&map_w2);
    tma_prefetch_desc(&map_x);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(a_full, 1); mbar_init(a_empty, 1);
    for (int i = 0; i < K::NS1; ++i) { mbar_init(&w1_full[i], 1); mbar_init(&w1_empty[i], CL); }
    for (int i = 0; i < K::NS2; ++i) { mbar_init(&w2_full[i], 1); mbar_init(&w2_empty[i], CL); }
    for (int i = 0; i < NSB; ++i) { mbar_init(&s_full[i], 1); mbar_init(&p_full[i], 4); }
    for (int i = 0; i < K::NOB; ++i) { mbar_init(&o_full[i], 1); mbar_init(&o_free[i], 4); }
    for (int i = 0; i < 4 * K::NBOX; ++i) mbar_init(&box_full[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_ptr, 512); tmem_relinquish(); }
  tc_fence_before();
  if (CL == 2) cluster_sync_all(); else __syncthreads();     // the pair's barriers are initialised before any remote arrival
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int crank = CL == 2 ? (int)cluster_ctarank() : 0;
  constexpr uint16_t kMask = CL == 2 ? 3 : 1;

  if (warp == 0) {
    // ===================== TMA producer 1: the A tile and the W1 ring =====================
    if (lane == 0) {
      int i1 = 0, nt = 0;
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
        mbar_wait(a_empty, (nt & 1) ^ 1);
        mbar_arrive_expect_tx(a_full, K::KB * A_KB_BYTES);
#pragma unroll
        for (int kb = 0; kb < K::KB; ++kb) tma_load_3d(smem + K::OFF_A + kb * A_KB_BYTES, &map_a, a_full, kb * 64, t * BM, 0);
        for (int h = 0; h < K::NH; ++h, ++i1) {
          const int s = i1 % K::NS1;
          mbar_wait(&w1_empty[s], ((i1 / K::NS1) & 1) ^ 1);
          mbar_arrive_expect_tx(&w1_full[s], K::W1_STAGE);
#pragma unroll
          for (int kb = 0; kb < K::KB; ++kb) {
            uint8_t* dst = smem + K::OFF_W1 + s * K::W1_STAGE + kb * W1_BYTES;
            if (CL == 2)     // this CTA's half of the block (32 hidden units), into both rings
              tma_load_3d_mc(dst + crank * (W1_BYTES / 2), &map_w1, &w1_full[s], kb * 64, h * BH + crank * (BH / 2), 0, kMask);
            else
              tma_load_3d(dst, &map_w1, &w1_full[s], kb * 64, h * BH, 0);
          }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== TMA producer 2: the W2 ring (its own thread: a full W2 ring must not hold back W1) =====================
    if (lane == 0) {
      int i2 = 0;
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x) {
        for (int blk = 0; blk < K::NH; ++blk, ++i2) {
          const int s = i2 % K::NS2;
          mbar_wait(&w2_empty[s], ((i2 / K::NS2) & 1) ^ 1);
          mbar_arrive_expect_tx(&w2_full[s], K::W2_BYTES);
#pragma unroll
          for (int j = 0; j < K::NSPLIT; ++j) {
            if (CL == 2)     // 72 of the 144 rows of every piece (9 whole swizzle atoms), into both rings
              tma_load_3d_mc(smem + K::OFF_W2 + s * K::W2_BYTES + (j * 144 + crank * 72) * 128, &map_w2, &w2_full[s], blk * 64, j * 144 + crank * 72, 0, kMask);
            else
              tma_load_3d(smem + K::OFF_W2 + s * K::W2_BYTES + j * 144 * 128, &map_w2, &w2_full[s], blk * 64, j * 144, 0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp, one elected lane issues) =====================
    // The issuing thread is a resource of its own: every instruction it spends on index arithmetic delays the next tcgen05.mma
    // (the first version of this loop needed ~400 clocks per MMA against the 50 - 72 the tensor pipe takes).  Ring positions are
    // carried as (stage, phase) counters, descriptors are built once per shared-memory block and advanced by constants, and the
    // loop is run by the whole warp with warp-uniform values (common.cuh umma_*_w), so that the descriptors live in uniform
    // registers.  Order per tile: fc1(0), fc1(1), then for every chunk c: fc1(c + 2), wait P(c), fc2(c) - with three S buffers the
    // tensor pipe always holds the next two chunks' fc1 while the GELU warps work, and S_b is reused by chunk c + 3 in order
    // behind fc2(c), which read P(c) out of it.
    {
      constexpr uint32_t idesc_fc1 = make_idesc_ex(128, BH, 0);
      constexpr uint32_t idesc_fc2 = make_idesc_ex(128, 144, 0);
      const uint32_t sbase = smem_u32(smem);
      int s1 = 0, s2 = 0, nt = 0;
      uint32_t ph1 = 0, ph2 = 0;        // phases of the W1 / W2 rings
      int bi = 0, bc = 0;               // S buffer of the next fc1 / of the next fc2
      uint32_t php = 0;                 // phase of p_full[bc] (flips when bc wraps)
      auto issue_fc1 = [&]() {          // one hidden chunk: S_bi = A W1[h]^T over the KB k-blocks of the resident A tile
        const uint32_t d = tmem_base + COL_S + BH * bi;
        mbar_wait(&w1_full[s1], ph1);
        tc_fence_after();
        const uint64_t da0 = make_smem_desc_sw128(sbase + K::OFF_A);
        const uint64_t db0 = make_smem_desc_sw128(sbase + K::OFF_W1 + s1 * K::W1_STAGE);
#pragma unroll
        for (int kb = 0; kb < K::KB; ++kb) {
          constexpr int kFull = C / 64;
          const int ksteps = kb < kFull ? 4 : (C - kFull * 64) / 16;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (k < ksteps)
              umma_bf16_w(d, da0 + (uint64_t)(kb * A_KB_BYTES / 16 + 2 * k), db0 + (uint64_t)(kb * W1_BYTES / 16 + 2 * k), idesc_fc1, (kb | k) != 0);
        }
        if (CL == 2) umma_commit_mc_w(&w1_empty[s1], kMask); else umma_commit_w(&w1_empty[s1]);
        if (++s1 == K::NS1) { s1 = 0; ph1 ^= 1; }
        umma_commit_w(&s_full[bi]);
        if (++bi == NSB) bi = 0;
      };
      for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
        const int ob = nt % K::NOB;
        const uint32_t o_col = tmem_base + COL_O + ob * C;
        mbar_wait(a_full, nt & 1);
        tc_fence_after();
        FM_TR(0, 1000 + nt);
        issue_fc1();
        issue_fc1();
#pragma unroll 1
        for (int c = 0; c < K::NH; ++c) {
          if (c + 2 < K::NH) issue_fc1();
          FM_TR(0, 100 + c);
          if (c + 2 == K::NH - 1) umma_commit_w(a_empty);                    // the last fc1 of the tile has been issued: A may be refilled
          mbar_wait(&p_full[bc], php);
          FM_TR(0, 200 + c);
          if (c == 0) mbar_wait(&o_free[ob], ((nt / K::NOB) & 1) ^ 1);      // this accumulator's previous tile has been read out
          mbar_wait(&w2_full[s2], ph2);
          tc_fence_after();
          FM_TR(0, 300 + c);
          const uint32_t pa = tmem_base + COL_S + BH * bc;                   // P(c): 64 hidden units = 32 columns of bf16 pairs
          const uint64_t dw = make_smem_desc_sw128(sbase + K::OFF_W2 + s2 * K::W2_BYTES);
#pragma unroll
          for (int k = 0; k < BH / 16; ++k) {
#pragma unroll
            for (int j = 0; j < K::NSPLIT; ++j)
              umma_bf16_ts_w(o_col + j * 144, pa + k * 8, dw + 2 * k + (uint64_t)(j * 144 * 128 / 16), idesc_fc2, (c | k) != 0);
          }
          if (CL == 2) umma_commit_mc_w(&w2_empty[s2], kMask); else umma_commit_w(&w2_empty[s2]);
          if (++s2 == K::NS2) { s2 = 0; ph2 ^= 1; }
          if (++bc == NSB) { bc = 0; php ^= 1; }
        }
        umma_commit_w(&o_full[ob]);
      }
    }
  } else if (warp >= 4 && warp < 12) {
    // ===================== GELU warps =====================
    // Chunks are numbered through the whole kernel (n = tile count * NH + c); group `grp` takes the chunks of its parity, chunk n
    // lives in S buffer n % 3.
    const int q = warp & 3, grp = (warp - 4) >> 2;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    int nt = 0;
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
      const long row = (long)t * BM + q * 32 + lane;
      float rstd = 1.f;
      if (row < p.M) {
        const float2* st = reinterpret_cast<const float2*>(p.stats_in) + row * p.parts_in;
        float s1 = 0.f, s2 = 0.f;
        for (int k = 0; k < p.parts_in; ++k) { const float2 v = __ldg(st + k); s1 += v.x; s2 += v.y; }
        const float mu = s1 * p.inv_c;
        rstd = rsqrtf(fmaxf(fmaf(s2, p.inv_c, -mu * mu), 0.f) + p.eps);
      }
      const uint32_t n0 = (uint32_t)nt * K::NH;
      for (int c = (int)((n0 + grp) & 1); c < K::NH; c += 2) {
        const uint32_t n = n0 + c, buf = n % NSB;
        const uint32_t ts = tmem_base + COL_S + BH * buf + lane_sel;
        if (q == 0) FM_TR(1 + grp, 100 + c);
        mbar_wait(&s_full[buf], (n / NSB) & 1);
        tc_fence_after();
        if (q == 0) FM_TR(1 + grp, 200 + c);
        uint32_t r0[32], r1[32];
        tmem_ld_x32(ts, r0);
        tmem_ld_x32(ts + 32, r1);
        // 32 columns at a time: P (16 columns of bf16 pairs) goes back over columns of S this warp has already read
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t (&r)[32] = half == 0 ? r0 : r1;
          uint32_t pk[32];
          const float4* bsrc = reinterpret_cast<const float4*>(p.b1 + c * BH + half * 32);
          float4 bv[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) bv[j] = __ldg(bsrc + j);
          if (half == 0) tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float v0, v1, v2, v3;
            gelu_fast_bf16x2(fmaf(__uint_as_float(r[4 * j]), rstd, bv[j].x), fmaf(__uint_as_float(r[4 * j + 1]), rstd, bv[j].y), v0, v1);
            gelu_fast_bf16x2(fmaf(__uint_as_float(r[4 * j + 2]), rstd, bv[j].z), fmaf(__uint_as_float(r[4 * j + 3]), rstd, bv[j].w), v2, v3);
            pk[2 * j] = pack_bf16x2(v0, v1);
            pk[2 * j + 1] = pack_bf16x2(v2, v3);
          }
          tmem_st_x16(ts + half * 16, pk);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&p_full[buf]);
        if (q == 0) FM_TR(1 + grp, 300 + c);
      }
    }
  } else if (warp >= 12) {
    // ===================== read-out warps =====================
    const int q = warp & 3;
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t box0 = smem_u32(smem + K::OFF_BOX) + q * K::NBOX * BOX_BYTES;
    uint64_t* my_full = box_full + q * K::NBOX;
    const uint32_t row_off = (uint32_t)lane * 128, xr = (uint32_t)(lane & 7) << 4;
    uint32_t uses[K::NBOX];                                   // residual boxes received per slot (phase of my_full[slot])
#pragma unroll
    for (int i = 0; i < K::NBOX; ++i) uses[i] = 0;
    auto request = [&](int slot, int tile, int ch) {          // lane 0: residual chunk `ch` of `tile` -> box `slot`
      mbar_arrive_expect_tx(&my_full[slot], BOX_BYTES);
      tma_load_3d_addr(box0 + slot * BOX_BYTES, &map_x, &my_full[slot], ch * 32, tile * BM + q * 32, 0);
    };
    if (lane == 0 && (int)blockIdx.x < p.num_tiles) {
#pragma unroll
      for (int i = 0; i < K::NBOX; ++i) request(i, blockIdx.x, i);
    }
    int nt = 0;
    for (int t = blockIdx.x; t < p.num_tiles; t += gridDim.x, ++nt) {
      const int ob = nt % K::NOB;
      const long row = (long)t * BM + q * 32 + lane;
      const uint32_t to = tmem_base + COL_O + ob * C + lane_sel;
      bf16* xbr = p.xb_out + row * C;
      float s1 = 0.f, s2 = 0.f;
      if (q == 0) FM_TR(3, 100 + nt);
      mbar_wait(&o_full[ob], (nt / K::NOB) & 1);
      tc_fence_after();
      if (q == 0) FM_TR(3, 200 + nt);
#pragma unroll
      for (int ch = 0; ch < K::NCH; ++ch) {
        constexpr int kFull = C / 32;                          // chunks with all 32 columns inside the row
        const int slot = ch % K::NBOX;
        const int nq = ch < kFull ? 8 : (C - kFull * 32) / 4;  // float4 groups of this chunk: 8, or 4 in the half chunk
        uint32_t r[32];
        if (ch < kFull) tmem_ld_x32(to + ch * 32, r); else tmem_ld_x16(to + ch * 32, r);
        tmem_ld_wait();
        if (ch == K::NCH - 1) {                                // O has left tensor memory: fc2 of a later tile may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&o_free[ob]);
        }
        mbar_wait(&my_full[slot], uses[slot] & 1);
        ++uses[slot];
        const uint32_t box = box0 + slot * BOX_BYTES + row_off;
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j < nq) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.b2 + ch * 32) + j);
            const uint32_t addr = box + (((uint32_t)j << 4) ^ xr);
            const float4 x = ld_shared_f4(addr);
            float4 v;
            v.x = __uint_as_float(r[4 * j]) + b.x + x.x;
            v.y = __uint_as_float(r[4 * j + 1]) + b.y + x.y;
            v.z = __uint_as_float(r[4 * j + 2]) + b.z + x.z;
            v.w = __uint_as_float(r[4 * j + 3]) + b.w + x.w;
            s1 += (v.x + v.y) + (v.z + v.w);
            s2 = fmaf(v.x, v.x, s2); s2 = fmaf(v.y, v.y, s2); s2 = fmaf(v.z, v.z, s2); s2 = fmaf(v.w, v.w, s2);
            st_shared_u4(addr, __float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w));
            pk[2 * j] = pack_bf16x2(v.x, v.y);
            pk[2 * j + 1] = pack_bf16x2(v.z, v.w);
          }
        }
        if (row < p.M) {                                       // bf16 copy: 64 B (32 B in the half chunk) per row, straight from registers
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (2 * j < nq) *reinterpret_cast<uint4*>(xbr + ch * 32 + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(&map_x, box0 + slot * BOX_BYTES, ch * 32, t * BM + q * 32, 0);     // clips the columns / rows outside x
          bulk_commit();
          if (K::NBOX < K::NCH && ch + K::NBOX < K::NCH) {     // ring: the box is refilled as soon as its store has read it
            bulk_wait_read<0>();
            request(slot, t, ch + K::NBOX);
          }
        }
      }
      if (q == 0) FM_TR(3, 300 + nt);
      if (row < p.M) reinterpret_cast<float2*>(p.stats_out)[row] = make_float2(s1, s2);
      if (lane == 0 && t + (int)gridDim.x < p.num_tiles) {     // next tile's residual rows: a whole tile of lead time
        bulk_wait_read<0>();
#pragma unroll
        for (int i = 0; i < K::NBOX; ++i) request(i, t + gridDim.x, i);
      }
    }
    if (lane == 0) bulk_wait_read<0>();                        // staging smem must outlive the last store's read
  }

  tc_fence_before();
  if (CL == 2) cluster_sync_all(); else __syncthreads();     // the peer's rings and barriers stay alive until both CTAs are done
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

// x [rows][C] fp32 row-major, box {32 columns, 32 rows, 1}, 128B swizzle: loads zero-fill and stores clip what lies outside
static int fm_map_x(CUtensorMap* m, float* ptr, long rows, long C) {
  EncodeTiledFnM enc = fm_encode_fn();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return 1; }
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (C * 4) % 16) { set_error("mlp_fused: x not 16 B aligned"); return 1; }
  const cuuint64_t dims[3] = {(cuuint64_t)C, (cuuint64_t)rows, 1};
  const cuuint64_t strides[2] = {(cuuint64_t)C * 4, (cuuint64_t)C * 4 * (cuuint64_t)rows};
  const cuuint32_t box[3] = {32, 32, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("mlp_fused: cuTensorMapEncodeTiled(x) failed (%d)", (int)r); return 1; }
  return 0;
}

template <int C, int CL>
static int launch_mlp_fused(const CUtensorMap& ma, const CUtensorMap& m1, const CUtensorMap& m2, const CUtensorMap& mx,
                            const FusedMlpParams& p, cudaStream_t stream) {
  using K = fm::Cfg<C>;
  static PerDeviceFlag attr;
  auto kern = mlp_fused_kernel<C, CL>;
  if (!attr.is_set()) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, K::SMEM);
    if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(mlp_fused<%d,%d>): %s", C, CL, cudaGetErrorString(e)); return 1; }
    attr.set();
  }
  int grid = p.num_tiles < num_sms() ? p.num_tiles : num_sms();
  if (CL == 2) grid &= ~1;                      // whole pairs; both CTAs of a pair then walk the same number of tiles
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(fm::THREADS); cfg.dynamicSmemBytes = K::SMEM; cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, m1, m2, mx, p);
  if (e != cudaSuccess) { set_error("cudaLaunchKernelEx(mlp_fused<%d,%d>): %s", C, CL, cudaGetErrorString(e)); return 1; }
  return 0;
}

// Stage 1 only (C = 144).  At C = 288 the same design is bound by MMA issue (fc1 as N = 64 instructions) and by the single O
// accumulator that fits next to three S buffers, and lands at the speed of the two HBM-bound GEMMs it would replace (measured:
// 1.11 - 1.15 ms against 0.99 - 1.24 ms per block per 32 images, profiles/r2u_mlp_fused_trace.md), so stage 2 keeps the GEMM pair.
bool mlp_fused_supported(int C) { return C == 144; }

// x[M][C] (fp32, in place) += fc2(GELU(rstd(row) * (xb W1''^T) + b1)) + b2; xb_out = bf16(x); stats_out[M][1][2] (one part per row)
int mlp_fused(const bf16* xb, const float* stats_in, int parts_in, float eps, const bf16* w1, const float* b1, const bf16* w2,
              const float* b2, float* x, bf16* xb_out, float* stats_out, long M, int C, cudaStream_t stream) {
  if (!mlp_fused_supported(C) || M <= 0) { set_error("mlp_fused: unsupported C=%d M=%ld", C, M); return 1; }
  CUtensorMap ma, m1, m2, mx;
  const int num_tiles = (int)((M + fm::BM - 1) / fm::BM);
  const bool pairs = num_tiles % 2 == 0 && num_tiles >= 2;      // CTA pairs sharing the weight stream need tiles in twos
  if (fm_map(&ma, xb, M, C, C, fm::BM) || fm_map(&m1, w1, 4L * C, C, C, pairs ? fm::BH / 2 : fm::BH) ||
      fm_map(&m2, w2, C, 4L * C, 4L * C, pairs ? 72 : 144) || fm_map_x(&mx, x, M, C)) return 1;
  FusedMlpParams p;
  p.stats_in = stats_in; p.parts_in = parts_in; p.eps = eps; p.inv_c = 1.0f / (float)C;
  p.b1 = b1; p.b2 = b2; p.xb_out = xb_out; p.stats_out = stats_out;
  p.M = (int)M; p.num_tiles = (int)((M + fm::BM - 1) / fm::BM);
  p.trace = nullptr;
  static long long* trace_dev = nullptr;
  const bool tracing = getenv("SAM2B_MLP_TRACE") != nullptr;
  if (tracing) {
    if (!trace_dev) cudaMalloc(&trace_dev, 4 * 256 * 2 * sizeof(long long));
    cudaMemsetAsync(trace_dev, 0, 4 * 256 * 2 * sizeof(long long), stream);
    p.trace = trace_dev;
  }
  char nm[64];
  if (g_prof_on) snprintf(nm, sizeof(nm), "mlp_fused m%ld c%d", M, C);
  ProfRange pr(stream, nm, 16.0 * (double)M * C * C, (double)M * C * (2.0 + 4.0 + 4.0 + 2.0));
  const int rc = pairs ? launch_mlp_fused<144, 2>(ma, m1, m2, mx, p, stream) : launch_mlp_fused<144, 1>(ma, m1, m2, mx, p, stream);
  if (rc) return rc;
  SAM2B_CHECK_LAUNCH("mlp_fused");
  if (tracing) {       // development: dump CTA 0's event trace of this launch
    static long long host[4 * 256 * 2];
    cudaStreamSynchronize(stream);
    cudaMemcpy(host, trace_dev, sizeof(host), cudaMemcpyDeviceToHost);
    static const char* roles[4] = {"mma", "gelu0", "gelu1", "readout"};
    long long t0 = 0;
    for (int r = 0; r < 4; ++r) for (int i = 0; i < 256; ++i) if (host[(r * 256 + i) * 2 + 1] && (!t0 || host[(r * 256 + i) * 2 + 1] < t0)) t0 = host[(r * 256 + i) * 2 + 1];
    for (int r = 0; r < 4; ++r) {
      fprintf(stderr, "MLPTRACE C=%d %s:", C, roles[r]);
      for (int i = 0; i < 256 && host[(r * 256 + i) * 2 + 1]; ++i) fprintf(stderr, " %lld@%lld", host[(r * 256 + i) * 2], host[(r * 256 + i) * 2 + 1] - t0);
      fprintf(stderr, "\n");
    }
  }
  return 0;
}

}  // namespace sam2b
