// K2/K3 (first implementation): Hiera MultiScaleAttention core for head_dim 72 (zero-padded to 80 in shared
// memory), reference hieradet.py:56-81 + window_partition/unpartition (backbones/utils.py:16-60) + q max-pool
// (hieradet.py:25-36,66-70).  Flash-style online softmax, fp32 statistics, bf16 mma.sync m16n8k16 tiles.
//
//  * window partition / unpartition are folded into the addressing (no permuted copies);
//  * q-pooling (2x2 max of q inside the window) is applied while loading Q;
//  * one CTA = 4 warps = 64 query rows; K/V stream through a double-buffered cp.async ring in chunks of 64 rows.
//    SHARED mode  (>= 64 queries per window): the 4 warps share each 64-key chunk of one window;
//    PRIVATE mode (16 or 4 queries per window): warp w owns window 4*tile+w and rows 16w..16w+15 of each chunk.
//
// This legacy-tensor-path kernel carries ~11 % of the encoder's FLOPs; the tcgen05 flash kernel for the three
// global blocks is the planned replacement.
#include <cstdio>
#include <cstdlib>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

static constexpr int HD = 72;
static constexpr int LDS = 88;        // smem row stride in bf16 (176 B: conflict-free ldmatrix)
static constexpr int ATT_THREADS = 128;

SAM2B_DEVINL void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
SAM2B_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
SAM2B_DEVINL void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

SAM2B_DEVINL void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
SAM2B_DEVINL void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
SAM2B_DEVINL void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

struct AttnParams {
  const bf16* qkv; bf16* out;
  int B, H, W, heads, ws, q_pool;
  int QW, KW;        // queries / keys per window
  int nwx;           // windows per row
  float scale_log2;  // (1/sqrt(72)) * log2(e)
};

SAM2B_DEVINL long key_token(const AttnParams& p, int b, int win, int j) {
  const int wy = win / p.nwx, wx = win - wy * p.nwx;
  const int ky = j / p.ws, kx = j - ky * p.ws;
  return ((long)b * p.H + wy * p.ws + ky) * p.W + wx * p.ws + kx;
}

template <bool PRIVATE>
__global__ void __launch_bounds__(ATT_THREADS) window_attn_hd72_kernel(const AttnParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  bf16* sQ = reinterpret_cast<bf16*>(smem_raw);             // [64][LDS]
  bf16* sK = sQ + 64 * LDS;                                 // [2][64][LDS]
  bf16* sV = sK + 2 * 64 * LDS;                             // [2][64][LDS]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int head = blockIdx.x, tile = blockIdx.y, b = blockIdx.z;      // head fastest: co-scheduled CTAs share sectors / DRAM pages
  const int C = p.heads * HD;
  const long row_stride = 3L * C;                           // elements between consecutive tokens in qkv
  const bf16* qbase = p.qkv + head * HD;
  const bf16* kbase = p.qkv + C + head * HD;
  const bf16* vbase = p.qkv + 2 * C + head * HD;

  // zero the pad columns [72,88) of every smem row once (cp.async / Q stores only touch [0,72))
  for (int i = tid; i < 5 * 64; i += ATT_THREADS) {
    uint4* z = reinterpret_cast<uint4*>(sQ + i * LDS + HD);
    z[0] = make_uint4(0, 0, 0, 0);
    z[1] = make_uint4(0, 0, 0, 0);
  }

  const int tiles_per_win = PRIVATE ? 1 : p.QW / 64;
  auto row_window = [&](int r) { return PRIVATE ? tile * 4 + (r >> 4) : tile / tiles_per_win; };
  auto row_qi = [&](int r) { return PRIVATE ? (r & 15) : (tile % tiles_per_win) * 64 + r; };

  auto load_kv = [&](int it, int buf) {
    for (int i = tid; i < 64 * 9; i += ATT_THREADS) {
      const int r = i / 9, c = i - r * 9;
      int win, j;
      if (PRIVATE) { win = tile * 4 + (r >> 4); j = it * 16 + (r & 15); }
      else { win = tile / tiles_per_win; j = it * 64 + r; }
      const long tok = key_token(p, b, win, j);
      cp_async16(sK + (buf * 64 + r) * LDS + c * 8, kbase + tok * row_stride + c * 8);
      cp_async16(sV + (buf * 64 + r) * LDS + c * 8, vbase + tok * row_stride + c * 8);
    }
    cp_async_commit();
  };

  const int n_iter = PRIVATE ? p.KW / 16 : p.KW / 64;
  load_kv(0, 0);

  // ---- Q tile (with optional 2x2 max pool) -> smem ----
  for (int i = tid; i < 64 * 9; i += ATT_THREADS) {
    const int r = i / 9, c = i - r * 9;
    const int win = row_window(r), qi = row_qi(r);
    uint4 v = make_uint4(0, 0, 0, 0);
    if (qi < p.QW) {
      if (!p.q_pool) {
        v = *reinterpret_cast<const uint4*>(qbase + key_token(p, b, win, qi) * row_stride + c * 8);
      } else {
        const int half = p.ws >> 1;
        const int wy = win / p.nwx, wx = win - wy * p.nwx;
        const int py = qi / half, px = qi - py * half;
        const long t00 = ((long)b * p.H + wy * p.ws + 2 * py) * p.W + wx * p.ws + 2 * px;
        const uint4 a0 = *reinterpret_cast<const uint4*>(qbase + t00 * row_stride + c * 8);
        const uint4 a1 = *reinterpret_cast<const uint4*>(qbase + (t00 + 1) * row_stride + c * 8);
        const uint4 a2 = *reinterpret_cast<const uint4*>(qbase + (t00 + p.W) * row_stride + c * 8);
        const uint4 a3 = *reinterpret_cast<const uint4*>(qbase + (t00 + p.W + 1) * row_stride + c * 8);
        const bf162* x0 = reinterpret_cast<const bf162*>(&a0);
        const bf162* x1 = reinterpret_cast<const bf162*>(&a1);
        const bf162* x2 = reinterpret_cast<const bf162*>(&a2);
        const bf162* x3 = reinterpret_cast<const bf162*>(&a3);
        bf162* o = reinterpret_cast<bf162*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) o[e] = __hmax2(__hmax2(x0[e], x1[e]), __hmax2(x2[e], x3[e]));
      }
    }
    *reinterpret_cast<uint4*>(sQ + r * LDS + c * 8) = v;
  }
  __syncthreads();

  // ---- Q fragments (A operand), 5 k-steps of 16 over the padded head_dim 80 ----
  uint32_t qa[5][4];
  {
    const uint32_t base = smem_u32(sQ + (warp * 16 + (lane & 15)) * LDS + (lane >> 4) * 8);
#pragma unroll
    for (int ks = 0; ks < 5; ++ks) ldsm_x4(base + ks * 32, qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
  }

  constexpr int NT = PRIVATE ? 2 : 8;   // key n-tiles (of 8) per chunk for this warp
  const int key_lo = PRIVATE ? warp * 16 : 0;
  float o_acc[10][4];
#pragma unroll
  for (int i = 0; i < 10; ++i) { o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f; }
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

  for (int it = 0; it < n_iter; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_iter) { load_kv(it + 1, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();

    // S = Q K^T
    float s[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
    const bf16* kb = sK + (buf * 64 + key_lo) * LDS;
#pragma unroll
    for (int ks = 0; ks < 5; ++ks) {
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        uint32_t b0, b1, b2, b3;
        const uint32_t addr = smem_u32(kb + (8 * (n + (lane >> 4)) + (lane & 7)) * LDS + ks * 16 + ((lane >> 3) & 1) * 8);
        ldsm_x4(addr, b0, b1, b2, b3);
        mma_bf16_16816(s[n], qa[ks], b0, b1);
        mma_bf16_16816(s[n + 1], qa[ks], b2, b3);
      }
    }
    // online softmax (rows g = lane/4 and g+8; this thread holds 2 columns of each n-tile)
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      mx[0] = fmaxf(mx[0], fmaxf(s[n][0], s[n][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[n][2], s[n][3]));
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
    }
    float corr[2], msc[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float m_new = fmaxf(m_run[h], mx[h] * p.scale_log2);
      corr[h] = exp2f(m_run[h] - m_new);
      m_run[h] = m_new;
      msc[h] = m_new;
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      s[n][0] = exp2f(s[n][0] * p.scale_log2 - msc[0]);
      s[n][1] = exp2f(s[n][1] * p.scale_log2 - msc[0]);
      s[n][2] = exp2f(s[n][2] * p.scale_log2 - msc[1]);
      s[n][3] = exp2f(s[n][3] * p.scale_log2 - msc[1]);
      rs[0] += s[n][0] + s[n][1];
      rs[1] += s[n][2] + s[n][3];
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) l_run[h] = l_run[h] * corr[h] + rs[h];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      o_acc[i][0] *= corr[0]; o_acc[i][1] *= corr[0];
      o_acc[i][2] *= corr[1]; o_acc[i][3] *= corr[1];
    }
    // O += P V
    const bf16* vb = sV + (buf * 64 + key_lo) * LDS;
#pragma unroll
    for (int kk = 0; kk < NT / 2; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dt = 0; dt < 10; dt += 2) {
        uint32_t b0, b1, b2, b3;
        const uint32_t addr = smem_u32(vb + (16 * kk + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + 8 * (dt + (lane >> 4)));
        ldsm_x4_t(addr, b0, b1, b2, b3);
        mma_bf16_16816(o_acc[dt], pa, b0, b1);
        mma_bf16_16816(o_acc[dt + 1], pa, b2, b3);
      }
    }
    __syncthreads();   // all warps done with buf before it is refilled by the next prefetch
  }

  // ---- finalise: O / l -> bf16 -> smem (own 16 rows of sQ) -> coalesced 16 B stores ----
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
  }
  const float inv0 = 1.f / l_run[0], inv1 = 1.f / l_run[1];
  {
    const int g = lane >> 2, tq = lane & 3;
    bf16* r0 = sQ + (warp * 16 + g) * LDS;
    bf16* r1 = sQ + (warp * 16 + g + 8) * LDS;
#pragma unroll
    for (int dt = 0; dt < 9; ++dt) {
      *reinterpret_cast<uint32_t*>(r0 + dt * 8 + tq * 2) = pack_bf16x2(o_acc[dt][0] * inv0, o_acc[dt][1] * inv0);
      *reinterpret_cast<uint32_t*>(r1 + dt * 8 + tq * 2) = pack_bf16x2(o_acc[dt][2] * inv1, o_acc[dt][3] * inv1);
    }
  }
  __syncwarp();
  const int Ho = p.q_pool ? p.H >> 1 : p.H, Wo = p.q_pool ? p.W >> 1 : p.W;
  const int wso = p.q_pool ? p.ws >> 1 : p.ws;
  for (int i = lane; i < 16 * 9; i += 32) {
    const int rr = i / 9, c = i - rr * 9;
    const int r = warp * 16 + rr;
    const int win = row_window(r), qi = row_qi(r);
    if (qi < p.QW) {
      const int wy = win / p.nwx, wx = win - wy * p.nwx;
      const int py = qi / wso, px = qi - py * wso;
      const long tok = ((long)b * Ho + wy * wso + py) * Wo + wx * wso + px;
      *reinterpret_cast<uint4*>(p.out + tok * C + head * HD + c * 8) = *reinterpret_cast<const uint4*>(sQ + r * LDS + c * 8);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// Persistent variant for the single-chunk cases (8x8 windows: 64 keys shared by 64 queries; 4x4 windows: four windows of
// 16 tokens per tile), no q-pooling.  These launches are pure HBM traffic (read qkv once, write out once), so what
// matters is bytes in flight: every CTA loops over work items and prefetches the Q / K / V rows of its NEXT item with
// cp.async into the other half of a double buffer while the tensor cores work on the current one.
// ------------------------------------------------------------------------------------------------------------
template <bool PRIVATE>
__global__ void __launch_bounds__(ATT_THREADS) window_attn_hd72_stream_kernel(const AttnParams p, int tiles_per_img, long n_items) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  bf16* sbuf = reinterpret_cast<bf16*>(smem_raw);           // [2][3][64][LDS]: Q, K, V of two items
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int C = p.heads * HD;
  const long row_stride = 3L * C;
  for (int i = tid; i < 6 * 64; i += ATT_THREADS) {          // pad columns [72,88) of every row stay zero
    uint4* z = reinterpret_cast<uint4*>(sbuf + i * LDS + HD);
    z[0] = make_uint4(0, 0, 0, 0);
    z[1] = make_uint4(0, 0, 0, 0);
  }
  auto decode = [&](long it, int& tile, int& head, int& b) {
    // head fastest: the heads of one window are adjacent 144 B segments of the same rows, so CTAs that run at the same
    // time share the 32 B sectors straddling two heads and the DRAM pages
    head = (int)(it % p.heads);
    const long r = it / p.heads;
    tile = (int)(r % tiles_per_img);
    b = (int)(r / tiles_per_img);
  };
  auto row_token = [&](int tile, int b, int r) {            // token index of row r (query row == key row in these cases)
    const int win = PRIVATE ? tile * 4 + (r >> 4) : tile;
    const int j = PRIVATE ? (r & 15) : r;
    return key_token(p, b, win, j);
  };
  auto prefetch = [&](long it, int buf) {
    int tile, head, b;
    decode(it, tile, head, b);
    bf16* dst = sbuf + buf * 3 * 64 * LDS;
    const bf16* src = p.qkv + head * HD;
    for (int i = tid; i < 64 * 9; i += ATT_THREADS) {
      const int r = i / 9, c = i - r * 9;
      const bf16* g = src + row_token(tile, b, r) * row_stride + c * 8;
      cp_async16(dst + r * LDS + c * 8, g);
      cp_async16(dst + (64 + r) * LDS + c * 8, g + C);
      cp_async16(dst + (128 + r) * LDS + c * 8, g + 2 * C);
    }
    cp_async_commit();
  };
  long it = blockIdx.x;
  if (it < n_items) prefetch(it, 0);
  int buf = 0;
  for (; it < n_items; it += gridDim.x, buf ^= 1) {
    const long nxt = it + gridDim.x;
    if (nxt < n_items) { prefetch(nxt, buf ^ 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    int tile, head, b;
    decode(it, tile, head, b);
    bf16* sQ = sbuf + buf * 3 * 64 * LDS;
    const bf16* sK = sQ + 64 * LDS;
    const bf16* sV = sQ + 128 * LDS;
    uint32_t qa[5][4];
    {
      const uint32_t base = smem_u32(sQ + (warp * 16 + (lane & 15)) * LDS + (lane >> 4) * 8);
#pragma unroll
      for (int ks = 0; ks < 5; ++ks) ldsm_x4(base + ks * 32, qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
    }
    constexpr int NT = PRIVATE ? 2 : 8;
    const int key_lo = PRIVATE ? warp * 16 : 0;
    float s[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
    const bf16* kb = sK + key_lo * LDS;
#pragma unroll
    for (int ks = 0; ks < 5; ++ks) {
#pragma unroll
      for (int n = 0; n < NT; n += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(smem_u32(kb + (8 * (n + (lane >> 4)) + (lane & 7)) * LDS + ks * 16 + ((lane >> 3) & 1) * 8), b0, b1, b2, b3);
        mma_bf16_16816(s[n], qa[ks], b0, b1);
        mma_bf16_16816(s[n + 1], qa[ks], b2, b3);
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      mx[0] = fmaxf(mx[0], fmaxf(s[n][0], s[n][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[n][2], s[n][3]));
    }
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 1));
      mx[h] = fmaxf(mx[h], __shfl_xor_sync(0xffffffffu, mx[h], 2));
      mx[h] *= p.scale_log2;
    }
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      s[n][0] = ex2_approx(fmaf(s[n][0], p.scale_log2, -mx[0]));
      s[n][1] = ex2_approx(fmaf(s[n][1], p.scale_log2, -mx[0]));
      s[n][2] = ex2_approx(fmaf(s[n][2], p.scale_log2, -mx[1]));
      s[n][3] = ex2_approx(fmaf(s[n][3], p.scale_log2, -mx[1]));
      rs[0] += s[n][0] + s[n][1];
      rs[1] += s[n][2] + s[n][3];
    }
    float o_acc[10][4];
#pragma unroll
    for (int i = 0; i < 10; ++i) { o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f; }
    const bf16* vb = sV + key_lo * LDS;
#pragma unroll
    for (int kk = 0; kk < NT / 2; ++kk) {
      uint32_t pa[4];
      pa[0] = pack_bf16x2(s[2 * kk][0], s[2 * kk][1]);
      pa[1] = pack_bf16x2(s[2 * kk][2], s[2 * kk][3]);
      pa[2] = pack_bf16x2(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      pa[3] = pack_bf16x2(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dt = 0; dt < 10; dt += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(smem_u32(vb + (16 * kk + (lane & 7) + ((lane >> 3) & 1) * 8) * LDS + 8 * (dt + (lane >> 4))), b0, b1, b2, b3);
        mma_bf16_16816(o_acc[dt], pa, b0, b1);
        mma_bf16_16816(o_acc[dt + 1], pa, b2, b3);
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      rs[h] += __shfl_xor_sync(0xffffffffu, rs[h], 1);
      rs[h] += __shfl_xor_sync(0xffffffffu, rs[h], 2);
    }
    const float inv0 = 1.f / rs[0], inv1 = 1.f / rs[1];
    __syncwarp();        // every lane of this warp has read its Q fragments: rows 16w..16w+15 of sQ become the output staging
    {
      const int g = lane >> 2, tq = lane & 3;
      bf16* r0 = sQ + (warp * 16 + g) * LDS;
      bf16* r1 = sQ + (warp * 16 + g + 8) * LDS;
#pragma unroll
      for (int dt = 0; dt < 9; ++dt) {
        *reinterpret_cast<uint32_t*>(r0 + dt * 8 + tq * 2) = pack_bf16x2(o_acc[dt][0] * inv0, o_acc[dt][1] * inv0);
        *reinterpret_cast<uint32_t*>(r1 + dt * 8 + tq * 2) = pack_bf16x2(o_acc[dt][2] * inv1, o_acc[dt][3] * inv1);
      }
    }
    __syncwarp();
    for (int i = lane; i < 16 * 9; i += 32) {
      const int rr = i / 9, c = i - rr * 9;
      const int r = warp * 16 + rr;
      *reinterpret_cast<uint4*>(p.out + row_token(tile, b, r) * C + head * HD + c * 8) = *reinterpret_cast<const uint4*>(sQ + r * LDS + c * 8);
    }
    __syncthreads();     // this buffer is refilled by the prefetch issued at the top of the next iteration
  }
}

int window_attention_hd72(const bf16* qkv, bf16* out, int B, int H, int W, int heads, int ws, int q_pool, cudaStream_t stream) {
  if (ws <= 0 || H % ws || W % ws || (q_pool && (ws & 1))) { set_error("window_attention: bad window %d for %dx%d", ws, H, W); return 1; }
  // 16x16 windows and global attention run on the tcgen05 flash kernel (attn_tcgen05.cu)
  if (flash_attention_hd72_supported(H, W, ws, q_pool) && !getenv("SAM2B_LEGACY_ATTN")) return flash_attention_hd72(qkv, out, B, H, W, heads, ws, stream);
  AttnParams p;
  p.qkv = qkv; p.out = out; p.B = B; p.H = H; p.W = W; p.heads = heads; p.ws = ws; p.q_pool = q_pool;
  p.KW = ws * ws; p.QW = q_pool ? p.KW / 4 : p.KW; p.nwx = W / ws;
  p.scale_log2 = 1.4426950408889634f / sqrtf((float)HD);
  const int nwin = (H / ws) * (W / ws);
  const size_t smem = 5 * 64 * LDS * sizeof(bf16);
  static PerDeviceFlag attr;
  if (!attr.is_set()) {
    cudaFuncSetAttribute(window_attn_hd72_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(window_attn_hd72_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    attr.set();
  }
  char nm[96];
  if (g_prof_on) snprintf(nm, sizeof(nm), "attn h%d heads%d ws%d pool%d b%d", H, heads, ws, q_pool, B);
  const double nq = (double)B * nwin * p.QW * heads;
  ProfRange pr(stream, nm, 4.0 * nq * p.KW * HD, ((double)B * H * W * 3 + (double)B * nwin * p.QW) * heads * HD * 2.0);
  if (!q_pool && ((p.KW == 64) || (p.KW == 16 && nwin % 4 == 0)) && !getenv("SAM2B_LEGACY_WINDOW")) {
    // single-chunk windows: persistent streaming kernel (3 CTAs per SM, next item prefetched while this one computes)
    const bool priv = p.KW == 16;
    const int tiles_per_img = priv ? nwin / 4 : nwin;
    const long n_items = (long)tiles_per_img * heads * B;
    const size_t smem2 = 6 * 64 * LDS * sizeof(bf16);
    static PerDeviceFlag attr2;
    if (!attr2.is_set()) {
      cudaFuncSetAttribute(window_attn_hd72_stream_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
      cudaFuncSetAttribute(window_attn_hd72_stream_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
      attr2.set();
    }
    const long max_ctas = (long)num_sms() * 3;
    const int grid = (int)(n_items < max_ctas ? n_items : max_ctas);
    if (priv) window_attn_hd72_stream_kernel<true><<<grid, ATT_THREADS, smem2, stream>>>(p, tiles_per_img, n_items);
    else window_attn_hd72_stream_kernel<false><<<grid, ATT_THREADS, smem2, stream>>>(p, tiles_per_img, n_items);
  } else if (p.QW >= 64) {
    if (p.QW % 64 || p.KW % 64) { set_error("window_attention: QW=%d KW=%d not multiples of 64", p.QW, p.KW); return 1; }
    dim3 grid(heads, nwin * (p.QW / 64), B);
    window_attn_hd72_kernel<false><<<grid, ATT_THREADS, smem, stream>>>(p);
  } else {
    if (p.QW > 16 || p.KW % 16 || nwin % 4) { set_error("window_attention: unsupported small window QW=%d KW=%d nwin=%d", p.QW, p.KW, nwin); return 1; }
    dim3 grid(heads, nwin / 4, B);
    window_attn_hd72_kernel<true><<<grid, ATT_THREADS, smem, stream>>>(p);
  }
  SAM2B_CHECK_LAUNCH("window_attn_hd72");
  return 0;
}

}  // namespace sam2b
