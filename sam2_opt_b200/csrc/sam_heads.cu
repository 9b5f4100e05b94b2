// SAM heads: prompt encoder (points), two-way mask decoder, up-scaling + hyper-network masks,
// IoU / object-score / object-pointer heads -- the kernels around the tcgen05 GEMMs that engine.cu's sam_heads_mb() strings
// together.  Reference (paths relative to the reference tree's sam2/sam2/): modeling/sam/prompt_encoder.py:70-231,
// modeling/sam/transformer.py:51-294, modeling/sam/mask_decoder.py:116-316, modeling/sam2_base_official.py:338-494.
//
// Numerics: the reference keeps these heads in fp32 (the multimask arg-max runs over near-tied predicted IoUs).  The image
// side (4096 rows per object) goes through the bf16 tensor cores with BOTH operands split into two bf16 terms,
// x = hi + lo: A' = [hi | lo | hi], B' = [hi | hi | lo] over 3K columns gives hi*hi + lo*hi + hi*lo with fp32 accumulation,
// i.e. a relative error of ~2^-16 per product instead of bf16's 2^-8 (the kernels here emit A' directly).  Everything on the
// token side (<= 32 tokens per object) and every soft-max, LayerNorm and GELU is plain fp32.
#include <math.h>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

namespace {

__device__ __forceinline__ void split2(float x, bf16& hi, bf16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float gelu_exact(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

// ---------------------------------------------------------------------------------------------------------------------
// src = image_embedding + dense prompt embedding (mask_decoder.py:262-266), token-major: keys fp32 [n*4096][256] and the
// split operand ks [n*4096][768].  pix is addressed as pix[b*sn + c*sc + p*sp] (NCHW: sp == 1; the E3 output
// [4096][n][256]: sc == 1; sn == 0 repeats one image for every prompt); dense [n][256][4096] or nullptr (no_mask_embed [256]).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void heads_prep_src_kernel(const float* __restrict__ pix, long sn, long sc, long sp, const float* __restrict__ dense,
                                      const float* __restrict__ nomask, float* __restrict__ keys, bf16* __restrict__ ks) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, p0 = blockIdx.x * 32, c0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
  const bool chw = (sp == 1);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = c0 + ty + 8 * i, p = p0 + tx;
    float v = 0.f;
    if (chw) v = pix[b * sn + c * sc + p];
    if (dense) v += dense[((long)b * 256 + c) * 4096 + p];
    tile[ty + 8 * i][tx] = v;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = p0 + ty + 8 * i, c = c0 + tx;
    float v = tile[tx][ty + 8 * i];
    if (!chw) v = pix[b * sn + c * sc + p * sp] + v;
    if (!dense) v += nomask[c];
    const long row = (long)b * 4096 + p;
    keys[row * 256 + c] = v;
    bf16 hi, lo;
    split2(v, hi, lo);
    ks[row * 768 + c] = hi; ks[row * 768 + 256 + c] = lo; ks[row * 768 + 512 + c] = hi;
  }
}

// LayerNorm over 256 channels (one warp per row), fp32 out + split operand out (either may be nullptr)
__global__ void ln256_split_kernel(const float* __restrict__ x, const float* __restrict__ g, const float* __restrict__ be, float eps,
                                   float* __restrict__ y, bf16* __restrict__ ys, long rows) {
  const long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  const float4 a = *reinterpret_cast<const float4*>(x + row * 256 + lane * 4);
  const float4 c = *reinterpret_cast<const float4*>(x + row * 256 + 128 + lane * 4);
  float v[8] = {a.x, a.y, a.z, a.w, c.x, c.y, c.z, c.w};
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  const float mean = warp_sum(s) * (1.0f / 256.0f);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] -= mean; q += v[i] * v[i]; }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / 256.0f) + eps);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int col = h * 128 + lane * 4;
    const float4 gg = *reinterpret_cast<const float4*>(g + col), bb = *reinterpret_cast<const float4*>(be + col);
    float o[4] = {v[h * 4 + 0] * rstd * gg.x + bb.x, v[h * 4 + 1] * rstd * gg.y + bb.y, v[h * 4 + 2] * rstd * gg.z + bb.z,
                  v[h * 4 + 3] * rstd * gg.w + bb.w};
    if (y) *reinterpret_cast<float4*>(y + row * 256 + col) = make_float4(o[0], o[1], o[2], o[3]);
    if (ys) {
      __align__(8) bf16 hi[4], lo[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) split2(o[i], hi[i], lo[i]);
      *reinterpret_cast<uint2*>(ys + row * 768 + col) = *reinterpret_cast<const uint2*>(hi);
      *reinterpret_cast<uint2*>(ys + row * 768 + 256 + col) = *reinterpret_cast<const uint2*>(lo);
      *reinterpret_cast<uint2*>(ys + row * 768 + 512 + col) = *reinterpret_cast<const uint2*>(hi);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Tokens of every object (prompt_encoder.py:124-150 `_embed_points` with the padding point, mask_decoder.py:235-247):
// rows 0..5 = obj_score / iou / 4 mask tokens, rows 6..6+P-1 = the points, last row = the padding point (label -1).
// tok [n][T][256] (queries) and qpe (the same values: the reference's query_pe is the initial point embedding).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void heads_tokens_kernel(const float* __restrict__ coords, const int* __restrict__ labels, int P,
                                    const float* __restrict__ gauss /*[2][128]*/, const float* __restrict__ not_a_point,
                                    const float* __restrict__ pt_emb0, const float* __restrict__ pt_emb1,
                                    const float* __restrict__ pt_emb2, const float* __restrict__ pt_emb3,
                                    const float* __restrict__ obj_tok, const float* __restrict__ iou_tok,
                                    const float* __restrict__ mask_toks /*[4][256]*/, const float* __restrict__ sparse /*[n][S][256]*/,
                                    int S, float* __restrict__ tok, float* __restrict__ qpe) {
  // sparse != nullptr: the caller's own sparse prompt embeddings (MaskDecoder.forward's argument) follow the 6 output tokens
  const int b = blockIdx.x, c = threadIdx.x, T = sparse ? 6 + S : 6 + P + 1;
  float* t0 = tok + (long)b * T * 256;
  float* q0 = qpe + (long)b * T * 256;
  for (int r = 0; r < T; ++r) {
    float v;
    if (r == 0) v = obj_tok[c];
    else if (r == 1) v = iou_tok[c];
    else if (r < 6) v = mask_toks[(r - 2) * 256 + c];
    else if (sparse) v = sparse[((long)b * S + (r - 6)) * 256 + c];
    else {
      const int i = r - 6;
      const int lab = (i < P) ? labels[b * P + i] : -1;
      if (lab == -1) v = not_a_point[c];
      else {
        // pe_layer._pe_encoding (position_encoding.py:146-154): coords in [0,1] -> [-1,1] -> @ gaussian matrix -> 2 pi -> sin | cos
        const float x = 2.0f * ((coords[(b * P + i) * 2 + 0] + 0.5f) / 1024.0f) - 1.0f;
        const float y = 2.0f * ((coords[(b * P + i) * 2 + 1] + 0.5f) / 1024.0f) - 1.0f;
        const int j = c & 127;
        const float ang = 6.283185307179586f * (x * gauss[j] + y * gauss[128 + j]);
        v = (c < 128) ? sinf(ang) : cosf(ang);
        const float* e = lab == 0 ? pt_emb0 : lab == 1 ? pt_emb1 : lab == 2 ? pt_emb2 : pt_emb3;
        if (lab >= 0 && lab <= 3) v += e[c];
      }
    }
    t0[r * 256 + c] = v;
    q0[r * 256 + c] = v;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// PromptEncoder.mask_downscaling (prompt_encoder.py:97-109,152-155): the dense embedding of a mask prompt [n,1,256,256] ->
// conv 2x2 s2 (1 -> 4) + LayerNorm2d + GELU -> conv 2x2 s2 (4 -> 16) + LayerNorm2d + GELU -> conv 1x1 (16 -> 256) = [n,256,64,64].
// One thread per output position: its 4 x 4 input patch, 4 x 4 and 16 intermediate values in registers; the 1x1 weights
// [256][16] in shared memory.  fp32 throughout, exact (erf) GELU.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) mask_downscale_kernel(const float* __restrict__ mask, const float* __restrict__ w0,
                                                             const float* __restrict__ b0, const float* __restrict__ g1,
                                                             const float* __restrict__ be1, const float* __restrict__ w3,
                                                             const float* __restrict__ b3, const float* __restrict__ g4,
                                                             const float* __restrict__ be4, const float* __restrict__ w6,
                                                             const float* __restrict__ b6, float* __restrict__ out) {
  __shared__ float w6s[256 * 16];
  __shared__ float w3s[16 * 16];     // [c2][c][dy][dx]
  __shared__ float small[16 + 4 + 4 + 4 + 16 + 16 + 16];
  float* w0s = small;                // [4][2][2]
  float* b0s = w0s + 16; float* g1s = b0s + 4; float* be1s = g1s + 4;
  float* b3s = be1s + 4; float* g4s = b3s + 16; float* be4s = g4s + 16;
  const int tid = threadIdx.x, b = blockIdx.y;
  for (int i = tid; i < 4096; i += 256) w6s[i] = w6[i];
  w3s[tid] = w3[tid];
  if (tid < 16) { w0s[tid] = w0[tid]; b3s[tid] = b3[tid]; g4s[tid] = g4[tid]; be4s[tid] = be4[tid]; }
  if (tid < 4) { b0s[tid] = b0[tid]; g1s[tid] = g1[tid]; be1s[tid] = be1[tid]; }
  __syncthreads();
  const int pos = blockIdx.x * 256 + tid, y = pos >> 6, x = pos & 63;
  const float* src = mask + (long)b * 65536 + (4 * y) * 256 + 4 * x;
  float in[4][4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const float4 v = *reinterpret_cast<const float4*>(src + r * 256);
    in[r][0] = v.x; in[r][1] = v.y; in[r][2] = v.z; in[r][3] = v.w;
  }
  float a1[4][2][2];                 // [c][py][px] after LayerNorm2d + GELU
#pragma unroll
  for (int py = 0; py < 2; ++py)
#pragma unroll
    for (int px = 0; px < 2; ++px) {
      float v[4], mean = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        float acc = b0s[c];
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) acc = fmaf(w0s[c * 4 + dy * 2 + dx], in[2 * py + dy][2 * px + dx], acc);
        v[c] = acc; mean += acc;
      }
      mean *= 0.25f;
      float var = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) { v[c] -= mean; var = fmaf(v[c], v[c], var); }
      const float rstd = 1.0f / sqrtf(var * 0.25f + 1e-6f);
#pragma unroll
      for (int c = 0; c < 4; ++c) a1[c][py][px] = gelu_exact(g1s[c] * (v[c] * rstd) + be1s[c]);
    }
  float a2[16], mean = 0.f;
#pragma unroll
  for (int c2 = 0; c2 < 16; ++c2) {
    float acc = b3s[c2];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
      for (int dy = 0; dy < 2; ++dy)
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) acc = fmaf(w3s[c2 * 16 + c * 4 + dy * 2 + dx], a1[c][dy][dx], acc);
    a2[c2] = acc; mean += acc;
  }
  mean *= (1.0f / 16.0f);
  float var = 0.f;
#pragma unroll
  for (int c2 = 0; c2 < 16; ++c2) { a2[c2] -= mean; var = fmaf(a2[c2], a2[c2], var); }
  const float rstd = 1.0f / sqrtf(var * (1.0f / 16.0f) + 1e-6f);
#pragma unroll
  for (int c2 = 0; c2 < 16; ++c2) a2[c2] = gelu_exact(g4s[c2] * (a2[c2] * rstd) + be4s[c2]);
  float* dst = out + (long)b * 256 * 4096 + pos;
  for (int co = 0; co < 256; ++co) {
    const float4* wr = reinterpret_cast<const float4*>(w6s + co * 16);
    float acc = b6[co];
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      const float4 wv = wr[k4];
      acc = fmaf(wv.x, a2[4 * k4], acc); acc = fmaf(wv.y, a2[4 * k4 + 1], acc);
      acc = fmaf(wv.z, a2[4 * k4 + 2], acc); acc = fmaf(wv.w, a2[4 * k4 + 3], acc);
    }
    dst[(long)co * 4096] = acc;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Small fp32 linear layers on the token side, several independent problems ("jobs") per launch:
//   Y[r][j] = act( sum_k (X[r][k] + X2[r][k]) * W[j][k] + b[j] ),  r < rows, j < N, K % 128 == 0.
// A job may be one K slice of a larger problem (split K: the partial sums are added by tok_sum_res_ln_kernel).
// Tile: 16 rows x 64 columns per CTA, 128-wide K chunks through shared memory.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) tok_linear_kernel(const __grid_constant__ LinJobs jobs) {
  const LinJob& J = jobs.j[blockIdx.z];
  const int n0 = blockIdx.x * 64, r0 = blockIdx.y * 16;
  if (n0 >= J.N || r0 >= J.rows) return;
  // 128-wide K chunks; rows of 132 floats keep every float4 16-byte aligned and the LDS.128 of 8 consecutive rows on 32
  // distinct banks.  128 threads, each 2 rows x 4 columns (5.3 FMA per shared-memory load).  The next chunk is fetched into
  // registers while the current one is multiplied (these problems are tiny: what a launch costs is its chain of dependent
  // memory round trips, not its arithmetic).
  constexpr int KC = 128, LD = KC + 4;
  __shared__ __align__(16) float Xs[16 * LD];
  __shared__ __align__(16) float Ws[64 * LD];
  const int tid = threadIdx.x, tr = tid >> 4, tc = tid & 15;
  float4 xr[4], wr[16];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = tid + 128 * i, r = idx >> 5, k4 = idx & 31;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + r < J.rows) {
        v = *reinterpret_cast<const float4*>(J.X + (long)(r0 + r) * J.ldx + k0 + 4 * k4);
        if (J.X2) {
          const float4 u = *reinterpret_cast<const float4*>(J.X2 + (long)(r0 + r) * J.ldx2 + k0 + 4 * k4);
          v.x += u.x; v.y += u.y; v.z += u.z; v.w += u.w;
        }
      }
      xr[i] = v;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int idx = tid + 128 * i, j = idx >> 5, k4 = idx & 31;
      wr[i] = (n0 + j < J.N) ? *reinterpret_cast<const float4*>(J.W + (long)(n0 + j) * J.ldw + k0 + 4 * k4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
  fetch(0);
  float bias[4];                  // fetched with the first chunk: a load at the end would be one more exposed memory round trip
#pragma unroll
  for (int jj = 0; jj < 4; ++jj) bias[jj] = (J.b && n0 + tc + 16 * jj < J.N) ? J.b[n0 + tc + 16 * jj] : 0.f;
  for (int k0 = 0; k0 < J.K; k0 += KC) {
#pragma unroll
    for (int i = 0; i < 4; ++i) { const int idx = tid + 128 * i; *reinterpret_cast<float4*>(Xs + (idx >> 5) * LD + 4 * (idx & 31)) = xr[i]; }
#pragma unroll
    for (int i = 0; i < 16; ++i) { const int idx = tid + 128 * i; *reinterpret_cast<float4*>(Ws + (idx >> 5) * LD + 4 * (idx & 31)) = wr[i]; }
    __syncthreads();
    if (k0 + KC < J.K) fetch(k0 + KC);
#pragma unroll 4
    for (int k4 = 0; k4 < KC / 4; ++k4) {
      const float4 x0 = *reinterpret_cast<const float4*>(Xs + tr * LD + 4 * k4);
      const float4 x1 = *reinterpret_cast<const float4*>(Xs + (tr + 8) * LD + 4 * k4);
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const float4 w = *reinterpret_cast<const float4*>(Ws + (tc + 16 * jj) * LD + 4 * k4);
        acc[0][jj] = fmaf(x0.x, w.x, acc[0][jj]); acc[0][jj] = fmaf(x0.y, w.y, acc[0][jj]);
        acc[0][jj] = fmaf(x0.z, w.z, acc[0][jj]); acc[0][jj] = fmaf(x0.w, w.w, acc[0][jj]);
        acc[1][jj] = fmaf(x1.x, w.x, acc[1][jj]); acc[1][jj] = fmaf(x1.y, w.y, acc[1][jj]);
        acc[1][jj] = fmaf(x1.z, w.z, acc[1][jj]); acc[1][jj] = fmaf(x1.w, w.w, acc[1][jj]);
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = r0 + tr + 8 * rr;
    if (r >= J.rows) continue;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = n0 + tc + 16 * jj;
      if (j >= J.N) continue;
      float v = acc[rr][jj] + bias[jj];
      if (J.act == 1) v = fmaxf(v, 0.f);
      else if (J.act == 2) v = 1.0f / (1.0f + expf(-v));
      J.Y[(long)r * J.ldy + j] = v;
    }
  }
}

// Y[r][:] = LN( sum_s part[s][r][:] + bias + residual[r][:] )  over 256 columns (one warp per row; LN optional: g == nullptr)
__global__ void tok_sum_res_ln_kernel(const float* __restrict__ part, long part_stride, int S, const float* __restrict__ bias,
                                      const float* __restrict__ res, const float* __restrict__ g, const float* __restrict__ be,
                                      float eps, float* __restrict__ y, int rows) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int lane = threadIdx.x & 31;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = (i >> 2) * 128 + lane * 4 + (i & 3);
    float a = 0.f;
    for (int s = 0; s < S; ++s) a += part[s * part_stride + (long)row * 256 + c];
    a += bias[c];
    if (res) a += res[(long)row * 256 + c];
    v[i] = a;
  }
  if (g) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += v[i];
    const float mean = warp_sum(s) * (1.0f / 256.0f);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { v[i] -= mean; q += v[i] * v[i]; }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / 256.0f) + eps);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = (i >> 2) * 128 + lane * 4 + (i & 3);
      v[i] = v[i] * rstd * g[c] + be[c];
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) y[(long)row * 256 + (i >> 2) * 128 + lane * 4 + (i & 3)] = v[i];
}

// Self-attention over the T <= 32 tokens of one object (transformer.py:222-294 with 8 heads of 32): one CTA per object,
// one warp per head, lane = query.  q, k, v, o: [n*T][256].
__global__ void __launch_bounds__(256) tok_self_attn_kernel(const float* __restrict__ q, const float* __restrict__ k,
                                                            const float* __restrict__ v, int T, float* __restrict__ o) {
  __shared__ float sm[32 * 256];
  const int b = blockIdx.x, tid = threadIdx.x, h = tid >> 5, t = tid & 31;
  const long base = (long)b * T * 256;
  for (int i = tid; i < T * 256; i += 256) sm[i] = k[base + i];
  float qr[32], s[32];
  if (t < T) {
#pragma unroll
    for (int d = 0; d < 32; ++d) qr[d] = q[base + t * 256 + h * 32 + d] * 0.17677669529663687f;   // 1 / sqrt(32)
  } else {
#pragma unroll
    for (int d = 0; d < 32; ++d) qr[d] = 0.f;
  }
  __syncthreads();
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float a = -INFINITY;
    if (j < T) {
      a = 0.f;
#pragma unroll
      for (int d = 0; d < 32; ++d) a = fmaf(qr[d], sm[j * 256 + h * 32 + d], a);
    }
    s[j] = a;
    m = fmaxf(m, a);
  }
  float l = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) { s[j] = (j < T) ? expf(s[j] - m) : 0.f; l += s[j]; }
  __syncthreads();
  for (int i = tid; i < T * 256; i += 256) sm[i] = v[base + i];
  __syncthreads();
  float acc[32];
#pragma unroll
  for (int d = 0; d < 32; ++d) acc[d] = 0.f;
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    if (j < T) {
#pragma unroll
      for (int d = 0; d < 32; ++d) acc[d] = fmaf(s[j], sm[j * 256 + h * 32 + d], acc[d]);
    }
  }
  if (t < T) {
    const float inv = 1.0f / l;
#pragma unroll
    for (int d = 0; d < 32; ++d) o[base + t * 256 + h * 32 + d] = acc[d] * inv;
  }
}

// Token -> image cross-attention (8 heads of 16 over the 4096 image keys).  One CTA per (tile of 32 keys, object); one thread
// per (token, head) pair runs its own soft-max over the tile (scores in registers, no cross-thread reduction at all) and
// leaves (max, sum, 16 un-normalised outputs); t2i_combine_kernel merges the 128 tiles.  kv: fp32 rows of `ld` floats with k
// at column 0 and v at column 128; q [n*T][128]; part [n][128][T][8][18].
constexpr int kT2iKeys = 32, kT2iTiles = 4096 / kT2iKeys;
__global__ void __launch_bounds__(256) t2i_partial_kernel(const float* __restrict__ kv, int ld, const float* __restrict__ q, int T,
                                                          float* __restrict__ part) {
  // [key][head][17]: the 8 heads a warp reads at once fall on 8 different banks
  __shared__ float ks[kT2iKeys * 8 * 17];
  __shared__ float vs[kT2iKeys * 8 * 17];
  const int tile = blockIdx.x, b = blockIdx.y, tid = threadIdx.x;
  const float* base = kv + ((long)b * 4096 + tile * kT2iKeys) * ld;
  for (int i = tid; i < kT2iKeys * 64; i += blockDim.x) {          // 64 float4 per key: 32 of k, 32 of v
    const int j = i >> 6, c4 = i & 63;
    const float4 a = *reinterpret_cast<const float4*>(base + (long)j * ld + 4 * c4);
    float* dst = (c4 < 32 ? ks : vs) + (j * 8 + ((c4 & 31) >> 2)) * 17 + 4 * (c4 & 3);
    dst[0] = a.x; dst[1] = a.y; dst[2] = a.z; dst[3] = a.w;
  }
  const int h = tid & 7, t = tid >> 3;
  float qr[16];
  if (t < T) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(q + ((long)b * T + t) * 128 + h * 16 + 4 * i);
      qr[4 * i] = a.x * 0.25f; qr[4 * i + 1] = a.y * 0.25f; qr[4 * i + 2] = a.z * 0.25f; qr[4 * i + 3] = a.w * 0.25f;
    }
  }
  __syncthreads();
  if (t >= T) return;
  float sc[kT2iKeys], m = -INFINITY;
#pragma unroll
  for (int j = 0; j < kT2iKeys; ++j) {
    const float* kk = ks + (j * 8 + h) * 17;
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int d = 0; d < 16; d += 2) { a0 = fmaf(qr[d], kk[d], a0); a1 = fmaf(qr[d + 1], kk[d + 1], a1); }
    sc[j] = a0 + a1;
    m = fmaxf(m, sc[j]);
  }
  float l = 0.f, acc[16];
#pragma unroll
  for (int d = 0; d < 16; ++d) acc[d] = 0.f;
#pragma unroll
  for (int j = 0; j < kT2iKeys; ++j) {
    const float* vv = vs + (j * 8 + h) * 17;
    const float e = expf(sc[j] - m);
    l += e;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[d] = fmaf(e, vv[d], acc[d]);
  }
  float* dst = part + ((((long)b * kT2iTiles + tile) * T + t) * 8 + h) * 18;
  dst[0] = m; dst[1] = l;
#pragma unroll
  for (int d = 0; d < 16; ++d) dst[2 + d] = acc[d];
}

// merge the key tiles: out [n*T][128], column = head * 16 + d.  One warp per (object, token, head): every lane folds 4 of the
// 128 tiles, then the lanes merge their (max, sum, outputs) pairwise.
__global__ void __launch_bounds__(256) t2i_combine_kernel(const float* __restrict__ part, int T, float* __restrict__ out, int total_warps) {
  const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (wid >= total_warps) return;
  const int h = wid & 7, bt = wid >> 3, b = bt / T, t = bt % T;
  const float* p = part + ((((long)b * kT2iTiles) * T + t) * 8 + h) * 18;
  const long ss = (long)T * 8 * 18;
  float M = -INFINITY, L = 0.f, O[16];
#pragma unroll
  for (int d = 0; d < 16; ++d) O[d] = 0.f;
#pragma unroll
  for (int i = 0; i < kT2iTiles / 32; ++i) {
    const float* q = p + (lane + 32 * i) * ss;
    const float m = q[0], Mn = fmaxf(M, m), e0 = expf(M - Mn), e1 = expf(m - Mn);
    L = L * e0 + q[1] * e1;
#pragma unroll
    for (int d = 0; d < 16; ++d) O[d] = O[d] * e0 + q[2 + d] * e1;
    M = Mn;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m = __shfl_xor_sync(0xffffffffu, M, o), l = __shfl_xor_sync(0xffffffffu, L, o);
    const float Mn = fmaxf(M, m), e0 = expf(M - Mn), e1 = expf(m - Mn);
    L = L * e0 + l * e1;
#pragma unroll
    for (int d = 0; d < 16; ++d) O[d] = O[d] * e0 + __shfl_xor_sync(0xffffffffu, O[d], o) * e1;
    M = Mn;
  }
  if (lane < 16) {
    float v = O[0];
#pragma unroll
    for (int d = 1; d < 16; ++d) v = (lane == d) ? O[d] : v;
    out[(long)bt * 128 + h * 16 + lane] = v / L;
  }
}

// Image -> token cross-attention: every pixel attends to the T tokens (8 heads of 16).  One thread per (pixel, head);
// q = columns [q_col, q_col + 128) of the fp32 rows `kvq` (ld floats); tk, tv [n*T][128] (token keys / values).
// Output: the split operand os [n*4096][384] = [hi | lo | hi] of the 128 attention outputs (A of the out_proj GEMM).
__global__ void __launch_bounds__(256) i2t_attn_kernel(const float* __restrict__ kvq, int ld, int q_col, const float* __restrict__ tk,
                                                       const float* __restrict__ tv, int T, bf16* __restrict__ os) {
  // token keys / values as [t][head][17]: the 8 heads a warp reads at once fall on 8 different banks
  __shared__ float ks[32 * 8 * 17];
  __shared__ float vs[32 * 8 * 17];
  const int b = blockIdx.y, tid = threadIdx.x;
  for (int i = tid; i < T * 128; i += 256) {
    const int j = (i >> 4) * 17 + (i & 15);
    ks[j] = tk[(long)b * T * 128 + i];
    vs[j] = tv[(long)b * T * 128 + i];
  }
  __syncthreads();
  const int h = tid & 7, p = blockIdx.x * 32 + (tid >> 3);
  const long row = (long)b * 4096 + p;
  const float* qp = kvq + row * ld + q_col + h * 16;
  float qr[16];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 a = *reinterpret_cast<const float4*>(qp + 4 * i);
    qr[4 * i] = a.x * 0.25f; qr[4 * i + 1] = a.y * 0.25f; qr[4 * i + 2] = a.z * 0.25f; qr[4 * i + 3] = a.w * 0.25f;
  }
  float m = -INFINITY, l = 0.f, acc[16];
#pragma unroll
  for (int d = 0; d < 16; ++d) acc[d] = 0.f;
  // two passes over the (few) tokens: exact max first, then the weighted sum
  for (int t = 0; t < T; ++t) {
    const float* kk = ks + (t * 8 + h) * 17;
    float a = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) a = fmaf(qr[d], kk[d], a);
    m = fmaxf(m, a);
  }
  for (int t = 0; t < T; ++t) {
    const float* kk = ks + (t * 8 + h) * 17;
    const float* vv = vs + (t * 8 + h) * 17;
    float a = 0.f;
#pragma unroll
    for (int d = 0; d < 16; ++d) a = fmaf(qr[d], kk[d], a);
    const float e = expf(a - m);
    l += e;
#pragma unroll
    for (int d = 0; d < 16; ++d) acc[d] = fmaf(e, vv[d], acc[d]);
  }
  const float inv = 1.0f / l;
  __align__(16) bf16 hi[16], lo[16];
#pragma unroll
  for (int d = 0; d < 16; ++d) split2(acc[d] * inv, hi[d], lo[d]);
  bf16* o = os + row * 384 + h * 16;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    reinterpret_cast<uint4*>(o)[i] = reinterpret_cast<const uint4*>(hi)[i];
    reinterpret_cast<uint4*>(o + 128)[i] = reinterpret_cast<const uint4*>(lo)[i];
    reinterpret_cast<uint4*>(o + 256)[i] = reinterpret_cast<const uint4*>(hi)[i];
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// Up-scaling tail (mask_decoder.py:283-300): u1 [n*4096][256 = (dy,dx,c)] is ConvTranspose2d(256 -> 64, k2 s2) of the
// final keys (a GEMM); per 128x128 position: + feat_s1, LayerNorm2d(64), GELU, ConvTranspose2d(64 -> 32, k2 s2), + feat_s0, GELU,
// then the 4 hyper-network dot products -> masks [n][4][256][256].  The 32-channel 256x256 map never exists.
// One CTA per 256 positions of the 128x128 grid: first a thread per position (up to the GELU), then a thread per (tap, 4
// positions) for the second transposed convolution; w2 [64][4 = (ey,ex)][32] and the biases live in shared memory.
// stab [n][2]: number of pixels of mask 0 above +delta / above -delta (mask_decoder.py:320-335), only if stab != nullptr.
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) heads_upscale_kernel(const float* __restrict__ u1, const float* __restrict__ b1 /*[64]*/,
                                                            const float* __restrict__ s1, long s1_bs, const float* __restrict__ ln_g,
                                                            const float* __restrict__ ln_b, const float* __restrict__ w2,
                                                            const float* __restrict__ b2 /*[32]*/, const float* __restrict__ s0,
                                                            long s0_bs, const float* __restrict__ hyper /*[n][4][32]*/,
                                                            float* __restrict__ masks, int* __restrict__ stab, float delta) {
  extern __shared__ float smem[];
  float* w2s = smem;              // [64][4][32]
  float* hs = w2s + 64 * 128;     // [4][32]
  float* b2s = hs + 128;          // [32]
  float* xs = b2s + 32;           // [64][256]: the 64-channel vector of each of the CTA's 256 positions after LayerNorm2d + GELU
  const int b = blockIdx.y, tid = threadIdx.x;
  for (int i = tid; i < 64 * 128; i += 256) w2s[i] = w2[i];
  if (tid < 128) hs[tid] = hyper[(long)b * 128 + tid];
  if (tid < 32) b2s[tid] = b2[tid];
  __syncthreads();
  const int pos = blockIdx.x * 256 + tid;          // 0 .. 16383 on the 128 x 128 grid
  const int y1 = pos >> 7, x1 = pos & 127;
  const long pix = (long)b * 4096 + (y1 >> 1) * 64 + (x1 >> 1);
  const float* up = u1 + pix * 256 + ((y1 & 1) * 2 + (x1 & 1)) * 64;
  const float* f1 = s1 + b * s1_bs + (long)y1 * 128 + x1;
  float x[64];
  float sum = 0.f;
#pragma unroll
  for (int c4 = 0; c4 < 16; ++c4) {
    const float4 a = *reinterpret_cast<const float4*>(up + 4 * c4);
    x[4 * c4 + 0] = a.x + b1[4 * c4 + 0] + f1[(long)(4 * c4 + 0) * 16384];
    x[4 * c4 + 1] = a.y + b1[4 * c4 + 1] + f1[(long)(4 * c4 + 1) * 16384];
    x[4 * c4 + 2] = a.z + b1[4 * c4 + 2] + f1[(long)(4 * c4 + 2) * 16384];
    x[4 * c4 + 3] = a.w + b1[4 * c4 + 3] + f1[(long)(4 * c4 + 3) * 16384];
  }
#pragma unroll
  for (int c = 0; c < 64; ++c) sum += x[c];
  const float mean = sum * (1.0f / 64.0f);
  float var = 0.f;
#pragma unroll
  for (int c = 0; c < 64; ++c) { x[c] -= mean; var = fmaf(x[c], x[c], var); }
  const float rstd = 1.0f / sqrtf(var * (1.0f / 64.0f) + 1e-6f);
  // GELU: gelu_fast (common.cuh) = x * Phi(x) with Phi(-|x|) as 2^poly6, max |error| 2e-7 -- fp32-grade at a third of erff's cost
#pragma unroll
  for (int c = 0; c < 64; ++c) xs[c * 256 + tid] = gelu_fast(ln_g[c] * (x[c] * rstd) + ln_b[c]);
  __syncthreads();
  // Second transposed convolution: every thread takes ONE of the four output taps q (uniform per warp) for FOUR positions, so
  // that a weight read from shared memory (a broadcast LDS.128 costs four wavefronts) feeds 16 FMAs instead of 4: with one
  // position per thread the kernel sat at the shared-memory pipe (ncu: 66 % l1tex, short-scoreboard stalls), 0.26 ms at 16 objects.
  int cnt_hi = 0, cnt_lo = 0;
  const int q = tid >> 6, pg = tid & 63;
  uint64_t acc2[4][16];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int o = 0; o < 16; ++o) acc2[i][o] = f2_pack(b2s[2 * o], b2s[2 * o + 1]);
#pragma unroll 4
  for (int c = 0; c < 64; ++c) {
    uint64_t xx[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float xc = xs[c * 256 + pg + 64 * i]; xx[i] = f2_pack(xc, xc); }
    const ulonglong2* wr = reinterpret_cast<const ulonglong2*>(w2s + (c * 4 + q) * 32);
#pragma unroll
    for (int o4 = 0; o4 < 8; ++o4) {
      const ulonglong2 wv = wr[o4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc2[i][2 * o4] = f2_fma(xx[i], wv.x, acc2[i][2 * o4]);
        acc2[i][2 * o4 + 1] = f2_fma(xx[i], wv.y, acc2[i][2 * o4 + 1]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p2 = blockIdx.x * 256 + pg + 64 * i, yy = p2 >> 7, xq = p2 & 127;
    const int y2 = 2 * yy + (q >> 1), x2 = 2 * xq + (q & 1);
    const float* f0 = s0 + b * s0_bs + (long)y2 * 256 + x2;
    float m[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int o = 0; o < 16; ++o) {
      float a0, a1;
      f2_unpack(acc2[i][o], a0, a1);
      const float g0 = gelu_fast(a0 + f0[(long)(2 * o) * 65536]), g1 = gelu_fast(a1 + f0[(long)(2 * o + 1) * 65536]);
#pragma unroll
      for (int k = 0; k < 4; ++k) m[k] = fmaf(hs[k * 32 + 2 * o + 1], g1, fmaf(hs[k * 32 + 2 * o], g0, m[k]));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) masks[(((long)b * 4 + k) * 256 + y2) * 256 + x2] = m[k];
    cnt_hi += (m[0] > delta);
    cnt_lo += (m[0] > -delta);
  }
  if (stab) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      cnt_hi += __shfl_xor_sync(0xffffffffu, cnt_hi, o);
      cnt_lo += __shfl_xor_sync(0xffffffffu, cnt_lo, o);
    }
    if ((tid & 31) == 0) { atomicAdd(stab + b * 2, cnt_hi); atomicAdd(stab + b * 2 + 1, cnt_lo); }
  }
}

// Per object: which of the 4 masks is returned / tracked, the IoU outputs and the object pointer
// (mask_decoder.py:192-224,337-382; sam2_base_official.py:435-494).  One CTA of 256 threads per object.
// sel [n][4]: sel[0] = index of the best (tracked) mask among the 4, sel[1] = is_obj.
__global__ void heads_select_kernel(const float* __restrict__ iou /*[n][4]*/, const float* __restrict__ obj /*[n]*/,
                                    const float* __restrict__ ptr4 /*[n][4][256]*/, const int* __restrict__ stab,
                                    const float* __restrict__ no_obj_ptr, int multimask, int dynamic, float thresh,
                                    float* __restrict__ iou_out, float* __restrict__ ptr_out, float* __restrict__ obj_out,
                                    int* __restrict__ sel, int decoder_only) {
  const int b = blockIdx.x, c = threadIdx.x;
  const float i0 = iou[b * 4], i1 = iou[b * 4 + 1], i2 = iou[b * 4 + 2], i3 = iou[b * 4 + 3];
  int best = 1;                                   // torch.argmax: the first maximum
  float bi = i1;
  if (i2 > bi) { bi = i2; best = 2; }
  if (i3 > bi) { bi = i3; best = 3; }
  const bool is_obj = obj[b] > 0.f;
  int mask_idx, tok_idx;
  if (multimask) {
    mask_idx = best; tok_idx = best;
    if (c < 3) iou_out[b * 3 + c] = (c == 0) ? i1 : (c == 1) ? i2 : i3;
  } else {
    mask_idx = 0; tok_idx = 0;
    float io = i0;
    if (dynamic) {
      const float ai = (float)stab[b * 2], au = (float)stab[b * 2 + 1];
      const float score = au > 0.f ? ai / au : 1.0f;
      if (!(score >= thresh)) { mask_idx = best; io = bi; }
    }
    if (c == 0) iou_out[b] = io;
  }
  const float lam = is_obj ? 1.0f : 0.0f;
  ptr_out[(long)b * 256 + c] = lam * ptr4[((long)b * 4 + tok_idx) * 256 + c] + (1.0f - lam) * no_obj_ptr[c];
  // decoder_only: the plain SamMaskDecoder.forward outputs, without _forward_sam_heads' NO_OBJ_SCORE fill
  if (c == 0) { sel[b * 4] = mask_idx; sel[b * 4 + 1] = (is_obj || decoder_only) ? 1 : 0; obj_out[b] = obj[b]; }
}

// low_multi [n][M][256][256] = is_obj ? masks[b][k] : NO_OBJ_SCORE  (k = 1..3 for multimask, the selected mask otherwise);
// low_best [n][1][256][256] = the selected mask.  4 pixels per thread.
__global__ void heads_write_masks_kernel(const float* __restrict__ masks, const int* __restrict__ sel, int multimask,
                                         float no_obj_score, float* __restrict__ low_multi, float* __restrict__ low_best) {
  const int b = blockIdx.y;
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int best = sel[b * 4];
  const bool is_obj = sel[b * 4 + 1] != 0;
  const float4 fill = make_float4(no_obj_score, no_obj_score, no_obj_score, no_obj_score);
  const float* mb = masks + (long)b * 4 * 65536;
  if (multimask) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float4 v = is_obj ? *reinterpret_cast<const float4*>(mb + (long)(k + 1) * 65536 + i) : fill;
      if (low_multi) *reinterpret_cast<float4*>(low_multi + ((long)b * 3 + k) * 65536 + i) = v;
      if (low_best && k + 1 == best) *reinterpret_cast<float4*>(low_best + (long)b * 65536 + i) = v;
    }
  } else {
    const float4 v = is_obj ? *reinterpret_cast<const float4*>(mb + (long)best * 65536 + i) : fill;
    if (low_multi) *reinterpret_cast<float4*>(low_multi + (long)b * 65536 + i) = v;
    if (low_best) *reinterpret_cast<float4*>(low_best + (long)b * 65536 + i) = v;
  }
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------------------------------------------------
int heads_prep_src(const float* pix, long sn, long sc, long sp, const float* dense, const float* nomask, float* keys, bf16* ks,
                   int n, cudaStream_t st) {
  if (sp != 1 && sc != 1) { set_error("sam heads: image embedding must be contiguous along pixels or along channels"); return 1; }
  ProfRange pr(st, "heads_prep_src", 0, (double)n * 4096 * 256 * (4 + 4 + 6));
  heads_prep_src_kernel<<<dim3(128, 8, n), dim3(32, 8), 0, st>>>(pix, sn, sc, sp, dense, nomask, keys, ks);
  SAM2B_CHECK_LAUNCH("heads_prep_src");
  return 0;
}

int mask_downscale(const float* mask, const float* const w[11], float* out, int n, cudaStream_t st) {
  ProfRange pr(st, "heads_mask_downscale", 2.0 * n * 4096 * (16 * 4 + 256 + 4096), (double)n * (65536 + 256 * 4096) * 4);
  mask_downscale_kernel<<<dim3(16, n), 256, 0, st>>>(mask, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], w[8], w[9], out);
  SAM2B_CHECK_LAUNCH("mask_downscale");
  return 0;
}

int ln256_split(const float* x, const float* g, const float* b, float eps, float* y, bf16* ys, long rows, cudaStream_t st) {
  ProfRange pr(st, "heads_ln256_split", 0, (double)rows * 256 * (4 + 4 + 6));
  ln256_split_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(x, g, b, eps, y, ys, rows);
  SAM2B_CHECK_LAUNCH("ln256_split");
  return 0;
}

int heads_tokens(const float* coords, const int* labels, int P, int n, const float* gauss, const float* not_a_point,
                 const float* const pt_emb[4], const float* obj_tok, const float* iou_tok, const float* mask_toks,
                 const float* sparse, int S, float* tok, float* qpe, cudaStream_t st) {
  ProfRange pr(st, "heads_tokens", 0, 0);
  heads_tokens_kernel<<<n, 256, 0, st>>>(coords, labels, P, gauss, not_a_point, pt_emb[0], pt_emb[1], pt_emb[2], pt_emb[3], obj_tok,
                                         iou_tok, mask_toks, sparse, S, tok, qpe);
  SAM2B_CHECK_LAUNCH("heads_tokens");
  return 0;
}

int tok_linear(const LinJobs& jobs, int njobs, cudaStream_t st) {
  if (njobs <= 0 || njobs > LinJobs::kMax) { set_error("tok_linear: %d jobs (1..%d)", njobs, LinJobs::kMax); return 1; }
  int max_n = 0, max_r = 0;
  double flops = 0;
  for (int i = 0; i < njobs; ++i) {
    const LinJob& j = jobs.j[i];
    if (j.K % 128 || (j.ldx % 4) || (j.ldw % 4) || (j.X2 && j.ldx2 % 4) || j.rows <= 0 || j.N <= 0) {
      set_error("tok_linear: job %d rows=%d N=%d K=%d (K %% 128 == 0, leading dimensions %% 4 == 0)", i, j.rows, j.N, j.K);
      return 1;
    }
    max_n = j.N > max_n ? j.N : max_n;
    max_r = j.rows > max_r ? j.rows : max_r;
    flops += 2.0 * j.rows * j.N * j.K;
  }
  ProfRange pr(st, "heads_tok_linear", flops, 0);
  tok_linear_kernel<<<dim3((max_n + 63) / 64, (max_r + 15) / 16, njobs), 128, 0, st>>>(jobs);
  SAM2B_CHECK_LAUNCH("tok_linear");
  return 0;
}

int tok_sum_res_ln(const float* part, long part_stride, int S, const float* bias, const float* res, const float* g,
                   const float* be, float eps, float* y, int rows, cudaStream_t st) {
  ProfRange pr(st, "heads_tok_sum_res_ln", 0, 0);
  tok_sum_res_ln_kernel<<<(rows + 7) / 8, 256, 0, st>>>(part, part_stride, S, bias, res, g, be, eps, y, rows);
  SAM2B_CHECK_LAUNCH("tok_sum_res_ln");
  return 0;
}

int tok_self_attn(const float* q, const float* k, const float* v, int n, int T, float* o, cudaStream_t st) {
  ProfRange pr(st, "heads_tok_self_attn", 0, 0);
  tok_self_attn_kernel<<<n, 256, 0, st>>>(q, k, v, T, o);
  SAM2B_CHECK_LAUNCH("tok_self_attn");
  return 0;
}

int t2i_attention(const float* kv, int ld, const float* q, int n, int T, float* part, float* out, cudaStream_t st) {
  {
    ProfRange pr(st, "heads_t2i_partial", 4.0 * n * T * 4096 * 128, (double)n * 4096 * 256 * 4);
    t2i_partial_kernel<<<dim3(kT2iTiles, n), (8 * T + 31) / 32 * 32, 0, st>>>(kv, ld, q, T, part);
    SAM2B_CHECK_LAUNCH("t2i_partial");
  }
  ProfRange pr(st, "heads_t2i_combine", 0, 0);
  const int total_warps = n * T * 8;
  t2i_combine_kernel<<<(total_warps + 7) / 8, 256, 0, st>>>(part, T, out, total_warps);
  SAM2B_CHECK_LAUNCH("t2i_combine");
  return 0;
}

int i2t_attention(const float* kvq, int ld, int q_col, const float* tk, const float* tv, int n, int T, bf16* os, cudaStream_t st) {
  ProfRange pr(st, "heads_i2t_attn", 4.0 * n * T * 4096 * 128, (double)n * 4096 * (128 * 4 + 384 * 2));
  i2t_attn_kernel<<<dim3(128, n), 256, 0, st>>>(kvq, ld, q_col, tk, tv, T, os);
  SAM2B_CHECK_LAUNCH("i2t_attn");
  return 0;
}

int heads_upscale(const float* u1, const float* b1, const float* s1, long s1_bs, const float* ln_g, const float* ln_b,
                  const float* w2, const float* b2, const float* s0, long s0_bs, const float* hyper, float* masks, int* stab,
                  float delta, int n, cudaStream_t st) {
  static PerDeviceFlag attr;
  const int smem = (64 * 128 + 128 + 32 + 64 * 256) * 4;
  if (!attr.is_set()) {
    if (cudaFuncSetAttribute(heads_upscale_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) {
      set_error("heads_upscale: cannot set the shared-memory size");
      return 1;
    }
    attr.set();
  }
  ProfRange pr(st, "heads_upscale", 2.0 * n * 16384 * (64.0 * 128 + 4 * 32), (double)n * (4096.0 * 256 + 4 * 65536) * 4);
  heads_upscale_kernel<<<dim3(64, n), 256, smem, st>>>(u1, b1, s1, s1_bs, ln_g, ln_b, w2, b2, s0, s0_bs, hyper, masks, stab, delta);
  SAM2B_CHECK_LAUNCH("heads_upscale");
  return 0;
}

int heads_select(const float* iou, const float* obj, const float* ptr4, const int* stab, const float* no_obj_ptr, int multimask,
                 int dynamic, float thresh, float* iou_out, float* ptr_out, float* obj_out, int* sel, int decoder_only, int n,
                 cudaStream_t st) {
  ProfRange pr(st, "heads_select", 0, 0);
  heads_select_kernel<<<n, 256, 0, st>>>(iou, obj, ptr4, stab, no_obj_ptr, multimask, dynamic, thresh, iou_out, ptr_out, obj_out, sel,
                                         decoder_only);
  SAM2B_CHECK_LAUNCH("heads_select");
  return 0;
}

int heads_write_masks(const float* masks, const int* sel, int multimask, float no_obj_score, float* low_multi, float* low_best,
                      int n, cudaStream_t st) {
  ProfRange pr(st, "heads_write_masks", 0, (double)n * 65536 * 4 * 5);
  heads_write_masks_kernel<<<dim3(64, n), 256, 0, st>>>(masks, sel, multimask, no_obj_score, low_multi, low_best);
  SAM2B_CHECK_LAUNCH("heads_write_masks");
  return 0;
}

}  // namespace sam2b
