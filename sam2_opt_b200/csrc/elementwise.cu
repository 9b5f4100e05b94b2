// Memory-bound kernels of the path (SURVEY.md §8a' K5/K6/K7/K8/K11): LayerNorm rows with fp32 statistics
// and bf16 output (GEMM operand), casts, 2x2 max-pool of the skip path, patch-embed im2col, FPN top-down add,
// row softmax and RoPE.  All are coalesced / 16-byte vectorised; reductions are warp shuffles.
#include <cstdio>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

// ------------------------------------------------------------------------------------------
// LayerNorm over the last dim (reference: nn.LayerNorm in hieradet.py:104,122 eps 1e-6;
// memory_attention.py:44-46 eps 1e-5).  One warp per row, row kept in registers, two-pass variance.
// ------------------------------------------------------------------------------------------
template <int MAXV>  // max float4 per lane
__global__ void __launch_bounds__(256) layernorm_rows_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps,
                                                             bf16* __restrict__ y16, float* __restrict__ y32, long rows, int C) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nv = C >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + row * C);
  float4 v[MAXV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nv) {
      v[i] = xr[idx];
      s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    } else {
      v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
  const float mean = warp_sum(s) / (float)C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nv) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int idx = lane + i * 32;
    if (idx < nv) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + idx);
      const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + idx);
      float4 o;
      o.x = (v[i].x - mean) * rstd * g.x + b.x;
      o.y = (v[i].y - mean) * rstd * g.y + b.y;
      o.z = (v[i].z - mean) * rstd * g.z + b.z;
      o.w = (v[i].w - mean) * rstd * g.w + b.w;
      if (y16) {
        uint2 pk;
        pk.x = pack_bf16x2(o.x, o.y);
        pk.y = pack_bf16x2(o.z, o.w);
        reinterpret_cast<uint2*>(y16 + row * C)[idx] = pk;
      }
      if (y32) reinterpret_cast<float4*>(y32 + row * C)[idx] = o;
    }
  }
}

int layernorm_rows(const float* x, const float* gamma, const float* beta, float eps, bf16* y16, float* y32, long rows,
                   int C, cudaStream_t stream) {
  if (C % 4 || C > 1152) { set_error("layernorm_rows: unsupported C=%d", C); return 1; }
  const int nv = C / 4;
  const int wpb = 8;
  const unsigned grid = (unsigned)((rows + wpb - 1) / wpb);
  char nm[64];
  if (g_prof_on) snprintf(nm, sizeof(nm), "layernorm c%d", C);
  ProfRange pr(stream, nm, 0, (double)rows * C * (4.0 + (y16 ? 2.0 : 0.0) + (y32 ? 4.0 : 0.0)));
  if (nv <= 64) layernorm_rows_kernel<2><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y16, y32, rows, C);
  else if (nv <= 96) layernorm_rows_kernel<3><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y16, y32, rows, C);
  else if (nv <= 160) layernorm_rows_kernel<5><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y16, y32, rows, C);
  else layernorm_rows_kernel<9><<<grid, 256, 0, stream>>>(x, gamma, beta, eps, y16, y32, rows, C);
  SAM2B_CHECK_LAUNCH("layernorm_rows");
  return 0;
}

// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) cast_f32_bf16_kernel(const float4* __restrict__ x, uint2* __restrict__ y, long n4) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    uint2 pk;
    pk.x = pack_bf16x2(v.x, v.y);
    pk.y = pack_bf16x2(v.z, v.w);
    y[i] = pk;
  }
}
int cast_f32_bf16(const float* x, bf16* y, long n, cudaStream_t stream) {
  if (n % 4) { set_error("cast_f32_bf16: n=%ld not a multiple of 4", n); return 1; }
  const long n4 = n / 4;
  long blocks = (n4 + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "cast_bf16", 0, (double)n * 6.0);
  cast_f32_bf16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<uint2*>(y), n4);
  SAM2B_CHECK_LAUNCH("cast_f32_bf16");
  return 0;
}

// ------------------------------------------------------------------------------------------
// 2x2 max-pool, NHWC fp32 (reference: do_pool / MaxPool2d(2,2) on the skip path, hieradet.py:25-36,140-142)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) maxpool2x2_nhwc_kernel(const float4* __restrict__ x, float4* __restrict__ y, int B, int H,
                                                              int W, int C4) {
  const int Ho = H >> 1, Wo = W >> 1;
  const long total = (long)B * Ho * Wo * C4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4);
    long t = i / C4;
    const int ox = (int)(t % Wo); t /= Wo;
    const int oy = (int)(t % Ho);
    const int b = (int)(t / Ho);
    const long base = (((long)b * H + 2 * oy) * W + 2 * ox) * C4 + c;
    const float4 a = x[base], bq = x[base + C4], cq = x[base + (long)W * C4], d = x[base + (long)W * C4 + C4];
    float4 o;
    o.x = fmaxf(fmaxf(a.x, bq.x), fmaxf(cq.x, d.x));
    o.y = fmaxf(fmaxf(a.y, bq.y), fmaxf(cq.y, d.y));
    o.z = fmaxf(fmaxf(a.z, bq.z), fmaxf(cq.z, d.z));
    o.w = fmaxf(fmaxf(a.w, bq.w), fmaxf(cq.w, d.w));
    y[i] = o;
  }
}
int maxpool2x2_nhwc(const float* x, float* y, int B, int H, int W, int C, cudaStream_t stream) {
  if (C % 4 || H % 2 || W % 2) { set_error("maxpool2x2_nhwc: bad shape"); return 1; }
  const long total = (long)B * (H / 2) * (W / 2) * (C / 4);
  long blocks = (total + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "maxpool2x2", 0, (double)B * H * W * C * 5.0);
  maxpool2x2_nhwc_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), B, H, W, C / 4);
  SAM2B_CHECK_LAUNCH("maxpool2x2_nhwc");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Patch-embed im2col (reference: PatchEmbed Conv2d(3,144,k7,s4,p3), backbones/utils.py:63-93;
// optional (x-mean)/std of SAM2Transforms, utils/transforms.py:27-31) -> bf16 rows [B*256*256][160] (K 147 padded to 160).
// ------------------------------------------------------------------------------------------
// One CTA per 64 consecutive output pixels of one output row.  Work item = (pixel, channel, ky): the 7 taps of that row
// of the 7x7 window are x = 4 ox - 3 .. 4 ox + 3, i.e. lanes 1-3 of the aligned float4 at 4 ox - 4 and the float4 at 4 ox:
// two coalesced 16 B loads per item (consecutive pixels -> consecutive float4s).  The bf16 row tile [64][160] is staged in
// shared memory (row pitch 81 words: conflict-free 2-byte stores) and written out with coalesced 4-byte stores.
static constexpr int PI_PIX = 64;
static constexpr int PI_PITCH = 162;      // bf16 elements per staged row
__global__ void __launch_bounds__(256) patch_im2col_kernel(const float* __restrict__ img, uint32_t* __restrict__ cols, int B, int S,
                                                           int normalise) {
  __shared__ __align__(16) bf16 tile[PI_PIX * PI_PITCH];
  const int So = S >> 2;
  const int tiles_x = So / PI_PIX;
  const float mean[3] = {0.485f, 0.456f, 0.406f};
  const float istd[3] = {1.f / 0.229f, 1.f / 0.224f, 1.f / 0.225f};
  const long n_tiles = (long)B * So * tiles_x;
  for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const int tx = (int)(t % tiles_x);
    const long r = t / tiles_x;
    const int oy = (int)(r % So), b = (int)(r / So);
    const int ox0 = tx * PI_PIX;
    for (int i = threadIdx.x; i < 21 * PI_PIX; i += 256) {
      const int px = i & (PI_PIX - 1), q = i >> 6;          // q = c * 7 + ky
      const int c = q / 7, ky = q - c * 7;
      const int y = oy * 4 - 3 + ky, ox = ox0 + px;
      float4 lo = make_float4(0.f, 0.f, 0.f, 0.f), hi = lo;
      bool ok_lo = false, ok_hi = false;
      if (y >= 0 && y < S) {
        const float* row = img + (((long)b * 3 + c) * S + y) * S + 4 * ox;
        ok_hi = true;
        hi = __ldg(reinterpret_cast<const float4*>(row));
        if (ox > 0) { ok_lo = true; lo = __ldg(reinterpret_cast<const float4*>(row - 4)); }
      }
      float v[7] = {lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      if (normalise) {
#pragma unroll
        for (int e = 0; e < 7; ++e) {
          const bool ok = e < 3 ? ok_lo : ok_hi;            // zero padding stays zero (it pads the NORMALISED image)
          v[e] = ok ? (v[e] - mean[c]) * istd[c] : 0.f;
        }
      }
      bf16* dst = tile + px * PI_PITCH + q * 7;
#pragma unroll
      for (int e = 0; e < 7; ++e) dst[e] = __float2bfloat16(v[e]);
    }
    for (int i = threadIdx.x; i < 13 * PI_PIX; i += 256) {   // K padding 147 -> 160
      const int px = i & (PI_PIX - 1), e = i >> 6;
      tile[px * PI_PITCH + 147 + e] = __float2bfloat16(0.f);
    }
    __syncthreads();
    const uint32_t* tw = reinterpret_cast<const uint32_t*>(tile);
    uint32_t* out = cols + (((long)b * So + oy) * So + ox0) * 80;
    for (int i = threadIdx.x; i < PI_PIX * 80; i += 256) {
      const int px = i / 80, w = i - px * 80;
      out[i] = tw[px * (PI_PITCH / 2) + w];
    }
    __syncthreads();
  }
}
int patch_im2col(const float* img, bf16* cols, int B, int S, int normalise, cudaStream_t stream) {
  if (S % (4 * PI_PIX) || (reinterpret_cast<uintptr_t>(img) & 15)) { set_error("patch_im2col: S=%d must be a multiple of %d, image 16 B aligned", S, 4 * PI_PIX); return 1; }
  const long total = (long)B * (S / 4) * (S / 4) * 20;
  const long n_tiles = (long)B * (S / 4) * (S / 4 / PI_PIX);
  long blocks = n_tiles;
  if (blocks > (long)num_sms() * 8) blocks = (long)num_sms() * 8;
  ProfRange pr(stream, "patch_im2col", 0, (double)B * S * S * 3 * 4.0 + (double)total * 16.0);
  patch_im2col_kernel<<<(unsigned)blocks, 256, 0, stream>>>(img, reinterpret_cast<uint32_t*>(cols), B, S, normalise);
  SAM2B_CHECK_LAUNCH("patch_im2col");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Pre-processing: uint8 HWC frames -> fp32 NCHW in [0,1], optionally (x - mean) / std.
// Reference: SAM2Transforms.__call__ (sam2/utils/transforms.py:38-46: ToTensor = /255, Resize is the identity at the
// model resolution, Normalize = sub(mean).div(std)); same IEEE operations in the same order -> bit-exact.
// One thread per 4 pixels of a row: 12 input bytes (three aligned 32-bit loads), one float4 store per colour plane.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) preprocess_u8_kernel(const uint32_t* __restrict__ in, float4* __restrict__ out, long quads,
                                                            long quads_per_img, int normalise) {
  const float mean[3] = {0.485f, 0.456f, 0.406f};
  const float stdv[3] = {0.229f, 0.224f, 0.225f};
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (long)gridDim.x * blockDim.x) {
    const uint32_t w0 = __ldg(in + 3 * i), w1 = __ldg(in + 3 * i + 1), w2 = __ldg(in + 3 * i + 2);
    const uint8_t b[12] = {(uint8_t)w0, (uint8_t)(w0 >> 8), (uint8_t)(w0 >> 16), (uint8_t)(w0 >> 24),
                           (uint8_t)w1, (uint8_t)(w1 >> 8), (uint8_t)(w1 >> 16), (uint8_t)(w1 >> 24),
                           (uint8_t)w2, (uint8_t)(w2 >> 8), (uint8_t)(w2 >> 16), (uint8_t)(w2 >> 24)};
    const long img = i / quads_per_img, q = i - img * quads_per_img;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      float v[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        v[e] = __fdiv_rn((float)b[3 * e + c], 255.0f);
        if (normalise) v[e] = __fdiv_rn(__fsub_rn(v[e], mean[c]), stdv[c]);
      }
      out[(img * 3 + c) * quads_per_img + q] = make_float4(v[0], v[1], v[2], v[3]);
    }
  }
}
int preprocess_u8_hwc(const uint8_t* in, float* out, int n, int H, int W, int normalise, cudaStream_t stream) {
  if (n <= 0 || H <= 0 || W <= 0 || W % 4 || (reinterpret_cast<uintptr_t>(in) & 3) || (reinterpret_cast<uintptr_t>(out) & 15)) {
    set_error("preprocess_u8_hwc: n=%d H=%d W=%d (W must be a multiple of 4, input 4 B / output 16 B aligned)", n, H, W);
    return 1;
  }
  const long qpi = (long)H * W / 4, quads = qpi * n;
  long blocks = (quads + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "preprocess_u8", 0, (double)quads * (12.0 + 48.0));
  preprocess_u8_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(in), reinterpret_cast<float4*>(out), quads, qpi,
                                                             normalise);
  SAM2B_CHECK_LAUNCH("preprocess_u8");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Post-processing: low-resolution mask logits -> output resolution, bilinear, align_corners = False,
// optionally thresholded in the same pass.  Reference: SAM2Transforms.postprocess_masks (sam2/utils/transforms.py:119,
// F.interpolate(masks, orig_hw, mode="bilinear", align_corners=False)), then `masks > mask_threshold`
// (sam2_image_predictor.py:300-302; sam2_video_predictor_official.py `video_res_masks > 0`).
// Arithmetic = torch's upsample_bilinear2d; the source coordinate is one fma (as torch's builds contract it), the blend is
// separately rounded fp32 operations:
//   src = max(fma(scale, dst + 0.5, -0.5), 0), scale = (float)in / out;  i0 = (int)src;  i1 = i0 + (i0 < in - 1);
//   l1 = src - i0, l0 = 1 - l1;  out = hl0 * (wl0 * v00 + wl1 * v01) + hl1 * (wl0 * v10 + wl1 * v11).
// HBM-bound on the store: 4 B (+ 1 B mask) per output pixel; the low-resolution source (256 KB per mask) stays in L1/L2.
// VEC = 4 output columns per thread: one 16-byte store + one 4-byte mask store per row (VEC = 1 when W % 4 != 0).
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void bilinear_coord(int dst, float scale, int in_size, int& i0, int& i1, float& l0, float& l1) {
  float src = __fmaf_rn(scale, __fadd_rn((float)dst, 0.5f), -0.5f);
  src = src < 0.f ? 0.f : src;
  i0 = (int)src;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = __fsub_rn(src, (float)i0);
  l1 = l1 < 0.f ? 0.f : (l1 > 1.f ? 1.f : l1);
  l0 = __fsub_rn(1.f, l1);
}
// Each thread owns VEC output columns (their source columns and weights are computed once) and walks UP_ROWS output rows (16: 0.074 -> 0.066 ms against 8 for 64 masks -> 1024^2);
// the horizontally blended source rows `top` / `bot` are recomputed only when the source row pair changes (every 4th output
// row at 256 -> 1024), which leaves ~1 fma-class operation per output pixel per row and makes the kernel store-bound.
constexpr int UP_ROWS = 16;
template <int VEC>
__global__ void __launch_bounds__(256) upsample_bilinear_mask_kernel(const float* __restrict__ low, float* __restrict__ out,
                                                                     uint8_t* __restrict__ mask, int n, int h, int w, int H, int W,
                                                                     float sh, float sw, float threshold) {
  const unsigned gpr = (unsigned)(W / VEC);                     // column groups per output row
  const unsigned tiles = (unsigned)((H + UP_ROWS - 1) / UP_ROWS);
  const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;     // (row tile, column group) of one image
  if (t >= gpr * tiles) return;
  const unsigned tile = t / gpr;
  const int x0 = (int)(t - tile * gpr) * VEC;
  int c0[VEC], c1[VEC];
  float wl0[VEC], wl1[VEC];
#pragma unroll
  for (int e = 0; e < VEC; ++e) bilinear_coord(x0 + e, sw, w, c0[e], c1[e], wl0[e], wl1[e]);
  const int y_begin = (int)tile * UP_ROWS, y_end = y_begin + UP_ROWS < H ? y_begin + UP_ROWS : H;
  for (int img = blockIdx.y; img < n; img += gridDim.y) {
    const float* src = low + (long)img * h * w;
    float top[VEC], bot[VEC];
    int py0 = -1, py1 = -1;
    for (int y = y_begin; y < y_end; ++y) {
      int y0, y1; float hl0, hl1;
      bilinear_coord(y, sh, h, y0, y1, hl0, hl1);
      if (y0 != py0) {
        if (y0 == py1) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) top[e] = bot[e];
        } else {
          const float* r0 = src + (long)y0 * w;
#pragma unroll
          for (int e = 0; e < VEC; ++e) top[e] = __fadd_rn(__fmul_rn(wl0[e], __ldg(r0 + c0[e])), __fmul_rn(wl1[e], __ldg(r0 + c1[e])));
        }
      }
      if (y1 != py1) {
        if (y1 == y0) {
#pragma unroll
          for (int e = 0; e < VEC; ++e) bot[e] = top[e];
        } else {
          const float* r1 = src + (long)y1 * w;
#pragma unroll
          for (int e = 0; e < VEC; ++e) bot[e] = __fadd_rn(__fmul_rn(wl0[e], __ldg(r1 + c0[e])), __fmul_rn(wl1[e], __ldg(r1 + c1[e])));
        }
      }
      py0 = y0; py1 = y1;
      float v[VEC];
#pragma unroll
      for (int e = 0; e < VEC; ++e) v[e] = __fadd_rn(__fmul_rn(hl0, top[e]), __fmul_rn(hl1, bot[e]));
      const long o = ((long)img * H + y) * W + x0;
      if constexpr (VEC == 4) {
        if (out) __stcs(reinterpret_cast<float4*>(out + o), make_float4(v[0], v[1], v[2], v[3]));
        if (mask) __stcs(reinterpret_cast<uint32_t*>(mask + o), (v[0] > threshold ? 1u : 0u) | (v[1] > threshold ? 0x100u : 0u) |
                                                                  (v[2] > threshold ? 0x10000u : 0u) | (v[3] > threshold ? 0x1000000u : 0u));
      } else {
        if (out) __stcs(out + o, v[0]);
        if (mask) mask[o] = v[0] > threshold ? 1 : 0;
      }
    }
  }
}
int upsample_bilinear_mask(const float* low, float* out, uint8_t* mask, long n, int h, int w, int H, int W, float threshold, cudaStream_t stream) {
  if (n <= 0 || n > 0x7fffffffL || h <= 0 || w <= 0 || H <= 0 || W <= 0 || !low || (!out && !mask)) {
    set_error("upsample_bilinear_mask: n=%ld h=%d w=%d H=%d W=%d (needs a source and at least one of out / mask)", n, h, w, H, W);
    return 1;
  }
  const float sh = (float)h / (float)H, sw = (float)w / (float)W;
  const bool vec = W % 4 == 0 && !(reinterpret_cast<uintptr_t>(out) & 15) && !(reinterpret_cast<uintptr_t>(mask) & 3);
  const long tiles = (H + UP_ROWS - 1) / UP_ROWS;
  const long per_img = tiles * (vec ? W / 4 : W);
  if (per_img > 0x7fffffffL) { set_error("upsample_bilinear_mask: output %d x %d too large", H, W); return 1; }
  dim3 grid((unsigned)((per_img + 255) / 256), (unsigned)(n < 65535 ? n : 65535));
  ProfRange pr(stream, "upsample_bilinear_mask", 0, (double)n * H * W * ((out ? 4.0 : 0.0) + (mask ? 1.0 : 0.0)) + (double)n * h * w * 4.0);
  if (vec) upsample_bilinear_mask_kernel<4><<<grid, 256, 0, stream>>>(low, out, mask, (int)n, h, w, H, W, sh, sw, threshold);
  else upsample_bilinear_mask_kernel<1><<<grid, 256, 0, stream>>>(low, out, mask, (int)n, h, w, H, W, sh, sw, threshold);
  SAM2B_CHECK_LAUNCH("upsample_bilinear_mask");
  return 0;
}

// ------------------------------------------------------------------------------------------
// FPN top-down: dst += nearest_x2(src)   (reference: image_encoder.py:115-127, fpn_interp_model nearest)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) upsample2x_add_kernel(float4* __restrict__ dst, const float2* __restrict__ src, long total4, int W4,
                                                             int H, int Ws2) {
  // dst viewed as float4 (4 consecutive x) ; src as float2 (2 consecutive x)
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += (long)gridDim.x * blockDim.x) {
    const int x4 = (int)(i % W4);
    long t = i / W4;
    const int y = (int)(t % H);
    const long bc = t / H;
    const float2 s = src[(bc * (H >> 1) + (y >> 1)) * Ws2 + x4];
    float4 d = dst[i];
    d.x += s.x; d.y += s.x; d.z += s.y; d.w += s.y;
    dst[i] = d;
  }
}
int upsample2x_add_nchw(float* dst, const float* src, int B, int C, int H, int W, cudaStream_t stream) {
  if (W % 4 || H % 2) { set_error("upsample2x_add: bad shape"); return 1; }
  const long total4 = (long)B * C * H * (W / 4);
  long blocks = (total4 + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "upsample2x_add", 0, (double)total4 * 16.0 * 2.25);
  upsample2x_add_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<float4*>(dst), reinterpret_cast<const float2*>(src), total4, W / 4, H, W / 4);
  SAM2B_CHECK_LAUNCH("upsample2x_add");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Sine position encoding of one feature level (PositionEmbeddingSine, reference modeling/position_encoding.py:90-124 with
// normalize = True, scale = 2 pi, temperature 10000): out [C][H][W]; channels [0, C/2) encode the row, [C/2, C) the column;
// inside a half, channel j holds sin (j even) / cos (j odd) of pos / 10000^(2 floor(j / 2) / (C / 2)),
// pos = (index + 1) / (extent + 1e-6) * 2 pi.  A constant of the model: computed once per (C, H, W) and cached by the caller.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) position_sine_kernel(float* __restrict__ out, int C, int H, int W) {
  const long total = (long)C * H * W;
  const int npf = C / 2;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int x = (int)(i % W);
    const int y = (int)((i / W) % H);
    const int c = (int)(i / ((long)W * H));
    const int j = c < npf ? c : c - npf;
    const float pos = c < npf ? (float)(y + 1) / ((float)H + 1e-6f) * 6.283185307179586f
                              : (float)(x + 1) / ((float)W + 1e-6f) * 6.283185307179586f;
    const float dim_t = powf(10000.0f, (float)(2 * (j / 2)) / (float)npf);
    const float a = pos / dim_t;
    out[i] = (j & 1) ? cosf(a) : sinf(a);
  }
}
int position_sine(float* out, int C, int H, int W, cudaStream_t stream) {
  if (!out || C <= 0 || (C & 3) || H <= 0 || W <= 0) { set_error("position_sine: bad arguments (C=%d must be a positive multiple of 4, H=%d, W=%d)", C, H, W); return 1; }
  const long total = (long)C * H * W;
  long blocks = (total + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "position_sine", 0, (double)total * 4.0);
  position_sine_kernel<<<(unsigned)blocks, 256, 0, stream>>>(out, C, H, W);
  SAM2B_CHECK_LAUNCH("position_sine");
  return 0;
}

// ------------------------------------------------------------------------------------------
// Row softmax fp32 -> bf16 (memory-attention scores).  One CTA per row; the row is staged in shared
// memory so HBM/L2 sees it once.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) softmax_rows_kernel(const float* __restrict__ s, long lds, bf16* __restrict__ p, long ldp, int cols) {
  extern __shared__ float row[];
  __shared__ float red[16];
  const float* sr = s + (long)blockIdx.x * lds;
  bf16* pr = p + (long)blockIdx.x * ldp;
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
  const int c4 = cols >> 2;
  float m = -INFINITY;
  for (int i = tid; i < c4; i += nt) {
    const float4 v = reinterpret_cast<const float4*>(sr)[i];
    reinterpret_cast<float4*>(row)[i] = v;
    m = fmaxf(m, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
  }
  for (int i = (c4 << 2) + tid; i < cols; i += nt) { row[i] = sr[i]; m = fmaxf(m, sr[i]); }
  m = warp_max(m);
  if (lane == 0) red[wid] = m;
  __syncthreads();
  m = red[0];
  for (int i = 1; i < nw; ++i) m = fmaxf(m, red[i]);
  __syncthreads();
  float sum = 0.f;
  for (int i = tid; i < cols; i += nt) {
    const float e = __expf(row[i] - m);
    row[i] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[wid] = sum;
  __syncthreads();
  sum = 0.f;
  for (int i = 0; i < nw; ++i) sum += red[i];
  const float inv = 1.0f / sum;
  const int ldp2 = (int)(ldp >> 1);
  for (int i = tid; i < ldp2; i += nt) {
    const int c = 2 * i;
    const float a = c < cols ? row[c] * inv : 0.f;
    const float b = c + 1 < cols ? row[c + 1] * inv : 0.f;
    reinterpret_cast<uint32_t*>(pr)[i] = pack_bf16x2(a, b);
  }
}
int softmax_rows(const float* s, long lds, bf16* p, long ldp, long rows, int cols, cudaStream_t stream) {
  const size_t smem = (size_t)cols * 4 + 16;
  if (smem > 200 * 1024 || (lds % 4) || (ldp % 2) || ldp < cols) { set_error("softmax_rows: unsupported cols=%d lds=%ld ldp=%ld", cols, lds, ldp); return 1; }
  static PerDeviceFlag attr;
  if (!attr.is_set()) {
    cudaFuncSetAttribute(softmax_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr.set();
  }
  ProfRange pr(stream, "softmax_rows", 0, (double)rows * cols * 6.0);
  softmax_rows_kernel<<<(unsigned)rows, 512, smem, stream>>>(s, lds, p, ldp, cols);
  SAM2B_CHECK_LAUNCH("softmax_rows");
  return 0;
}

}  // namespace sam2b
