// Kernels of the MemoryEncoder sub-graph (boundary E4; reference modeling/memory_encoder.py): the bandwidth-bound
// front of the MaskDownSampler (conv3x3 s2 + LayerNorm2d + GELU fused per stage, fp32, channels-last so the
// per-pixel LayerNorm2d is thread/quad local), im2col for its GEMM-sized last stage, the NCHW->NHWC operand cast,
// and the CXBlock depthwise 7x7 + LayerNorm2d.  All 1x1 / pointwise convs run on the tcgen05 GEMM (engine.cu).
#include <cstdio>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

// ---- LayerNorm + GELU rows (C <= 256) ----------------------------------------------------------------------
__global__ void __launch_bounds__(256) layernorm_gelu_rows_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                                  const float* __restrict__ beta, float eps, bf16* __restrict__ y,
                                                                  long rows, int C) {
  const int lane = threadIdx.x & 31;
  const long row = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nv = C >> 2;
  float4 v[2];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = lane + i * 32;
    v[i] = idx < nv ? reinterpret_cast<const float4*>(x + row * C)[idx] : make_float4(0.f, 0.f, 0.f, 0.f);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  const float mean = warp_sum(s) / (float)C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    if (lane + i * 32 < nv) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
  }
  const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int idx = lane + i * 32;
    if (idx < nv) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + idx);
      const float4 b = __ldg(reinterpret_cast<const float4*>(beta) + idx);
      uint2 pk;
      pk.x = pack_bf16x2(gelu_erf((v[i].x - mean) * rstd * g.x + b.x), gelu_erf((v[i].y - mean) * rstd * g.y + b.y));
      pk.y = pack_bf16x2(gelu_erf((v[i].z - mean) * rstd * g.z + b.z), gelu_erf((v[i].w - mean) * rstd * g.w + b.w));
      reinterpret_cast<uint2*>(y + row * C)[idx] = pk;
    }
  }
}
int layernorm_gelu_rows(const float* x, const float* gamma, const float* beta, float eps, bf16* y, long rows, int C, cudaStream_t stream) {
  if (C % 4 || C > 256) { set_error("layernorm_gelu_rows: unsupported C=%d", C); return 1; }
  ProfRange pr(stream, "layernorm_gelu", 0, (double)rows * C * 6.0);
  layernorm_gelu_rows_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, stream>>>(x, gamma, beta, eps, y, rows, C);
  SAM2B_CHECK_LAUNCH("layernorm_gelu_rows");
  return 0;
}

// ---- MaskDownSampler stages -------------------------------------------------------------------------------
template <int N>
SAM2B_DEVINL void ln_gelu_inplace(float (&o)[N], const float* ln_w, const float* ln_b, int c0, float mean, float rstd) {
#pragma unroll
  for (int i = 0; i < N; ++i) o[i] = gelu_erf((o[i] - mean) * rstd * ln_w[c0 + i] + ln_b[c0 + i]);
}

// stage 1: 1 -> 4 channels, thread per output pixel
__global__ void __launch_bounds__(256) mask_ds1_kernel(const float* __restrict__ in, float4* __restrict__ out, int n, int S,
                                                       const float* __restrict__ w, const float* __restrict__ b,
                                                       const float* __restrict__ ln_w, const float* __restrict__ ln_b) {
  __shared__ float sw[36 + 12];
  if (threadIdx.x < 36) sw[threadIdx.x] = w[threadIdx.x];
  else if (threadIdx.x < 40) sw[threadIdx.x] = b[threadIdx.x - 36];
  else if (threadIdx.x < 44) sw[threadIdx.x] = ln_w[threadIdx.x - 40];
  else if (threadIdx.x < 48) sw[threadIdx.x] = ln_b[threadIdx.x - 44];
  __syncthreads();
  const int So = S >> 1;
  const long total = (long)n * So * So;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int ox = (int)(i % So);
    const int oy = (int)((i / So) % So);
    const long img = i / ((long)So * So);
    const float* src = in + img * S * S;
    float o[4] = {sw[36], sw[37], sw[38], sw[39]};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int y = 2 * oy - 1 + ky;
      if (y < 0 || y >= S) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int x = 2 * ox - 1 + kx;
        if (x < 0 || x >= S) continue;
        const float v = __ldg(src + (long)y * S + x);
#pragma unroll
        for (int c = 0; c < 4; ++c) o[c] = fmaf(v, sw[c * 9 + ky * 3 + kx], o[c]);
      }
    }
    const float mean = 0.25f * ((o[0] + o[1]) + (o[2] + o[3]));
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) q += (o[c] - mean) * (o[c] - mean);
    const float rstd = rsqrtf(0.25f * q + 1e-6f);
    ln_gelu_inplace<4>(o, sw + 40, sw + 44, 0, mean, rstd);
    out[i] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// stage 1 straight from the low-resolution mask logits [n][256][256]: the 1024^2 input of the down-sampler,
//   mask_for_mem = sigmoid(bilinear_x4(low)) * 20 - 10        (binarize: (bilinear_x4(low) > 0) * 20 - 10)
// (SAM2Base._encode_new_memory, sam2_base_official.py:1000-1013, on the high-res mask of _forward_sam_heads :447-452), is
// evaluated per tap instead of being materialised: bilinear weights in the arithmetic of upsample_bilinear_mask (fp32, no
// contraction), then the same conv3x3 s2 + LayerNorm2d + GELU as mask_ds1_kernel.
__device__ __forceinline__ void ds_bilinear_coord(int dst, int in_size, int& i0, int& i1, float& l0, float& l1) {
  float src = __fmaf_rn(0.25f, __fadd_rn((float)dst, 0.5f), -0.5f);
  src = src < 0.f ? 0.f : src;
  i0 = (int)src;
  if (i0 > in_size - 1) i0 = in_size - 1;
  i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
  l1 = __fsub_rn(src, (float)i0);
  l1 = l1 < 0.f ? 0.f : (l1 > 1.f ? 1.f : l1);
  l0 = __fsub_rn(1.f, l1);
}
__global__ void __launch_bounds__(256) mask_ds1_low_kernel(const float* __restrict__ low, float4* __restrict__ out, int n, int binarize,
                                                           const float* __restrict__ w, const float* __restrict__ b,
                                                           const float* __restrict__ ln_w, const float* __restrict__ ln_b) {
  __shared__ float sw[36 + 12];
  if (threadIdx.x < 36) sw[threadIdx.x] = w[threadIdx.x];
  else if (threadIdx.x < 40) sw[threadIdx.x] = b[threadIdx.x - 36];
  else if (threadIdx.x < 44) sw[threadIdx.x] = ln_w[threadIdx.x - 40];
  else if (threadIdx.x < 48) sw[threadIdx.x] = ln_b[threadIdx.x - 44];
  __syncthreads();
  const int S = 1024, So = 512, L = 256;
  const long total = (long)n * So * So;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int ox = (int)(i % So);
    const int oy = (int)((i / So) % So);
    const float* src = low + (i / ((long)So * So)) * L * L;
    int cx0[3], cx1[3];
    float wx0[3], wx1[3];
#pragma unroll
    for (int kx = 0; kx < 3; ++kx) {
      const int x = 2 * ox - 1 + kx;
      ds_bilinear_coord(x < 0 ? 0 : (x >= S ? S - 1 : x), L, cx0[kx], cx1[kx], wx0[kx], wx1[kx]);
    }
    float o[4] = {sw[36], sw[37], sw[38], sw[39]};
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int y = 2 * oy - 1 + ky;
      if (y < 0 || y >= S) continue;
      int y0, y1; float hl0, hl1;
      ds_bilinear_coord(y, L, y0, y1, hl0, hl1);
      const float* r0 = src + y0 * L;
      const float* r1 = src + y1 * L;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int x = 2 * ox - 1 + kx;
        if (x < 0 || x >= S) continue;
        const float top = __fadd_rn(__fmul_rn(wx0[kx], __ldg(r0 + cx0[kx])), __fmul_rn(wx1[kx], __ldg(r0 + cx1[kx])));
        const float bot = __fadd_rn(__fmul_rn(wx0[kx], __ldg(r1 + cx0[kx])), __fmul_rn(wx1[kx], __ldg(r1 + cx1[kx])));
        const float hi = __fadd_rn(__fmul_rn(hl0, top), __fmul_rn(hl1, bot));
        const float p = binarize ? (hi > 0.f ? 1.0f : 0.0f) : 1.0f / (1.0f + expf(-hi));
        const float v = __fadd_rn(__fmul_rn(p, 20.0f), -10.0f);
#pragma unroll
        for (int c = 0; c < 4; ++c) o[c] = fmaf(v, sw[c * 9 + ky * 3 + kx], o[c]);
      }
    }
    const float mean = 0.25f * ((o[0] + o[1]) + (o[2] + o[3]));
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) q += (o[c] - mean) * (o[c] - mean);
    const float rstd = rsqrtf(0.25f * q + 1e-6f);
    ln_gelu_inplace<4>(o, sw + 40, sw + 44, 0, mean, rstd);
    out[i] = make_float4(o[0], o[1], o[2], o[3]);
  }
}
int mask_ds1_from_low(const float* low, float* out, int n, int binarize, const float* w, const float* b, const float* ln_w,
                      const float* ln_b, cudaStream_t stream) {
  const long pix = (long)n * 512 * 512;
  ProfRange pr(stream, "mask_ds1_low", 2.0 * pix * 36, (double)pix * 16.0 + (double)n * 65536 * 4.0);
  long blocks = (pix + 255) / 256; if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  mask_ds1_low_kernel<<<(unsigned)blocks, 256, 0, stream>>>(low, reinterpret_cast<float4*>(out), n, binarize, w, b, ln_w, ln_b);
  SAM2B_CHECK_LAUNCH("mask_ds1_low");
  return 0;
}

// out[b][c][p] += no_obj[c] for the objects whose score says "not there" (sam2_base_official.py:1019-1024)
__global__ void add_no_obj_embed_kernel(float* __restrict__ x, const float* __restrict__ obj, const float* __restrict__ no_obj, long total) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int b = (int)(i / (64 * 4096)), c = (int)((i / 4096) % 64);
  if (!(obj[b] > 0.f)) x[i] += no_obj[c];
}
int add_no_obj_embed(float* x, const float* obj, const float* no_obj, int n, cudaStream_t stream) {
  const long total = (long)n * 64 * 4096;
  ProfRange pr(stream, "add_no_obj_embed", 0, (double)total * 8.0);
  add_no_obj_embed_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(x, obj, no_obj, total);
  SAM2B_CHECK_LAUNCH("add_no_obj_embed");
  return 0;
}

// stage 2: 4 -> 16 channels, NHWC in, thread per output pixel
__global__ void __launch_bounds__(128) mask_ds2_kernel(const float4* __restrict__ in, float4* __restrict__ out, int n, int H,
                                                       const float* __restrict__ w, const float* __restrict__ b,
                                                       const float* __restrict__ ln_w, const float* __restrict__ ln_b) {
  __shared__ float sw[9 * 4 * 16];   // [tap][cin][cout]
  __shared__ float sp[48];
  for (int i = threadIdx.x; i < 576; i += blockDim.x) {
    const int co = i / 36, r = i % 36, ci = r / 9, tap = r % 9;     // reference layout [co][ci][ky][kx]
    sw[(tap * 4 + ci) * 16 + co] = w[i];
  }
  if (threadIdx.x < 16) { sp[threadIdx.x] = b[threadIdx.x]; sp[16 + threadIdx.x] = ln_w[threadIdx.x]; sp[32 + threadIdx.x] = ln_b[threadIdx.x]; }
  __syncthreads();
  const int Ho = H >> 1;
  const long total = (long)n * Ho * Ho;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int ox = (int)(i % Ho);
    const int oy = (int)((i / Ho) % Ho);
    const long img = i / ((long)Ho * Ho);
    float o[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) o[c] = sp[c];
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int y = 2 * oy - 1 + ky;
      if (y < 0 || y >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int x = 2 * ox - 1 + kx;
        if (x < 0 || x >= H) continue;
        const float4 v = __ldg(in + (img * H + y) * H + x);
        const float* wt = sw + (ky * 3 + kx) * 64;
#pragma unroll
        for (int c = 0; c < 16; ++c) o[c] = fmaf(v.x, wt[c], fmaf(v.y, wt[16 + c], fmaf(v.z, wt[32 + c], fmaf(v.w, wt[48 + c], o[c]))));
      }
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) s += o[c];
    const float mean = s * (1.f / 16.f);
    float q = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) q += (o[c] - mean) * (o[c] - mean);
    const float rstd = rsqrtf(q * (1.f / 16.f) + 1e-6f);
    ln_gelu_inplace<16>(o, sp + 16, sp + 32, 0, mean, rstd);
#pragma unroll
    for (int c = 0; c < 4; ++c) out[i * 4 + c] = make_float4(o[4 * c], o[4 * c + 1], o[4 * c + 2], o[4 * c + 3]);
  }
}

// stage 3 (16 -> 64 channels) and stage 4 are dense enough for the tensor cores: im2col + tcgen05 GEMM + layernorm_gelu_rows (engine.cu)

int mask_ds_stage(int stage, const float* in, float* out, int n, int H_in, const float* w, const float* b, const float* ln_w,
                  const float* ln_b, cudaStream_t stream) {
  const int Ho = H_in / 2;
  const long pix = (long)n * Ho * Ho;
  char nm[32];
  snprintf(nm, sizeof(nm), "mask_ds%d", stage);
  if (stage == 1) {
    ProfRange pr(stream, nm, 2.0 * pix * 36, (double)pix * 32.0);
    long blocks = (pix + 255) / 256; if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
    mask_ds1_kernel<<<(unsigned)blocks, 256, 0, stream>>>(in, reinterpret_cast<float4*>(out), n, H_in, w, b, ln_w, ln_b);
  } else if (stage == 2) {
    ProfRange pr(stream, nm, 2.0 * pix * 576, (double)pix * (64.0 + 64.0));
    long blocks = (pix + 127) / 128; if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
    mask_ds2_kernel<<<(unsigned)blocks, 128, 0, stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out), n, H_in, w, b, ln_w, ln_b);
  } else {
    set_error("mask_ds_stage: bad stage %d", stage);
    return 1;
  }
  SAM2B_CHECK_LAUNCH("mask_ds_stage");
  return 0;
}

// ---- im2col 3x3 s2 p1, NHWC fp32 -> bf16 ---------------------------------------------------------------
__global__ void __launch_bounds__(256) im2col3x3s2_kernel(const float4* __restrict__ in, uint4* __restrict__ cols, int n, int H, int C8) {
  const int Ho = H >> 1;
  const long total = (long)n * Ho * Ho * 9 * C8;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C8);
    long t = i / C8;
    const int tap = (int)(t % 9); t /= 9;
    const int ox = (int)(t % Ho); t /= Ho;
    const int oy = (int)(t % Ho);
    const long img = t / Ho;
    const int y = 2 * oy - 1 + tap / 3, x = 2 * ox - 1 + tap % 3;
    uint4 pk = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < H && x >= 0 && x < H) {
      const float4* src = in + (((img * H + y) * H + x) * C8 + c) * 2;
      const float4 a = __ldg(src), b = __ldg(src + 1);
      pk.x = pack_bf16x2(a.x, a.y); pk.y = pack_bf16x2(a.z, a.w);
      pk.z = pack_bf16x2(b.x, b.y); pk.w = pack_bf16x2(b.z, b.w);
    }
    cols[i] = pk;
  }
}
int im2col3x3s2_nhwc(const float* in, bf16* cols, int n, int H, int C, cudaStream_t stream) {
  if (C % 8 || H % 2) { set_error("im2col3x3s2: bad shape"); return 1; }
  const long total = (long)n * (H / 2) * (H / 2) * 9 * (C / 8);
  long blocks = (total + 255) / 256; if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "im2col3x3s2", 0, (double)total * 16.0 + (double)n * H * H * C * 4.0);
  im2col3x3s2_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<uint4*>(cols), n, H, C / 8);
  SAM2B_CHECK_LAUNCH("im2col3x3s2");
  return 0;
}

// the same from a bf16 NHWC map (stage 4 after the GEMM form of stage 3): 8 channels = one 16-byte move
__global__ void __launch_bounds__(256) im2col3x3s2_bf16_kernel(const uint4* __restrict__ in, uint4* __restrict__ cols, int n, int H, int C8) {
  const int Ho = H >> 1;
  const long total = (long)n * Ho * Ho * 9 * C8;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C8);
    long t = i / C8;
    const int tap = (int)(t % 9); t /= 9;
    const int ox = (int)(t % Ho); t /= Ho;
    const int oy = (int)(t % Ho);
    const long img = t / Ho;
    const int y = 2 * oy - 1 + tap / 3, x = 2 * ox - 1 + tap % 3;
    uint4 pk = make_uint4(0, 0, 0, 0);
    if (y >= 0 && y < H && x >= 0 && x < H) pk = __ldg(in + ((img * H + y) * H + x) * C8 + c);
    cols[i] = pk;
  }
}
int im2col3x3s2_nhwc_bf16(const bf16* in, bf16* cols, int n, int H, int C, cudaStream_t stream) {
  if (C % 8 || H % 2) { set_error("im2col3x3s2_bf16: bad shape"); return 1; }
  const long total = (long)n * (H / 2) * (H / 2) * 9 * (C / 8);
  long blocks = (total + 255) / 256; if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "im2col3x3s2", 0, (double)total * 16.0 + (double)n * H * H * C * 2.0);
  im2col3x3s2_bf16_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(cols), n, H, C / 8);
  SAM2B_CHECK_LAUNCH("im2col3x3s2_bf16");
  return 0;
}

// ---- [n,C,HW] fp32 -> [n,HW,C] bf16 (32x32 smem tile transpose) ---------------------------------------------
__global__ void __launch_bounds__(256) nchw_to_nhwc_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, int C, int HW,
                                                                long in_batch_stride) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int hw0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  for (int r = ty; r < 32; r += 8) tile[r][tx] = in[b * in_batch_stride + (long)(c0 + r) * HW + hw0 + tx];
  __syncthreads();
  for (int r = ty; r < 32; r += 8) out[((long)b * HW + hw0 + r) * C + c0 + tx] = __float2bfloat16(tile[tx][r]);
}
int nchw_to_nhwc_bf16(const float* in, bf16* out, int n, int C, int HW, cudaStream_t stream, long in_batch_stride) {
  if (C % 32 || HW % 32) { set_error("nchw_to_nhwc_bf16: bad shape"); return 1; }
  ProfRange pr(stream, "nchw_to_nhwc_bf16", 0, (double)n * C * HW * 6.0);
  nchw_to_nhwc_bf16_kernel<<<dim3(HW / 32, C / 32, n), 256, 0, stream>>>(in, out, C, HW, in_batch_stride < 0 ? (long)C * HW : in_batch_stride);
  SAM2B_CHECK_LAUNCH("nchw_to_nhwc_bf16");
  return 0;
}

// ---- CXBlock: depthwise 7x7 + LayerNorm2d; one warp per 4 consecutive pixels of a row, 8 channels per lane ------------
// Register tiling over 4 pixels: a row of the 7x7 window spans 10 input columns for the 4 outputs (instead of 4 x 7) and each
// tap weight is loaded once per 4 pixels: 238 instead of 784 16-byte loads per 4 pixels (the kernel is L1-load bound).
__global__ void __launch_bounds__(256) dwconv7_ln_kernel(const float4* __restrict__ x, uint4* __restrict__ y, int n, int H,
                                                         const float4* __restrict__ dw_w, const float4* __restrict__ dw_b,
                                                         const float4* __restrict__ ln_w, const float4* __restrict__ ln_b) {
  const int lane = threadIdx.x & 31;
  const int gpr = H >> 2;                                   // 4-pixel groups per row
  const long grp = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (grp >= (long)n * H * gpr) return;
  const int px0 = (int)(grp % gpr) * 4, py = (int)((grp / gpr) % H);
  const long img = grp / ((long)gpr * H);
  float acc[4][8];
  {
    const float4 b0 = __ldg(dw_b + 2 * lane), b1 = __ldg(dw_b + 2 * lane + 1);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      acc[q][0] = b0.x; acc[q][1] = b0.y; acc[q][2] = b0.z; acc[q][3] = b0.w;
      acc[q][4] = b1.x; acc[q][5] = b1.y; acc[q][6] = b1.z; acc[q][7] = b1.w;
    }
  }
  for (int ky = 0; ky < 7; ++ky) {
    const int yy = py - 3 + ky;
    if (yy < 0 || yy >= H) continue;
    float w[7][8];
#pragma unroll
    for (int kx = 0; kx < 7; ++kx) {
      const float4 w0 = __ldg(dw_w + (ky * 7 + kx) * 64 + 2 * lane), w1 = __ldg(dw_w + (ky * 7 + kx) * 64 + 2 * lane + 1);
      w[kx][0] = w0.x; w[kx][1] = w0.y; w[kx][2] = w0.z; w[kx][3] = w0.w;
      w[kx][4] = w1.x; w[kx][5] = w1.y; w[kx][6] = w1.z; w[kx][7] = w1.w;
    }
    const float4* row = x + ((img * H + yy) * H) * 64 + 2 * lane;
#pragma unroll
    for (int col = 0; col < 10; ++col) {
      const int xx = px0 - 3 + col;
      if (xx < 0 || xx >= H) continue;
      const float4 v0 = __ldg(row + (long)xx * 64), v1 = __ldg(row + (long)xx * 64 + 1);
      const float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int kx = col - q;                              // output pixel px0 + q sees this column as tap kx
        if (kx >= 0 && kx < 7) {
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[q][e] = fmaf(v[e], w[kx][e], acc[q][e]);
        }
      }
    }
  }
  const float4 g0 = __ldg(ln_w + 2 * lane), g1 = __ldg(ln_w + 2 * lane + 1);
  const float4 b0 = __ldg(ln_b + 2 * lane), b1 = __ldg(ln_b + 2 * lane + 1);
  const long pix0 = (img * H + py) * H + px0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float* a = acc[q];
    const float mean = warp_sum((a[0] + a[1]) + (a[2] + a[3]) + (a[4] + a[5]) + (a[6] + a[7])) * (1.f / 256.f);
    float d[8];
    float s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { d[i] = a[i] - mean; s2 += d[i] * d[i]; }
    const float rstd = rsqrtf(warp_sum(s2) * (1.f / 256.f) + 1e-6f);
    uint4 pk;
    pk.x = pack_bf16x2(d[0] * rstd * g0.x + b0.x, d[1] * rstd * g0.y + b0.y);
    pk.y = pack_bf16x2(d[2] * rstd * g0.z + b0.z, d[3] * rstd * g0.w + b0.w);
    pk.z = pack_bf16x2(d[4] * rstd * g1.x + b1.x, d[5] * rstd * g1.y + b1.y);
    pk.w = pack_bf16x2(d[6] * rstd * g1.z + b1.z, d[7] * rstd * g1.w + b1.w);
    y[(pix0 + q) * 32 + lane] = pk;
  }
}
int dwconv7_ln_nhwc(const float* x, bf16* y, int n, int H, const float* dw_w, const float* dw_b, const float* ln_w,
                    const float* ln_b, cudaStream_t stream) {
  const long pix = (long)n * H * H;
  ProfRange pr(stream, "dwconv7_ln", 2.0 * pix * 256 * 49, (double)pix * 256 * 6.0);
  if (H % 4) { set_error("dwconv7_ln: H=%d must be a multiple of 4", H); return 1; }
  dwconv7_ln_kernel<<<(unsigned)((pix / 4 + 7) / 8), 256, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<uint4*>(y), n, H,
                                                                  reinterpret_cast<const float4*>(dw_w), reinterpret_cast<const float4*>(dw_b),
                                                                  reinterpret_cast<const float4*>(ln_w), reinterpret_cast<const float4*>(ln_b));
  SAM2B_CHECK_LAUNCH("dwconv7_ln");
  return 0;
}

}  // namespace sam2b
