// Antialiased bilinear resize of camera / video frames to the model resolution.
// Reference: SAM2ImagePredictor._set_image_ (sam2/sam2/sam2_image_predictor.py:173-196) -> torchvision Resize((1024, 1024))
// (sam2/utils/transforms.py:30) = F.interpolate(mode="bilinear", antialias=True, align_corners=False) on `images / 255.0`.
// The operator is a separable triangle filter whose support grows with the down-scaling factor; oracle/preprocess_oracle.py
// restates it and is bit-identical to torch's CPU operator; the kernels below repeat the same fp32 operations in the same
// order (first tap a plain product, then fused multiply-adds left to right), so they are bit-exact against the oracle.
//
// Three launches, all stream-ordered (no host tables, no host synchronisation):
//   aa_weights_kernel     per output index of one axis: first tap, tap count and the normalised weights
//   aa_horizontal_kernel  uint8 HWC (or fp32 CHW) frame -> fp32 planes [n][3][h][out_w]   (x / 255 through a 256-entry table)
//   aa_vertical_kernel    -> fp32 [n][3][out_h][out_w], optional (x - mean) / std, four output columns per thread
// The intermediate (12 bytes per input row per output column) is written and re-read once; at frame sizes (a 1800 x 1200
// frame: 14.7 MB) it stays in the 126 MB L2.  HBM-bound: algorithmic bytes = 3 per input pixel + 12 per output pixel.
#include <cstdio>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

struct AaAxis { int* first; int* count; float* w; int kmax; };

__global__ void __launch_bounds__(128) aa_weights_kernel(int in_size, int out_size, int kmax, int* __restrict__ first,
                                                         int* __restrict__ count, float* __restrict__ w) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= out_size) return;
  // every operation individually rounded (no contraction), as in the oracle
  const float scale = __fdiv_rn((float)in_size, (float)out_size);
  const bool down = scale >= 1.0f;
  const float support = down ? scale : 1.0f;
  const float invscale = down ? __fdiv_rn(1.0f, scale) : 1.0f;
  const float center = __fmul_rn(scale, __fadd_rn((float)i, 0.5f));
  int lo = (int)__fadd_rn(__fsub_rn(center, support), 0.5f);
  lo = lo > 0 ? lo : 0;
  int hi = (int)__fadd_rn(__fadd_rn(center, support), 0.5f);
  hi = hi < in_size ? hi : in_size;
  int n = hi - lo;
  n = n < 0 ? 0 : (n > kmax ? kmax : n);
  float* wi = w + (long)i * kmax;
  float total = 0.0f;
  for (int j = 0; j < n; ++j) {
    const float t = __fmul_rn(__fadd_rn(__fsub_rn((float)(j + lo), center), 0.5f), invscale);
    const float wj = fmaxf(0.0f, __fsub_rn(1.0f, fabsf(t)));
    wi[j] = wj;
    total = __fadd_rn(total, wj);
  }
  if (total != 0.0f)
    for (int j = 0; j < n; ++j) wi[j] = __fdiv_rn(wi[j], total);
  for (int j = n; j < kmax; ++j) wi[j] = 0.0f;
  first[i] = lo;
  count[i] = n;
}

// one thread per (frame, input row, output column): the three colour channels of that column
template <bool U8>
__global__ void __launch_bounds__(256) aa_horizontal_kernel(const void* __restrict__ src, float* __restrict__ mid, int n, int h, int w,
                                                            int out_w, const int* __restrict__ first, const int* __restrict__ count,
                                                            const float* __restrict__ wt, int kmax) {
  __shared__ float lut[256];
  if (U8) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = __fdiv_rn((float)i, 255.0f);   // ToTensor: x / 255
    __syncthreads();
  }
  const long total = (long)n * h * out_w;
  for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long)gridDim.x * blockDim.x) {
    const int xo = (int)(idx % out_w);
    const long ry = idx / out_w;            // frame * h + y
    const int y = (int)(ry % h);
    const long f = ry / h;
    const int lo = first[xo], cnt = count[xo];
    const float* wi = wt + (long)xo * kmax;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f;
    if (U8) {
      const uint8_t* p = reinterpret_cast<const uint8_t*>(src) + ((f * h + y) * (long)w + lo) * 3;
      for (int j = 0; j < cnt; ++j) {
        const float wj = wi[j];
        const float v0 = lut[p[3 * j]], v1 = lut[p[3 * j + 1]], v2 = lut[p[3 * j + 2]];
        if (j == 0) { a0 = __fmul_rn(v0, wj); a1 = __fmul_rn(v1, wj); a2 = __fmul_rn(v2, wj); }
        else { a0 = __fmaf_rn(v0, wj, a0); a1 = __fmaf_rn(v1, wj, a1); a2 = __fmaf_rn(v2, wj, a2); }
      }
    } else {
      const float* p = reinterpret_cast<const float*>(src) + (f * 3 * h + y) * (long)w + lo;
      const long plane = (long)h * w;
      for (int j = 0; j < cnt; ++j) {
        const float wj = wi[j];
        const float v0 = p[j], v1 = p[plane + j], v2 = p[2 * plane + j];
        if (j == 0) { a0 = __fmul_rn(v0, wj); a1 = __fmul_rn(v1, wj); a2 = __fmul_rn(v2, wj); }
        else { a0 = __fmaf_rn(v0, wj, a0); a1 = __fmaf_rn(v1, wj, a1); a2 = __fmaf_rn(v2, wj, a2); }
      }
    }
    const long o = (f * 3 * h + y) * (long)out_w + xo;
    const long plane_o = (long)h * out_w;
    mid[o] = a0; mid[o + plane_o] = a1; mid[o + 2 * plane_o] = a2;
  }
}

// one thread per (frame, channel, output row, 4 output columns)
__global__ void __launch_bounds__(256) aa_vertical_kernel(const float* __restrict__ mid, float* __restrict__ out, int n, int h, int out_h,
                                                          int out_w, const int* __restrict__ first, const int* __restrict__ count,
                                                          const float* __restrict__ wt, int kmax, int normalise) {
  const int w4 = out_w >> 2;
  const long total = (long)n * 3 * out_h * w4;
  for (long idx = (long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long)gridDim.x * blockDim.x) {
    const int x4 = (int)(idx % w4);
    long r = idx / w4;
    const int yo = (int)(r % out_h);
    const long fc = r / out_h;               // frame * 3 + channel
    const int c = (int)(fc % 3);
    const int lo = first[yo], cnt = count[yo];
    const float* wi = wt + (long)yo * kmax;
    const float4* p = reinterpret_cast<const float4*>(mid + (fc * h + lo) * (long)out_w) + x4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 0; j < cnt; ++j) {
      const float wj = wi[j];
      const float4 v = p[(long)j * w4];
      if (j == 0) { a.x = __fmul_rn(v.x, wj); a.y = __fmul_rn(v.y, wj); a.z = __fmul_rn(v.z, wj); a.w = __fmul_rn(v.w, wj); }
      else { a.x = __fmaf_rn(v.x, wj, a.x); a.y = __fmaf_rn(v.y, wj, a.y); a.z = __fmaf_rn(v.z, wj, a.z); a.w = __fmaf_rn(v.w, wj, a.w); }
    }
    if (normalise) {
      // Normalize: (x - mean) / std with the ImageNet constants (sam2/utils/transforms.py:27-28), separately rounded
      const float mean = c == 0 ? 0.485f : (c == 1 ? 0.456f : 0.406f);
      const float sd = c == 0 ? 0.229f : (c == 1 ? 0.224f : 0.225f);
      a.x = __fdiv_rn(__fsub_rn(a.x, mean), sd); a.y = __fdiv_rn(__fsub_rn(a.y, mean), sd);
      a.z = __fdiv_rn(__fsub_rn(a.z, mean), sd); a.w = __fdiv_rn(__fsub_rn(a.w, mean), sd);
    }
    reinterpret_cast<float4*>(out + (fc * out_h + yo) * (long)out_w)[x4] = a;
  }
}

static int aa_kmax(int in_size, int out_size) {
  const float scale = (float)in_size / (float)out_size;
  const float support = scale >= 1.0f ? scale : 1.0f;
  return (int)ceilf(support) * 2 + 1;
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

long resize_frames_workspace(int n, int h, int w, int out_h, int out_w) {
  if (n <= 0 || h <= 0 || w <= 0 || out_h <= 0 || out_w <= 0) return 0;
  const int kx = aa_kmax(w, out_w), ky = aa_kmax(h, out_h);
  size_t b = 0;
  b += align256((size_t)out_w * 8) + align256((size_t)out_w * kx * 4);
  b += align256((size_t)out_h * 8) + align256((size_t)out_h * ky * 4);
  b += align256((size_t)n * 3 * h * out_w * 4);
  return (long)b;
}

int resize_frames(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out, int out_h, int out_w, int normalise,
                  void* workspace, cudaStream_t stream) {
  if (!src || !out || !workspace || n <= 0 || h <= 0 || w <= 0 || out_h <= 0 || out_w <= 0 || (out_w & 3)) {
    set_error("resize_frames: bad arguments (n=%d h=%d w=%d out=%dx%d; out_w must be a multiple of 4)", n, h, w, out_h, out_w);
    return 1;
  }
  const int kx = aa_kmax(w, out_w), ky = aa_kmax(h, out_h);
  uint8_t* ws = reinterpret_cast<uint8_t*>(workspace);
  AaAxis ax, ay;
  ax.first = reinterpret_cast<int*>(ws); ax.count = ax.first + out_w; ws += align256((size_t)out_w * 8);
  ax.w = reinterpret_cast<float*>(ws); ws += align256((size_t)out_w * kx * 4); ax.kmax = kx;
  ay.first = reinterpret_cast<int*>(ws); ay.count = ay.first + out_h; ws += align256((size_t)out_h * 8);
  ay.w = reinterpret_cast<float*>(ws); ws += align256((size_t)out_h * ky * 4); ay.kmax = ky;
  float* mid = reinterpret_cast<float*>(ws);
  {
    ProfRange pr(stream, "aa_weights", 0, (double)(out_w * (kx + 2) + out_h * (ky + 2)) * 4.0);
    aa_weights_kernel<<<(out_w + 127) / 128, 128, 0, stream>>>(w, out_w, kx, ax.first, ax.count, ax.w);
    SAM2B_CHECK_LAUNCH("aa_weights(x)");
    aa_weights_kernel<<<(out_h + 127) / 128, 128, 0, stream>>>(h, out_h, ky, ay.first, ay.count, ay.w);
    SAM2B_CHECK_LAUNCH("aa_weights(y)");
  }
  const long cap = (long)num_sms() * 16;
  {
    const long total = (long)n * h * out_w;
    long blocks = (total + 255) / 256;
    if (blocks > cap) blocks = cap;
    ProfRange pr(stream, "aa_horizontal", 0, (double)n * h * ((double)w * (src_is_u8_hwc ? 3.0 : 12.0) + (double)out_w * 12.0));
    if (src_is_u8_hwc) aa_horizontal_kernel<true><<<(unsigned)blocks, 256, 0, stream>>>(src, mid, n, h, w, out_w, ax.first, ax.count, ax.w, kx);
    else aa_horizontal_kernel<false><<<(unsigned)blocks, 256, 0, stream>>>(src, mid, n, h, w, out_w, ax.first, ax.count, ax.w, kx);
    SAM2B_CHECK_LAUNCH("aa_horizontal");
  }
  {
    const long total = (long)n * 3 * out_h * (out_w / 4);
    long blocks = (total + 255) / 256;
    if (blocks > cap) blocks = cap;
    ProfRange pr(stream, "aa_vertical", 0, (double)n * 3 * ((double)h * out_w * 4.0 + (double)out_h * out_w * 4.0));
    aa_vertical_kernel<<<(unsigned)blocks, 256, 0, stream>>>(mid, out, n, h, out_h, out_w, ay.first, ay.count, ay.w, ky, normalise);
    SAM2B_CHECK_LAUNCH("aa_vertical");
  }
  return 0;
}

}  // namespace sam2b

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference): the same three
// launches as resize_frames() with the workspace in a file-scope device array instead of caller memory, for callers that
// cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the
// library holds the array until the process exits.  All calls on a device share it: calls that overlap on the GPU (two
// streams) race on it; calls on one stream are ordered.  Nothing in it has to start at zero - every call rewrites all it
// reads (tap tables, then the intermediate) - so the contents left by the previous call are harmless.
// ------------------------------------------------------------------------------------------------------------
#include <mutex>

#include "../../include/sam2b.h"

namespace {
static __device__ __align__(256) unsigned char g_resize_scratch[SAM2B_RESIZE_STATIC_MAX_BYTES];

// The array's name is a host-side stand-in in host code: the device address is looked up once per device, with that
// device current, and cached.
int resize_scratch_address(void** out) {
  static std::mutex mu;
  static void* addr[64] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    sam2b::set_error("sam2b_op_resize_frames_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return 1;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_resize_scratch);
    if (e != cudaSuccess || !p) {
      sam2b::set_error("sam2b_op_resize_frames_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return 1;
    }
    addr[dev] = p;
  }
  *out = addr[dev];
  return 0;
}
}  // namespace

extern "C" int sam2b_op_resize_frames_static(const void* src, int src_is_u8_hwc, int n, int h, int w, float* out_nchw, int out_h,
                                             int out_w, int normalise, void* stream) {
  const long need = sam2b::resize_frames_workspace(n, h, w, out_h, out_w);
  if (need <= 0) {
    sam2b::set_error("sam2b_op_resize_frames_static: bad sizes (n=%d h=%d w=%d out=%dx%d)", n, h, w, out_h, out_w);
    return 1;
  }
  if (need > (long)SAM2B_RESIZE_STATIC_MAX_BYTES) {
    sam2b::set_error("sam2b_op_resize_frames_static: n=%d frames of %dx%d -> %dx%d need %ld bytes of scratch, the static array holds %ld; "
                     "use sam2b_op_resize_frames with a caller workspace", n, h, w, out_h, out_w, need, (long)SAM2B_RESIZE_STATIC_MAX_BYTES);
    return 1;
  }
  void* scratch = nullptr;
  if (resize_scratch_address(&scratch)) return 1;
  return sam2b::resize_frames(src, src_is_u8_hwc, n, h, w, out_nchw, out_h, out_w, normalise, scratch, reinterpret_cast<cudaStream_t>(stream));
}
