// Memory-bound helpers of the MemoryAttention sub-graph (boundary E3; reference memory_attention.py:299-349,
// transformer.py:345-424): layout changes between the reference's seq-first [T,n,C] tensors and the batch-major
// GEMM operands, "memory + memory_pos" key inputs, and RoPE fused with the fp32 -> bf16 operand cast.
#include <cstdio>

#include "common.cuh"
#include "ops.h"

namespace sam2b {

__global__ void __launch_bounds__(256) memattn_prep_x_kernel(const float4* __restrict__ curr, const float4* __restrict__ pos,
                                                             float4* __restrict__ x, int T, int n, int C4) {
  const long total = (long)T * n * C4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4);
    long r = i / C4;
    const int b = (int)(r % n);
    const long t = r / n;
    const float4 a = curr[i], p = pos[i];
    x[((long)b * T + t) * C4 + c] = make_float4(a.x + 0.1f * p.x, a.y + 0.1f * p.y, a.z + 0.1f * p.z, a.w + 0.1f * p.w);
  }
}
// rows [3][n] = (spatial keys, pointer tokens, keys in total) of every object from DEVICE counts [2][n] = (frames, pointer
// tokens), clamped to the tensors' extents (the host cannot validate values it never sees)
__global__ void memattn_rows_kernel(const int* __restrict__ counts, int* __restrict__ rows, int n, int L_max, int P_max) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n) return;
  int L = counts[b], P = counts[n + b];
  L = L < 1 ? 1 : (L > L_max ? L_max : L);
  P = P < 0 ? 0 : (P > P_max ? P_max : P);
  P &= ~3;
  rows[b] = L * 4096; rows[n + b] = P; rows[2 * n + b] = L * 4096 + P;
}
int memattn_rows_from_device(const int* counts, int* rows, int n, int L_max, int P_max, cudaStream_t stream) {
  ProfRange pr(stream, "memattn_rows", 0, 0);
  memattn_rows_kernel<<<(n + 127) / 128, 128, 0, stream>>>(counts, rows, n, L_max, P_max);
  SAM2B_CHECK_LAUNCH("memattn_rows");
  return 0;
}

int memattn_prep_x(const float* curr, const float* curr_pos, float* x, int T, int n, int C, cudaStream_t stream) {
  const long total = (long)T * n * (C / 4);
  long blocks = (total + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "memattn_prep_x", 0, (double)total * 48.0);
  memattn_prep_x_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(curr), reinterpret_cast<const float4*>(curr_pos),
                                                             reinterpret_cast<float4*>(x), T, n, C / 4);
  SAM2B_CHECK_LAUNCH("memattn_prep_x");
  return 0;
}

__global__ void __launch_bounds__(256) batch_to_seq_kernel(const float4* __restrict__ x, float4* __restrict__ out, int T, int n, int C4) {
  const long total = (long)T * n * C4;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i % C4);
    long r = i / C4;
    const int b = (int)(r % n);
    const long t = r / n;
    out[i] = x[((long)b * T + t) * C4 + c];
  }
}
int batch_to_seq(const float* x, float* out, int T, int n, int C, cudaStream_t stream) {
  const long total = (long)T * n * (C / 4);
  long blocks = (total + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "batch_to_seq", 0, (double)total * 32.0);
  batch_to_seq_kernel<<<(unsigned)blocks, 256, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(out), T, n, C / 4);
  SAM2B_CHECK_LAUNCH("batch_to_seq");
  return 0;
}

__global__ void __launch_bounds__(256) memattn_prep_memory_kernel(const float4* __restrict__ mem, const float4* __restrict__ mpos, int L,
                                                                  const float4* __restrict__ excl, const float4* __restrict__ pexcl, int P,
                                                                  int n, const int* __restrict__ rows, uint2* __restrict__ kin,
                                                                  uint2* __restrict__ vin, long ldk) {
  // one thread per 4 channels: 16 threads per 64-wide row
  const long total = (long)n * ldk * 16;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int c = (int)(i & 15);
    const long r = i >> 4;
    const int b = (int)(r / ldk);
    const long j = r - (long)b * ldk;
    const long Ls = rows ? (long)rows[b] : (long)L * 4096;      // this object's spatial keys and pointer tokens
    const int Pb = rows ? rows[n + b] : P;
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f), p = m;
    if (j < Ls) {
      const long src = (j * n + b) * 16 + c;            // [L*4096][n][64]
      m = mem[src]; p = mpos[src];
    } else if (j < Ls + Pb) {
      const long src = ((j - Ls) * n + b) * 16 + c;     // [P][n][64]
      m = excl[src]; p = pexcl[src];
    }
    uint2 k, v;
    k.x = pack_bf16x2(m.x + p.x, m.y + p.y); k.y = pack_bf16x2(m.z + p.z, m.w + p.w);
    v.x = pack_bf16x2(m.x, m.y); v.y = pack_bf16x2(m.z, m.w);
    kin[i] = k; vin[i] = v;
  }
}
int memattn_prep_memory(const float* memory, const float* memory_pos, int L, const float* excl, const float* pos_excl, int P,
                        int n, const int* rows, bf16* kin, bf16* vin, long ldk, cudaStream_t stream) {
  const long total = (long)n * ldk * 16;
  long blocks = (total + 255) / 256;
  if (blocks > (long)num_sms() * 16) blocks = (long)num_sms() * 16;
  ProfRange pr(stream, "memattn_prep_memory", 0, (double)total * 48.0);
  memattn_prep_memory_kernel<<<(unsigned)blocks, 256, 0, stream>>>(
      reinterpret_cast<const float4*>(memory), reinterpret_cast<const float4*>(memory_pos), L, reinterpret_cast<const float4*>(excl),
      reinterpret_cast<const float4*>(pos_excl), P, n, rows, reinterpret_cast<uint2*>(kin), reinterpret_cast<uint2*>(vin), ldk);
  SAM2B_CHECK_LAUNCH("memattn_prep_memory");
  return 0;
}

}  // namespace sam2b
