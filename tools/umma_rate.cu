// Dev tool (not part of the library): issue rate of tcgen05.mma (M = 128, K = 16, bf16) as a function of N, of where the A
// operand lives (shared memory descriptor vs tensor memory) and of the B layout (K-major / MN-major).  One CTA per SM, one
// thread issues ITER instructions back to back and waits for the commit; prints clocks per instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sam2_opt_b200/csrc tools/umma_rate.cu -o gpurun_out/umma_rate
#include <cstdio>
#include <cstdlib>
#include "common.cuh"

using namespace sam2b;

namespace sam2b { void set_error(const char*, ...) {} }

template <int NACC>
__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int a_tmem, int b_mn, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  for (int i = threadIdx.x; i < (16384 + 65536) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&tmem_ptr, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_ptr;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
    const uint32_t a = smem_u32(smem), b = a + 16384;
    uint64_t da[4], db[4];
    uint32_t ta[4], d[4];
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      db[ks] = b_mn ? make_smem_desc(b + ks * 2048, 8192, 1024, 2) : make_smem_desc(b + ks * 32, 16, 1024, 2);
      da[ks] = make_smem_desc(a + ks * 32, 16, 1024, 2);
      ta[ks] = tb + 448 + 8 * ks;
      d[ks] = tb + (uint32_t)((ks % NACC) * N);          // NACC independent accumulators, round robin
    }
    long long t0 = clock64();
    if (a_tmem) {
#pragma unroll 1
      for (int i = 0; i < iters; i += 4) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) umma_bf16_ts(d[ks], ta[ks], db[ks], idesc, 1);
      }
    } else {
#pragma unroll 1
      for (int i = 0; i < iters; i += 4) {
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) umma_bf16(d[ks], da[ks], db[ks], idesc, 1);
      }
    }
    umma_commit(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tb, 512); }
}

int main() {
  long long* out;
  cudaMalloc(&out, 8);
  const int smem = 16384 + 65536 + 2048;
  cudaFuncSetAttribute(rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(rate_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 4096;
  for (int b_mn = 0; b_mn < 2; ++b_mn)
    for (int a_tmem = 0; a_tmem < 2; ++a_tmem)
      for (int N : {16, 64, 80, 128, 192, 256}) {
        if (b_mn && N > 64 && N % 64) continue;
        for (int nacc : {1, 2, 4}) {
          if (nacc * N > 448) continue;
          const int grid = 148;
          if (nacc == 1) rate_kernel<1><<<grid, 128, smem>>>(N, a_tmem, b_mn, iters, out);
          else if (nacc == 2) rate_kernel<2><<<grid, 128, smem>>>(N, a_tmem, b_mn, iters, out);
          else rate_kernel<4><<<grid, 128, smem>>>(N, a_tmem, b_mn, iters, out);
          long long h = 0;
          cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
          cudaError_t e = cudaDeviceSynchronize();
          printf("B %s  A %s  N=%3d  accumulators=%d: %7.1f clk / instruction (ideal %5.1f)  %s\n", b_mn ? "MN-major" : "K-major ",
                 a_tmem ? "tmem" : "smem", N, nacc, (double)h / iters, N / 2.0, e == cudaSuccess ? "" : cudaGetErrorString(e));
        }
      }
  return 0;
}
