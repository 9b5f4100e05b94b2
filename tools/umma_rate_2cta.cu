// Dev tool (not part of the library): issue rate of tcgen05.mma.cta_group::2 (a CTA pair, M = 256, K = 16, bf16, both operands
// from shared memory) as a function of N.  One thread of the leader CTA issues ITER instructions back to back.  Compare with
// tools/umma_rate.cu (cta_group::1: max(N / 2, ~50) clocks per instruction).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sam2_opt_b200/csrc tools/umma_rate_2cta.cu -o tools/umma_rate_2cta.bin
#include <cstdio>
#include <cstdlib>
#include "common.cuh"

using namespace sam2b;
namespace sam2b { void set_error(const char*, ...) {} }

__global__ void __launch_bounds__(128, 1) rate_kernel(int N, int iters, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const uint32_t rank = cluster_ctarank();
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc_2cta(&tmem_ptr, 512); tmem_relinquish_2cta(); }
  fence_proxy_async_smem();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tb = tmem_ptr;
  if (threadIdx.x == 0 && rank == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (16u << 24);   // M = 256
    const uint32_t a = smem_u32(smem), b = a + 16384;
    uint64_t da[4], db[4];
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      da[ks] = make_smem_desc(a + ks * 32, 16, 1024, 2);
      db[ks] = make_smem_desc(b + ks * 32, 16, 1024, 2);
    }
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; i += 4) {
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) umma_bf16_2cta(tb, da[ks], db[ks], idesc, 1);
    }
    umma_commit_2cta(&bar);
    mbar_wait(&bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) out[0] = t1 - t0;
  } else if (threadIdx.x == 0) {
    mbar_wait(&bar, 0);          // the multicast commit arrives on the peer's barrier too
  }
  tc_fence_before();
  cluster_sync_all();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc_2cta(tb, 512); }
}

int main() {
  long long* out;
  cudaMalloc(&out, 8);
  const int smem = 16384 + 32768 + 2048;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 4096;
  for (int N : {32, 64, 96, 128, 160, 192, 256}) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(148);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, rate_kernel, N, iters, out);
    long long h = 0;
    cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    cudaError_t e2 = cudaDeviceSynchronize();
    printf("cta_group::2  M=256 N=%3d: %7.1f clk / instruction (ideal %5.1f = N/2)  %s %s\n", N, (double)h / iters, N / 2.0,
           e == cudaSuccess ? "" : cudaGetErrorString(e), e2 == cudaSuccess ? "" : cudaGetErrorString(e2));
  }
  return 0;
}
