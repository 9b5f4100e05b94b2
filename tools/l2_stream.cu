// Dev tool (not part of the library): how fast can 148 SMs pull L2-resident data into shared memory with bulk async copies
// (the GEMM / attention operand path)?  Each CTA streams 16 KB chunks of a buffer that fits L2 through a ring of STAGES
// slots, no compute.  Prints aggregate TB/s for several ring depths and working sets.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I sam2_opt_b200/csrc tools/l2_stream.cu -o tools/l2_stream.bin
#include <cstdio>
#include <cstdlib>
#include "common.cuh"

using namespace sam2b;
namespace sam2b { void set_error(const char*, ...) {} }

constexpr int CHUNK = 16384;

__global__ void __launch_bounds__(128) stream_kernel(const uint8_t* __restrict__ buf, long n_chunks, int iters, int stages, int issuers,
                                                     int chunk) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar[4][12];
  const int w = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    for (int k = 0; k < 4; ++k)
      for (int i = 0; i < 12; ++i) mbar_init(&bar[k][i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0 && w < issuers) {
    // issuer w of `issuers`: its own ring of `stages` slots, every issuers-th chunk of this CTA's sequence
    uint8_t* ring = smem + (size_t)w * stages * chunk;
    long c = (long)blockIdx.x * issuers + w;
    const int total = iters / issuers;
    int issued = 0, done = 0;
    while (done < total) {
      while (issued < total && issued - done < stages) {
        const int s = issued % stages;
        mbar_arrive_expect_tx(&bar[w][s], chunk);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(ring + s * chunk)), "l"(buf + (c % n_chunks) * CHUNK), "r"(chunk), "r"(smem_u32(&bar[w][s])) : "memory");
        c += (long)gridDim.x * issuers;
        ++issued;
      }
      mbar_wait(&bar[w][done % stages], (done / stages) & 1);
      ++done;
    }
  }
}

int main() {
  uint8_t* buf;
  const long max_bytes = 1L << 30;
  cudaMalloc(&buf, max_bytes);
  cudaMemset(buf, 1, max_bytes);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 192 * 1024);
  const long ws_mb = 48;
  for (int chunk : {4096, 16384})
    for (int issuers : {1, 2, 4})
      for (int stages : {2, 6}) {
        const long n_chunks = ws_mb * (1L << 20) / CHUNK;
        const int iters = 4096;
        const int grid = 148;
        const size_t smem = (size_t)issuers * stages * chunk;
        stream_kernel<<<grid, 128, smem>>>(buf, n_chunks, 512, stages, issuers, chunk);     // warm L2
        cudaEventRecord(a);
        stream_kernel<<<grid, 128, smem>>>(buf, n_chunks, iters, stages, issuers, chunk);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        printf("L2-resident, 1 CTA/SM, chunk %5d B, %d issuing threads, ring %2d: %7.2f TB/s  (%s)\n", chunk, issuers, stages,
               (double)grid * iters * chunk / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
      }
  return 0;
}
