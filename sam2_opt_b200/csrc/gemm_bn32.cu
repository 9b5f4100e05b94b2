// tcgen05 GEMM instantiations for BLOCK_N = 32 (see gemm_kernel.cuh)
#include "gemm_kernel.cuh"

namespace sam2b {
SAM2B_GEMM_INSTANTIATE_GENERIC(32)
}  // namespace sam2b
