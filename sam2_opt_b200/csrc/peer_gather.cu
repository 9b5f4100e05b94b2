// Peer-memory plumbing for the one collective of the path: the gather of the image embeddings (feat0, feat1, feat2) onto
// every GPU of the box (the reference has no multi-GPU inference path at all, section 2a).
//
// B200-first design: every rank owns one gather buffer per feature level, allocated here with cudaMalloc and exported as a
// CUDA IPC handle; every other rank maps it once.  The FpnNeck GEMMs write this rank's results straight into its own slice
// of the local buffer, and the slice is then PUSHED into the same slice of every peer's buffer with plain device-to-device
// copies over NVLink 5 / NVSwitch.  Those copies run on the copy engines: they occupy no SM, so the persistent GEMM and
// attention kernels of the next step (one CTA per SM, statically scheduled) keep all 148 SMs while the previous step's
// embeddings are in flight - which an NCCL all-gather kernel on a side stream would not allow (its CTAs take SMs away
// from a persistent kernel and double the duration of every launch they overlap with).
#include <cstdio>
#include <cstring>

#include "../../include/sam2b.h"
#include "ops.h"

using namespace sam2b;

#define PG_OK(expr)                                                                                         \
  do {                                                                                                      \
    cudaError_t e__ = (expr);                                                                               \
    if (e__ != cudaSuccess) {                                                                               \
      set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__);              \
      return 1;                                                                                             \
    }                                                                                                       \
  } while (0)

namespace {
struct DevGuard {
  int prev = -1;
  bool switched = false;
  explicit DevGuard(int dev) {
    if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = (cudaSetDevice(dev) == cudaSuccess);
  }
  ~DevGuard() { if (switched) cudaSetDevice(prev); }
};
}  // namespace

extern "C" {

int sam2b_peer_alloc(int device, int64_t bytes, void** dev_ptr, unsigned char* ipc_handle_64) {
  if (!dev_ptr || !ipc_handle_64 || bytes <= 0) { set_error("sam2b_peer_alloc: bad arguments"); return 1; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  DevGuard g(device);
  void* p = nullptr;
  PG_OK(cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    set_error("cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
    return 1;
  }
  memcpy(ipc_handle_64, &h, 64);
  *dev_ptr = p;
  return 0;
}

int sam2b_peer_free(int device, void* dev_ptr) {
  if (!dev_ptr) return 0;
  DevGuard g(device);
  PG_OK(cudaFree(dev_ptr));
  return 0;
}

int sam2b_peer_open(int device, const unsigned char* ipc_handle_64, void** mapped_ptr) {
  if (!ipc_handle_64 || !mapped_ptr) { set_error("sam2b_peer_open: bad arguments"); return 1; }
  DevGuard g(device);
  cudaIpcMemHandle_t h;
  memcpy(&h, ipc_handle_64, 64);
  PG_OK(cudaIpcOpenMemHandle(mapped_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}

int sam2b_peer_close(int device, void* mapped_ptr) {
  if (!mapped_ptr) return 0;
  DevGuard g(device);
  PG_OK(cudaIpcCloseMemHandle(mapped_ptr));
  return 0;
}

int sam2b_peer_copy(void* dst, const void* src, int64_t bytes, void* stream) {
  if (!dst || !src || bytes < 0) { set_error("sam2b_peer_copy: bad arguments"); return 1; }
  if (bytes == 0) return 0;
  // unified addressing: the runtime routes a copy between local memory and a mapped peer allocation over NVLink on a copy engine
  PG_OK(cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDefault, reinterpret_cast<cudaStream_t>(stream)));
  return 0;
}

}  // extern "C"
