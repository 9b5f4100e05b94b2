// Connected-component labelling with areas for the hole-filling post-processor of the video predictor.
// Reference: sam2/csrc/connected_components.cu (the fork's only native kernel, block-based union-find, launched once per
// image from a host loop) as used by fill_holes_in_mask_scores (sam2/utils/misc.py:311-338): components of `mask <= 0`
// under 8-connectivity, every pixel of a component reporting the component's area; background pixels report label 0.
//
// This version is pixel-based and batched: four launches label all n masks at once.
//   1. cc_init      parent[i] = i for foreground pixels, -1 otherwise; area[i] = 0
//   2. cc_merge     union with the already-visited half of the 8-neighbourhood (W, NW, N, NE) through atomicMin on the
//                   parent array (lock-free union-find; the smaller index becomes the root)
//   3. cc_count     root = find(i) with path compression; area[root] += 1
//   4. cc_finalize  label = root + 1 (0 for background), count = area[root]
// Everything is integer work on n x H x W int32 arrays (HBM-bound, 256 KB per 256 x 256 mask): coalesced, grid sized in
// multiples of the SM count.
#include "common.cuh"
#include "ops.h"

namespace sam2b {

SAM2B_DEVINL int cc_find(const int* parent, int i) {
  int p = __ldcg(parent + i);          // L2 reads: parents change under concurrent atomicMin
  while (p != i) { i = p; p = __ldcg(parent + i); }
  return i;
}

SAM2B_DEVINL void cc_union(int* parent, int a, int b) {
  for (;;) {
    a = cc_find(parent, a);
    b = cc_find(parent, b);
    if (a == b) return;
    if (a > b) { const int t = a; a = b; b = t; }       // a < b: hang b's root under a
    const int old = atomicMin(parent + b, a);
    if (old == b) return;                               // b was still a root: linked
    b = old;                                            // somebody re-rooted b meanwhile: retry from there
  }
}

__global__ void __launch_bounds__(256) cc_init_kernel(const uint8_t* __restrict__ mask, int* __restrict__ parent, int* __restrict__ area,
                                                      long total, int hw) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    parent[i] = mask[i] ? (int)(i % hw) : -1;
    area[i] = 0;
  }
}

__global__ void __launch_bounds__(256) cc_merge_kernel(const uint8_t* __restrict__ mask, int* __restrict__ parent, long total, int H, int W) {
  const int hw = H * W;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    if (!mask[i]) continue;
    const long base = i / hw * hw;
    const int l = (int)(i - base), y = l / W, x = l - y * W;
    const uint8_t* m = mask + base;
    int* par = parent + base;
    if (x > 0 && m[l - 1]) cc_union(par, l, l - 1);
    if (y > 0) {
      if (m[l - W]) cc_union(par, l, l - W);
      if (x > 0 && m[l - W - 1]) cc_union(par, l, l - W - 1);
      if (x + 1 < W && m[l - W + 1]) cc_union(par, l, l - W + 1);
    }
  }
}

__global__ void __launch_bounds__(256) cc_count_kernel(int* __restrict__ parent, int* __restrict__ area, long total, int hw) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    if (parent[i] < 0) continue;
    const long base = i / hw * hw;
    const int root = cc_find(parent + base, (int)(i - base));
    parent[i] = root;                                   // path compression (still an ancestor for concurrent finds)
    atomicAdd(area + base + root, 1);
  }
}

__global__ void __launch_bounds__(256) cc_finalize_kernel(int* __restrict__ labels, int* __restrict__ counts, long total, int hw) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
    const int root = labels[i];                         // every foreground entry is a root index after cc_count
    if (root < 0) { labels[i] = 0; counts[i] = 0; continue; }
    const long base = i / hw * hw;
    const int a = counts[base + root];                  // the root's own entry already holds the area (rewritten with itself)
    counts[i] = a;
    labels[i] = root + 1;
  }
}

// labels / counts: int32 [n][H][W] outputs (labels doubles as the union-find parent array while the kernels run)
int connected_components(const uint8_t* mask, int* labels, int* counts, int n, int H, int W, cudaStream_t stream) {
  if (n <= 0 || H <= 0 || W <= 0 || (long)H * W > (1L << 30)) { set_error("connected_components: bad shape n=%d H=%d W=%d", n, H, W); return 1; }
  const long total = (long)n * H * W;
  long blocks = (total + 255) / 256;
  const long cap = (long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  const int hw = H * W;
  {
    ProfRange pr(stream, "cc_init", 0, (double)total * 9.0);
    cc_init_kernel<<<(unsigned)blocks, 256, 0, stream>>>(mask, labels, counts, total, hw);
    SAM2B_CHECK_LAUNCH("cc_init");
  }
  {
    ProfRange pr(stream, "cc_merge", 0, (double)total * 9.0);
    cc_merge_kernel<<<(unsigned)blocks, 256, 0, stream>>>(mask, labels, total, H, W);
    SAM2B_CHECK_LAUNCH("cc_merge");
  }
  {
    ProfRange pr(stream, "cc_count", 0, (double)total * 12.0);
    cc_count_kernel<<<(unsigned)blocks, 256, 0, stream>>>(labels, counts, total, hw);
    SAM2B_CHECK_LAUNCH("cc_count");
  }
  {
    ProfRange pr(stream, "cc_finalize", 0, (double)total * 16.0);
    cc_finalize_kernel<<<(unsigned)blocks, 256, 0, stream>>>(labels, counts, total, hw);
    SAM2B_CHECK_LAUNCH("cc_finalize");
  }
  return 0;
}

}  // namespace sam2b
