"""Dev tool (not a test): the tcgen05 GEMM against torch.matmul (cuBLASLt) on the stage-3 shapes, bf16 in / bf16 out, no
epilogue work on either side, back to back in one process (same clocks / power state)."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 16


def timeit(f, it=20):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


g = torch.Generator(device="cuda").manual_seed(0)
for name, m, N, K in [("s3.qkv", 4096, 1728, 576), ("s3.fc1", 4096, 2304, 576), ("s3.fc2", 4096, 576, 2304), ("s3.proj", 4096, 576, 576),
                      ("s4.fc1", 1024, 4608, 1152), ("big", 4096, 4096, 4096)]:
    M = m * n_img
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bt = b.t()
    ours = timeit(lambda: ops.gemm(a, b, out_dtype=torch.bfloat16))
    ref = timeit(lambda: torch.matmul(a, bt))
    fl = 2.0 * M * N * K
    print(f"{name:8s} M={M:7d} N={N:5d} K={K:5d}: ours {ours*1e3:8.1f} us {fl/ours/1e9:7.1f} TF/s | cuBLASLt {ref*1e3:8.1f} us {fl/ref/1e9:7.1f} TF/s", flush=True)
