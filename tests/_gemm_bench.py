"""Dev tool (not a test): CUDA-event timing of the tcgen05 GEMM on the trunk's shapes, per BLOCK_N and epilogue."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

PEAK = 1396.5
n_img = int(sys.argv[1]) if len(sys.argv) > 1 else 16
bns = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0]
# (name, M/img, N, K, act, residual, out)
CASES = [("s3.qkv", 4096, 1728, 576, None, False, "bf16"), ("s3.proj", 4096, 576, 576, None, True, "f32"),
         ("s3.fc1", 4096, 2304, 576, "gelu", False, "bf16"), ("s3.fc2", 4096, 576, 2304, None, True, "f32"),
         ("s2.qkv", 16384, 864, 288, None, False, "bf16"), ("s2.fc1", 16384, 1152, 288, "gelu", False, "bf16"),
         ("s2.fc2", 16384, 288, 1152, None, True, "f32"), ("s1.qkv", 65536, 432, 144, None, False, "bf16"),
         ("s1.fc1", 65536, 576, 144, "gelu", False, "bf16"), ("s1.fc2", 65536, 144, 576, None, True, "f32"),
         ("s4.qkv", 1024, 3456, 1152, None, False, "bf16"), ("s4.fc1", 1024, 4608, 1152, "gelu", False, "bf16"),
         ("s4.fc2", 1024, 1152, 4608, None, True, "f32")]
g = torch.Generator(device="cuda").manual_seed(0)
for name, m, N, K, act, res, out in CASES:
    M = m * n_img
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    r = torch.randn(M, N, device="cuda", generator=g) if res else None
    for bn in bns:
        try:
            f = lambda: ops.gemm(a, b, bias=bias, act=act, residual=r, out_dtype=torch.bfloat16 if out == "bf16" else torch.float32, block_n=bn)
            for _ in range(2):
                f()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                f()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 5
            tf = 2.0 * M * N * K / ms / 1e9
            print(f"{name:8s} M={M:7d} N={N:5d} K={K:5d} bn={bn:3d} {ms:8.3f} ms {tf:7.1f} TF/s {tf / PEAK:6.1%}", flush=True)
        except Exception as ex:  # noqa: BLE001
            print(name, bn, "failed", ex)
    del a, b, r
