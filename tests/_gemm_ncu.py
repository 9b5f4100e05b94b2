"""Dev tool (not a test): a handful of GEMM launches on the stage-3 shapes for an `ncu --set full` capture."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

bn = int(sys.argv[1]) if len(sys.argv) > 1 else 0
M = 65536
g = torch.Generator(device="cuda").manual_seed(0)
for N, K, act, res, out in [(1728, 576, None, False, torch.bfloat16), (2304, 576, "gelu", False, torch.bfloat16),
                            (576, 2304, None, True, torch.float32), (576, 576, None, True, torch.float32)]:
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    r = torch.randn(M, N, device="cuda", generator=g) if res else None
    for _ in range(3):
        ops.gemm(a, b, bias=bias, act=act, residual=r, out_dtype=out, block_n=bn)
    torch.cuda.synchronize()
print("ok")
