"""Dev tool (not a test): the memory/epilogue-bound stage-1/2 GEMM forms for an `ncu --set full` capture."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

g = torch.Generator(device="cuda").manual_seed(0)
for M, N, K, act, res, out in [(1048576, 576, 144, "gelu", False, torch.bfloat16), (262144, 1152, 288, "gelu", False, torch.bfloat16),
                               (1048576, 432, 144, None, False, torch.bfloat16), (1048576, 144, 576, None, True, torch.float32)]:
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    r = torch.randn(M, N, device="cuda", generator=g) if res else None
    for _ in range(2):
        ops.gemm(a, b, bias=bias, act=act, residual=r, out_dtype=out)
    torch.cuda.synchronize()
print("ok")
