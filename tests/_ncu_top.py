"""Dev tool (not a test): one launch of each dominant kernel on its real shape, for an `ncu --set full` capture.
Order: stage-3 GEMMs qkv, fc1(GELU), fc2(+res), proj(+res); hd72 attention 16x16 windows and global; hd256 attention."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

M = 65536
g = torch.Generator(device="cuda").manual_seed(0)
for N, K, act, res, out in [(1728, 576, None, False, torch.bfloat16), (2304, 576, "gelu", False, torch.bfloat16),
                            (576, 2304, None, True, torch.float32), (576, 576, None, True, torch.float32)]:
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    r = torch.randn(M, N, device="cuda", generator=g) if res else None
    for _ in range(2):
        ops.gemm(a, b, bias=bias, act=act, residual=r, out_dtype=out)
    torch.cuda.synchronize()
qkv = torch.randn(16, 64, 64, 3 * 576, device="cuda", generator=g).bfloat16()
for ws in (16, 64):
    for _ in range(2):
        ops.window_attention(qkv, 8, ws, False)
    torch.cuda.synchronize()
q = torch.randn(1, 4096, 256, device="cuda", generator=g).bfloat16()
k = torch.randn(1, 28736, 256, device="cuda", generator=g).bfloat16()
v = torch.randn(1, 28736, 256, device="cuda", generator=g).bfloat16()
for _ in range(2):
    ops.attention_hd256(q, k, v)
torch.cuda.synchronize()
print("ok")
