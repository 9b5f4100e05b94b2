"""Dev tool (not a test): hd72 attention timings per shape (CUDA events, 10 iterations)."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops


def timeit(f, it=10):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


quick = "--quick" in sys.argv        # for ncu: the two tcgen05 shapes at batch 16 only
g = torch.Generator(device="cuda").manual_seed(0)
shapes = [(64, 8, 16, False), (64, 8, 64, False), (256, 2, 8, False), (128, 4, 4, False), (32, 16, 8, False)]
for (H, heads, ws, pool) in shapes[:2] if quick else shapes:
    for B in ((16,) if quick else (1, 2, 4, 8, 16)):
        qkv = torch.randn(B, H, H, 3 * heads * 72, device="cuda", generator=g).bfloat16()
        ms = timeit(lambda: ops.window_attention(qkv, heads, ws, pool), 1 if quick else 10)
        keys = ws * ws
        fl = 4.0 * B * H * H * heads * keys * 72
        by = B * H * H * heads * 72 * 2 * 4
        print(f"H={H} heads={heads} ws={ws} B={B:2d}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TF/s  {by/ms/1e6:7.1f} GB/s", flush=True)
