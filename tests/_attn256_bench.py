"""Dev tool (not a test): hd256 attention (memory attention core) timings per shape (CUDA events)."""
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


quick = "--quick" in sys.argv
ns = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 and sys.argv[1][0].isdigit() else ((2,) if quick else (1, 4, 16))
g = torch.Generator(device="cuda").manual_seed(0)
for lk in (4096, 28736):
    for n in ns:
        q = torch.randn(n, 4096, 256, device="cuda", generator=g).bfloat16()
        k = torch.randn(n, lk, 256, device="cuda", generator=g).bfloat16()
        v = torch.randn(n, lk, 256, device="cuda", generator=g).bfloat16()
        ms = timeit(lambda: ops.attention_hd256(q, k, v), 2 if quick else 10)
        fl = 4.0 * n * 4096 * lk * 256
        print(f"n={n:2d} Lk={lk:5d}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TF/s", flush=True)
