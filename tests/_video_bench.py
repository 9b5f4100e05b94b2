"""Dev tool (not a test): CUDA-event timing of the video-path sub-graphs E3 (memory attention) and E4 (memory encoder)."""
import collections
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import engine as E
from sam2_opt_b200.synth_weights import make_weights

e = E.Engine(make_weights(0))
dev = e.device


def timeit(f, it=5):
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / it


for n, L, P in [(1, 7, 64), (1, 1, 4), (4, 7, 64), (16, 7, 64)]:
    curr = torch.randn(4096, n, 256, device=dev)
    cpos = torch.randn(4096, n, 256, device=dev)
    mem = torch.randn(L, 4096, n, 64, device=dev)
    mpos = torch.randn(L, 4096, n, 64, device=dev)
    ex = torch.randn(P, n, 64, device=dev)
    expos = torch.randn(P, n, 64, device=dev)
    ms = timeit(lambda: e.memory_attention(curr, mem, cpos, mpos, ex, expos))
    gf = n * (4 * (19.33 + 8.59) + 4 * (2 * 2 * 4096 * (L * 4096 + P) * 256 + 2 * 2 * (L * 4096 + P) * 64 * 256 + 2 * 2 * 4096 * 256 * 256) / 1e9)
    print(f"E3 n={n} L={L} P={P}: {ms:.2f} ms  ({gf / ms:.1f} TF/s algorithmic)")
    if (n, L) == (1, 7):
        E.profile_begin()
        e.memory_attention(curr, mem, cpos, mpos, ex, expos)
        agg = collections.OrderedDict()
        for name, m, fl, by in E.profile_end():
            a = agg.setdefault(name.split(" md")[0], [0, 0.0])
            a[0] += 1; a[1] += m
        for k, (c, m) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
            print(f"    {m:8.3f} ms x{c:<3d} {k}")
for n in (1, 16):
    pix = torch.randn(n, 256, 64, 64, device=dev)
    msk = torch.randn(n, 1, 1024, 1024, device=dev)
    ms = timeit(lambda: e.memory_encode(pix, msk))
    print(f"E4 n={n}: {ms:.3f} ms")
img = torch.randn(1, 3, 1024, 1024, device=dev)
print(f"E2 n=1: {timeit(lambda: e.image_encode(img, True, False)):.3f} ms")
