"""Dev tool (not a test): per-launch CUDA-event profile of one memory-encoder call (n objects)."""
import collections
import sys

import torch

sys.path.insert(0, "/root/repo")
from oracle import sam2_oracle as O
from sam2_opt_b200 import engine as E

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
e = E.Engine(O.make_weights(0))
pix = torch.randn(n, 256, 64, 64, device="cuda")
m = torch.randn(n, 1, 1024, 1024, device="cuda")
for _ in range(2):
    e.memory_encode(pix, m)
torch.cuda.synchronize()
E.profile_begin()
e.memory_encode(pix, m)
rows = E.profile_end()
agg = collections.OrderedDict()
for name, ms, fl, by in rows:
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += ms; a[2] += fl; a[3] += by
tot = sum(a[1] for a in agg.values())
print(f"n={n} total kernel time {tot:.3f} ms, {len(rows)} launches")
for name, (cnt, ms, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{ms:9.3f} ms {ms/tot:6.1%}  x{cnt:<3d} {fl/ms/1e9 if ms else 0:8.1f} TF/s {by/ms/1e6 if ms else 0:8.1f} GB/s  {name}")
