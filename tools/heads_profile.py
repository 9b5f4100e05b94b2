#!/usr/bin/env python
"""Dev tool: torch-profiler view of ONE steady-state tracked frame (n objects): which kernels outside libsam2b (SAM heads,
memory-bank assembly, mask up-sampling) take the time.   python tools/heads_profile.py [objects]"""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200.predictor import SAM2VideoPredictor  # noqa: E402
from sam2_opt_b200.synth import synth_video  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

nobj = int(sys.argv[1]) if len(sys.argv) > 1 else 16
g = max(1, int(nobj ** 0.5 + 0.999))
clicks = [[512.0, 512.0]] if nobj == 1 else [[128 + 768.0 * (o % g) / max(1, g - 1), 128 + 768.0 * (o // g) / max(1, g - 1)] for o in range(nobj)]
vp = SAM2VideoPredictor(make_weights(0), device="cuda:0", graph_heads=False)     # eager heads so that the profiler sees the ops
clip = synth_video(28, seed=1).pin_memory()
st = vp.init_state(clip)
for o in range(nobj):
    vp.add_new_points_or_box(st, 0, o, points=[clicks[o]], labels=[1])
gen = vp.propagate_in_video(st)
for _ in range(20):
    next(gen)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(4):
        next(gen)
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in rows if e.device_type.name == "CUDA" or e.device_time_total)
print(f"{nobj} objects, 4 frames; device time by kernel (us per frame)")
seen = 0
for e in rows:
    if e.device_time_total <= 0 or not e.key or e.key.startswith(("aten::", "cuda", "Memcpy", "Memset")) and False:
        continue
    print(f"  {e.device_time_total / 4:9.1f} us  x{e.count / 4:6.1f}  {e.key[:110]}")
    seen += 1
    if seen >= 45:
        break
vp.release()
