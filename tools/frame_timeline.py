#!/usr/bin/env python
"""Dev tool: is a steady-state tracked frame bound by the host or by the GPU?  Profiles a few frames of propagate_in_video with
the default predictor (graphed SAM heads, bank ring) and prints, per frame: wall time, GPU busy time (union of the kernel /
memcpy intervals), the host time spent inside CUDA launch calls and the launch counts.

    python tools/frame_timeline.py [objects] [frames_profiled]
"""
import os
import sys
import time

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200.predictor import SAM2VideoPredictor  # noqa: E402
from sam2_opt_b200.synth import synth_video  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

nobj = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nprof = int(sys.argv[2]) if len(sys.argv) > 2 else 8
g = max(1, int(nobj ** 0.5 + 0.999))
clicks = [[512.0, 512.0]] if nobj == 1 else [[128 + 768.0 * (o % g) / max(1, g - 1), 128 + 768.0 * (o // g) / max(1, g - 1)] for o in range(nobj)]
vp = SAM2VideoPredictor(make_weights(0), device="cuda:0", graph_track=os.environ.get("VP_GRAPH", "1") == "1")
# 17 warm frames, then the profiled ones start right after an encode-ahead block boundary (frame 32) so that no image
# encoding falls inside the window
clip = synth_video(33 + nprof, seed=1).pin_memory()
st = vp.init_state(clip)
for o in range(nobj):
    vp.add_new_points_or_box(st, 0, o, points=[clicks[o]], labels=[1])
gen = vp.propagate_in_video(st)
for _ in range(33):
    t, ids, masks = next(gen)
    (masks > 0).any().item()
torch.cuda.synchronize()
t0 = time.perf_counter()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(nprof):
        t, ids, masks = next(gen)
        (masks > 0).any().item()
    torch.cuda.synchronize()
wall = (time.perf_counter() - t0) / nprof * 1e3
iv, launches, launch_us, graphs, cpu_ops = [], 0, 0.0, 0, {}
for e in prof.events():
    if e.device_type.name == "CUDA":
        iv.append((e.time_range.start, e.time_range.end, e.name))
    elif e.name.startswith(("cudaLaunchKernel", "cuLaunchKernel", "cudaMemcpyAsync", "cudaMemsetAsync")):
        launches += 1
        launch_us += e.cpu_time_total
    elif e.name.startswith("cudaGraphLaunch"):
        graphs += 1
        launch_us += e.cpu_time_total
    elif e.name.startswith("aten::") and e.cpu_parent is None:
        cpu_ops[e.name] = cpu_ops.get(e.name, 0) + 1
iv.sort()
busy, end = 0.0, -1.0
for s, e_, _ in iv:
    if e_ <= end:
        continue
    busy += e_ - max(s, end)
    end = e_
print(f"{nobj} object(s), frames {t - nprof + 1}..{t} (profiler on): wall {wall:.3f} ms/frame, GPU busy {busy / nprof / 1e3:.3f} ms/frame, "
      f"{len(iv) / nprof:.0f} device activities/frame, {launches / nprof:.0f} launch calls + {graphs / nprof:.1f} graph launches per frame "
      f"({launch_us / nprof / 1e3:.3f} ms of host time inside them)")
by = {}
for s, e_, nm in iv:
    k = nm[:90]
    a = by.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += e_ - s
for k, (c, us) in sorted(by.items(), key=lambda kv: -kv[1][1])[:32]:
    print(f"  {us / nprof:9.1f} us  x{c / nprof:6.1f}  {k}")
print("top-level aten ops per frame:", ", ".join(f"{k[6:]} {v / nprof:.1f}" for k, v in sorted(cpu_ops.items(), key=lambda kv: -kv[1])[:24]))
vp.release()
