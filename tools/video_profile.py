#!/usr/bin/env python
"""Dev tool: where a steady-state tracked frame spends its time.  Runs propagate_in_video on a synthetic clip and, for
frames >= 17 (full 7-frame bank, 16 pointers), collects the library's per-launch CUDA-event profile; prints kernel time per
family per frame next to the wall time per frame (the difference is torch-side work: SAM heads graph, bank assembly, gaps).

    python tools/video_profile.py [objects] [frames]
"""
import collections
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200 import engine as E  # noqa: E402
from sam2_opt_b200.predictor import SAM2VideoPredictor  # noqa: E402
from sam2_opt_b200.synth import synth_video  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

nobj = int(sys.argv[1]) if len(sys.argv) > 1 else 1
T = int(sys.argv[2]) if len(sys.argv) > 2 else 40
g = max(1, int(nobj ** 0.5 + 0.999))
clicks = [[512.0, 512.0]] if nobj == 1 else [[128 + 768.0 * (o % g) / max(1, g - 1), 128 + 768.0 * (o // g) / max(1, g - 1)] for o in range(nobj)]
vp = SAM2VideoPredictor(make_weights(0), device="cuda:0", native_heads=os.environ.get("VP_NATIVE", "1") == "1", graph_track=False)
clip = synth_video(T, seed=1).pin_memory()


def run(profile_from=None):
    st = vp.init_state(clip)
    for o in range(nobj):
        vp.add_new_points_or_box(st, 0, o, points=[clicks[o]], labels=[1])
    stamps = [time.perf_counter()]
    rows = None
    for t, ids, masks in vp.propagate_in_video(st):
        (masks > 0).any().item()
        stamps.append(time.perf_counter())
        if profile_from is not None and t == profile_from - 1:
            torch.cuda.synchronize()
            E.profile_begin()
    if profile_from is not None:
        rows = E.profile_end()
    return stamps, rows


run()
stamps, _ = run()
per = [stamps[i + 1] - stamps[i] for i in range(len(stamps) - 1)]
print(f"{nobj} object(s), {T} frames: wall ms per frame, blocks of 8:", [round(1e3 * sum(per[i:i + 8]) / len(per[i:i + 8]), 2) for i in range(0, T, 8)])
s0 = 17
stamps, rows = run(profile_from=s0)
nfr = T - s0
fam = collections.OrderedDict()
for name, ms, fl, by in rows:
    k = name.split(" ")[0]
    if k == "gemm":
        k = "gemm " + " ".join(name.split(" ")[1:4])
    d = fam.setdefault(k, [0, 0.0, 0.0])
    d[0] += 1; d[1] += ms; d[2] += fl
tot = sum(v[1] for v in fam.values())
print(f"library kernel time per frame (frames {s0}..{T - 1}, profiling on): {tot / nfr:.3f} ms; wall per frame in this pass {1e3 * (stamps[-1] - stamps[s0]) / nfr:.3f} ms")
for k, (n, ms, fl) in sorted(fam.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:60s} {n / nfr:7.1f} launches/frame {ms / nfr:8.3f} ms/frame  {fl / max(ms, 1e-9) / 1e9:8.1f} TF/s")
vp.release()
