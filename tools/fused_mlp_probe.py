#!/usr/bin/env python
"""Dev tool: the stage-1/2 MLP kernels of one batch-B encoder pass with the two-GEMM schedule and with mlp_fused
(SAM2B_FUSED_MLP read per call), per-launch CUDA-event times from the library's profiler; plus the whole pass.
    python tools/fused_mlp_probe.py [batch]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200 import engine as E  # noqa: E402
from sam2_opt_b200.synth import synth_image_norm  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
eng = E.Engine(make_weights(0), torch.device("cuda:0"))
img = synth_image_norm(B, seed=1).cuda()
outs = {}
import statistics
times = {"0": [], "1": []}
for mode in ("0", "1"):
    os.environ["SAM2B_FUSED_MLP"] = mode
    for _ in range(2):
        outs[mode] = [t.clone() for t in eng.image_encode(img, True, False)]
torch.cuda.synchronize()
for rep in range(8):                     # interleaved A / B: the clocks drift under the power cap, medians over alternating runs
    for mode in ("0", "1"):
        os.environ["SAM2B_FUSED_MLP"] = mode
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(2):
            eng.image_encode(img, True, False)
        z.record()
        torch.cuda.synchronize()
        times[mode].append(a.elapsed_time(z) / 2)
for mode in ("0", "1"):
    os.environ["SAM2B_FUSED_MLP"] = mode
    ms = statistics.median(times[mode])
    print(f"SAM2B_FUSED_MLP={mode}: median {ms:.2f} ms per {B}-image pass ({B / ms * 1e3:.1f} images/s); runs {[round(x, 1) for x in times[mode]]}")
    E.profile_begin()
    eng.image_encode(img, True, False)
    rows = E.profile_end()
    agg = {}
    for name, ms_, fl, by in rows:
        if name.startswith("mlp_fused") or any(s_ in name for s_ in (" n576 k144", " n144 k576", " n1152 k288", " n288 k1152")):
            d = agg.setdefault(name, [0, 0.0])
            d[0] += 1; d[1] += ms_
    for k, (n, ms_) in agg.items():
        print(f"   {k:40s} x{n}  {ms_ / n:.3f} ms each")
r = [((a.double() - b.double()).norm() / b.double().norm()).item() for a, b in zip(outs["1"], outs["0"])]
print("fused vs two-GEMM outputs rel-L2:", [f"{x:.2e}" for x in r])
eng.release()
