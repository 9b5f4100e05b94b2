#!/usr/bin/env python
"""One small call of every boundary (E1 image encoder on one image, E3 memory attention with L = 2 / P = 8, E4 memory encoder)
for `compute-sanitizer` (SURVEY.md section 5 "race detection / sanitizers"):

    compute-sanitizer --tool memcheck  python tools/sanitize_run.py
    compute-sanitizer --tool racecheck python tools/sanitize_run.py

Inputs are tiny on purpose: every kernel of the three sub-graphs launches at least once, and the tools slow a launch 10-100x."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200.engine import Engine, launch_count  # noqa: E402
from sam2_opt_b200.synth import synth_image_norm, synth_memattn_inputs, synth_memenc_inputs  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402
from sam2_opt_b200 import ops, postprocess  # noqa: E402

which = sys.argv[1:] or ["e1", "e3", "e4", "ops"]
dev = torch.device("cuda:0")
eng = Engine(make_weights(0), dev)
l0 = launch_count()
if "e1" in which:
    outs = eng.image_encode(synth_image_norm(1, seed=7).to(dev), True, False)
    torch.cuda.synchronize()
    print("E1 ok", [float(o.abs().mean()) for o in outs])
if "e3" in which:
    ins = [t.to(dev) for t in synth_memattn_inputs(2, 8, 1, seed=3)]
    out = eng.memory_attention(*ins)
    torch.cuda.synchronize()
    print("E3 ok", float(out.abs().mean()))
if "e4" in which:
    pix, msk = synth_memenc_inputs(1, seed=4)
    x, pos = eng.memory_encode(pix.to(dev), msk.to(dev))
    torch.cuda.synchronize()
    print("E4 ok", float(x.abs().mean()))
if "ops" in which:
    u8 = torch.randint(0, 256, (1, 1024, 1024, 3), dtype=torch.uint8, device=dev)
    ops.preprocess_u8(u8, normalise=True)
    low = torch.randn(2, 1, 256, 256, device=dev)
    postprocess.upsample_masks(low, (540, 960))
    postprocess.fill_holes_in_mask_scores(low, 8)
    torch.cuda.synchronize()
    print("ops ok")
print("launches", launch_count() - l0)
eng.release()
