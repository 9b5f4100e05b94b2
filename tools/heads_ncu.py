#!/usr/bin/env python
"""Dev tool: a few sam2b_sam_heads calls (a tracked frame: no prompt, multimask) for `ncu` captures / launch lists.

    python tools/heads_ncu.py [objects] [calls]
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200.engine import Engine  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 3
eng = Engine(make_weights(0), torch.device("cuda:0"))
g = torch.Generator(device="cuda").manual_seed(0)
seq = torch.randn(4096, n, 256, device="cuda", generator=g)
pix = seq.permute(1, 2, 0).reshape(n, 256, 64, 64)                 # the E3 output layout
hr = [0.5 * torch.randn(1, 32, 256, 256, device="cuda", generator=g), 0.5 * torch.randn(1, 64, 128, 128, device="cuda", generator=g)]
ev = [torch.cuda.Event(enable_timing=True) for _ in range(calls + 1)]
ev[0].record()
for i in range(calls):
    out = eng.sam_heads(pix, hr, multimask_output=True)
    ev[i + 1].record()
torch.cuda.synchronize()
print(f"sam2b_sam_heads, {n} objects: ms per call", [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(calls)])
eng.release()
