"""Not a test (run by hand on the GPU box): sam2b_op_upsample_masks against torch's F.interpolate (+ threshold) on the video
predictor's per-frame load, CUDA events on the current stream, inputs + outputs (64 x 1024^2 x 5 B = 336 MB) larger than L2.
Algorithmic bytes per output pixel: 4 (logit) + 1 (mask); the 256^2 source is read once from HBM (0.25 B per output pixel)."""
import json
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
from sam2_opt_b200 import postprocess as PP  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


res = []
for n, hw in ((64, (1024, 1024)), (64, (720, 1280)), (16, (480, 854))):
    x = torch.randn(n, 1, 256, 256, device="cuda") * 8
    px = n * hw[0] * hw[1]
    t_both = timed(lambda: PP.upsample_masks(x, hw, threshold=0.0))
    t_mask = timed(lambda: PP.upsample_masks(x, hw, threshold=0.0, return_logits=False))
    t_torch = timed(lambda: F.interpolate(x, hw, mode="bilinear", align_corners=False) > 0)
    res.append({"masks": n, "out_hw": hw, "ms_logits_and_mask": t_both, "GBps_logits_and_mask": (px * 5 + x.numel() * 4) / t_both / 1e6,
                "ms_mask_only": t_mask, "GBps_mask_only": (px * 1 + x.numel() * 4) / t_mask / 1e6,
                "ms_torch_interpolate_then_gt": t_torch})
print(json.dumps({"upsample_masks": res}))
