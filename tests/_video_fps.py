"""Dev tool (not a test): frames/s of SAM2VideoPredictor.propagate_in_video (BASELINE.json configs 3 and 4 shapes)."""
import sys
import time

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200.predictor import SAM2VideoPredictor
from sam2_opt_b200.synth import synth_video
from sam2_opt_b200.synth_weights import make_weights

T = int(sys.argv[1]) if len(sys.argv) > 1 else 64
import os
vp = SAM2VideoPredictor(make_weights(0), encode_ahead=int(os.environ.get('VP_AHEAD', '16')),
                        native_heads=os.environ.get('VP_NATIVE', '1') == '1', graph_track=os.environ.get('VP_GRAPH', '1') == '1')
frames = synth_video(T, seed=1).pin_memory()


def run(n_obj, prof=False):
    st = vp.init_state(frames)
    g = int(n_obj ** 0.5 + 0.999)
    for o in range(n_obj):
        x, y = (512, 512) if n_obj == 1 else (128 + 768 * (o % g) / max(1, g - 1), 128 + 768 * (o // g) / max(1, g - 1))
        vp.add_new_points_or_box(st, 0, o, points=[[x, y]], labels=[1])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    stamps = []
    for t, ids, masks in vp.propagate_in_video(st):
        (masks > 0).any().item()         # the consumer reads every frame's result
        stamps.append(time.perf_counter())
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    steady = (stamps[-1] - stamps[16]) / (len(stamps) - 17) if len(stamps) > 20 else float("nan")
    print(f"objects={n_obj}: {T} frames in {dt:.3f} s -> {T / dt:.2f} frames/s overall, steady (frames>=16) "
          f"{1 / steady:.2f} frames/s = {n_obj / steady:.1f} object-frames/s", flush=True)


for _n in [int(x) for x in os.environ.get('VP_OBJS', '1,1,16').split(',')]:
    run(_n)
if "--prof" in sys.argv:
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as p:
        run(1)
    print(p.key_averages().table(sort_by="self_cuda_time_total", row_limit=25, max_name_column_width=60))
    print(p.key_averages().table(sort_by="self_cpu_time_total", row_limit=15, max_name_column_width=60))
