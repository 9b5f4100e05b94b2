"""Dev tool: torch-profiler breakdown of propagate_in_video with 16 objects."""
import sys
import torch
sys.path.insert(0, "/root/repo")
from torch.profiler import profile, ProfilerActivity
from sam2_opt_b200.predictor import SAM2VideoPredictor
from sam2_opt_b200.synth import synth_video
from sam2_opt_b200.synth_weights import make_weights
T, NOBJ = 32, int(sys.argv[1]) if len(sys.argv) > 1 else 16
vp = SAM2VideoPredictor(make_weights(0))
frames = synth_video(T, seed=1).pin_memory()
def run():
    st = vp.init_state(frames)
    for o in range(NOBJ):
        vp.add_new_points_or_box(st, 0, o, points=[[128 + 768 * (o % 4) / 3, 128 + 768 * (o // 4) / 3]], labels=[1])
    for t, ids, masks in vp.propagate_in_video(st):
        (masks > 0).any().item()
run()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as p:
    run()
print(p.key_averages().table(sort_by="self_cuda_time_total", row_limit=32, max_name_column_width=70))
print(p.key_averages().table(sort_by="self_cpu_time_total", row_limit=18, max_name_column_width=60))
