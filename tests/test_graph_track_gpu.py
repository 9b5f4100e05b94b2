"""GPU: the tracked frame replayed from one CUDA graph (predictor._GraphedTrack; with two objects the
graph also serves the frames where the bank is still filling: E3 at full size with device-side counts) against the same
predictor enqueuing every call (graph_track=False).  The two differ only in the order of the memory-bank slots,
i.e. in fp32 summation order inside the attention: per-frame mask IoU >= 0.998 (measured 0.9994; a real slot / age mix-up gave
0.977), object pointers rel-L2 <= 1e-3 (measured 1e-4).

Also covered: a workspace re-allocation between replays (a larger problem arrives: the captured addresses go stale and the
graph must be captured again) and a correction click followed by a second propagation (the tracker starts over)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def weights():
    from oracle import sam2_oracle as O
    return O.make_weights(0)


def _iou(a, b):
    a, b = a > 0, b > 0
    u = (a | b).sum().item()
    return 1.0 if u == 0 else (a & b).sum().item() / u


def _run(vp, frames, clicks, poke=None):
    st = vp.init_state(frames)
    for o, (x, y) in clicks.items():
        vp.add_new_points_or_box(st, 0, o, points=np.array([[x, y]], dtype=np.float32), labels=np.array([1], np.int32))
    for t, ids, masks in vp.propagate_in_video(st):
        if poke is not None and t == 24:
            poke()
    return st


def test_graph_replay_matches_enqueued_frames(weights):
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    torch.set_grad_enabled(False)
    T = 40
    frames = synth_video(T, seed=5)
    clicks = {1: (400.0, 520.0), 2: (700.0, 300.0)}
    vp = SAM2VideoPredictor(weights, graph_track=True)
    assert vp.graph_track

    def poke():   # a 24-prompt decoder call grows the SAM-heads workspaces while the tracker's graph is alive
        pix = torch.randn(1, 256, 64, 64, device=vp.device)
        hr = [torch.randn(1, 32, 256, 256, device=vp.device), torch.randn(1, 64, 128, 128, device=vp.device)]
        pts = torch.rand(24, 1, 2, device=vp.device) * 1024
        vp.engine.sam_heads(pix.expand(24, -1, -1, -1), hr, pts, torch.ones(24, 1, dtype=torch.int32, device=vp.device),
                            multimask_output=True, decoder_only=True, high_res=False)

    gen0 = vp.engine.workspace_generation()
    st_g = _run(vp, frames, clicks, poke)
    assert vp.engine.workspace_generation() > gen0
    trackers = list(st_g["gtrack"].values())
    assert len(trackers) == 1 and trackers[0].served == T - 1, "the graph path did not run from the first tracked frame on"
    vp_e = SAM2VideoPredictor(weights, graph_track=False)
    st_e = _run(vp_e, frames, clicks)
    worst, worst_ptr = 1.0, 0.0
    for o in clicks:
        for t in range(1, T):
            a, b = st_g["non_cond"][o][t], st_e["non_cond"][o][t]
            worst = min(worst, _iou(a["pred_masks"], b["pred_masks"]))
            worst_ptr = max(worst_ptr, ((a["obj_ptr"] - b["obj_ptr"]).norm() / b["obj_ptr"].norm()).item())
    print(f"graph vs enqueued: worst mask IoU {worst:.5f}, worst pointer rel-L2 {worst_ptr:.2e}")
    assert worst >= 0.998 and worst_ptr <= 1e-3

    # a correction click on a tracked frame, then a second propagation: both predictors must still agree
    for vp_, st_ in ((vp, st_g), (vp_e, st_e)):
        vp_.add_new_points_or_box(st_, 20, 1, points=np.array([[410.0, 500.0]], dtype=np.float32), labels=np.array([1], np.int32),
                                  clear_old_points=False)
        list(vp_.propagate_in_video(st_))
    worst = min(_iou(st_g["non_cond"][o][t]["pred_masks"], st_e["non_cond"][o][t]["pred_masks"]) for o in clicks for t in range(21, T))
    print(f"after a correction click: worst mask IoU {worst:.5f}")
    assert worst >= 0.998
    vp.release()
    vp_e.release()
