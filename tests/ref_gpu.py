"""Test helper: the UNMODIFIED reference (``/root/reference`` here, the ``oracle/_ref`` copy on the GPU box) instantiated on
the GPU in strict fp32 — the "fast oracle" of SURVEY.md §8c for the configurations the CPU oracle cannot finish in seconds.
TF32 is switched off for matmul AND cuDNN (cuDNN's defaults to on and would silently degrade the conv layers)."""
import contextlib

import numpy as np
import torch

from oracle import reference_loader as rl


def strict_fp32():
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    torch.set_float32_matmul_precision("highest")


def build(kind, weights, device="cuda"):
    """kind: 'base' (for SAM2ImagePredictor) or 'video' (SAM2VideoPredictor); strict load of the synthetic weights."""
    strict_fp32()
    m = rl.build_reference(kind)
    m.load_state_dict(weights, strict=True)
    return m.to(device).eval()


@contextlib.contextmanager
def synthetic_video(frames, height=1024, width=1024):
    """Make the reference's `init_state(video_path=...)` return `frames` [T,3,1024,1024] (normalised fp32, what its loader
    produces, utils/misc.py:213-277) instead of reading JPEGs."""
    import sam2.sam2_video_predictor_official as vp
    old = vp.load_video_frames
    vp.load_video_frames = lambda **k: (frames, height, width)
    try:
        yield
    finally:
        vp.load_video_frames = old


def propagate_low_res(model, frames, clicks, max_track=None):
    """Run the reference video predictor: `clicks` = {obj_id: (x, y)} positive clicks on frame 0; returns
    {frame: {obj_id: low-res mask logits [256,256] (cpu)}} read from the predictor's own per-object output dicts.
    max_track: stop after that many tracked frames (the clip length, which enters the pointer encodings, stays the same)."""
    with synthetic_video(frames):
        st = model.init_state("synthetic")
    for oid, (x, y) in clicks.items():
        model.add_new_points_or_box(st, 0, oid, points=np.array([[x, y]], dtype=np.float32), labels=np.array([1], np.int32))
    out = {}
    for fi, obj_ids, video_res in model.propagate_in_video(st, max_frame_num_to_track=max_track):
        out[fi] = {}
        for oid in obj_ids:
            idx = st["obj_id_to_idx"][oid]
            d = st["output_dict_per_obj"][idx]
            rec = d["cond_frame_outputs"].get(fi) or d["non_cond_frame_outputs"][fi]
            out[fi][oid] = rec["pred_masks"][0, 0].float().cpu().clone()
    return out


def iou(a_bool, b_bool):
    a_bool, b_bool = np.asarray(a_bool, dtype=bool), np.asarray(b_bool, dtype=bool)
    u = (a_bool | b_bool).sum()
    return 1.0 if u == 0 else float((a_bool & b_bool).sum()) / float(u)


def propagate_low_res_from_masks(model, frames, masks, max_track=None):
    """Same with MASK prompts on frame 0 (`add_new_mask`, what the reference's tools/vos_inference.py feeds): `masks` =
    {obj_id: bool array [H,W]}; returns ({frame: {obj_id: low-res logits}}, {obj_id: obj_ptr of frame 0})."""
    with synthetic_video(frames):
        st = model.init_state("synthetic")
    for oid, m in masks.items():
        model.add_new_mask(st, 0, oid, m)
    out, ptrs = {}, {}
    for fi, obj_ids, video_res in model.propagate_in_video(st, max_frame_num_to_track=max_track):
        out[fi] = {}
        for oid in obj_ids:
            d = st["output_dict_per_obj"][st["obj_id_to_idx"][oid]]
            rec = d["cond_frame_outputs"].get(fi) or d["non_cond_frame_outputs"][fi]
            out[fi][oid] = rec["pred_masks"][0, 0].float().cpu().clone()
            if fi == 0:
                ptrs[oid] = rec["obj_ptr"][0].float().cpu().clone()
    return out, ptrs
