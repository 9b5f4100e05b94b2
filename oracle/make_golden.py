"""Generate tests/golden/*.npz by running the UNMODIFIED reference (FasterProcess/sam2_opt,
/root/reference) torch fp32 path on CPU with the oracle's synthetic weights.

    python -m oracle.make_golden            # needs /root/reference (build container only)

The reference holds no golden vectors for this path, so these fixtures are
what pins the oracle: every array below is produced by reference code (module forward /
predictor API), never by ``oracle/sam2_oracle.py``.  The script also asserts that the oracle
reproduces each of them to fp32 round-off at generation time.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import reference_loader as rl  # noqa: E402
from oracle import sam2_oracle as O  # noqa: E402
from oracle.synth import (synth_image_u8, synth_image_norm, synth_video, synth_memattn_inputs,  # noqa: E402
                          synth_memenc_inputs, checksum)

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def sl(x, *idx):
    return x[idx].contiguous().numpy().astype(np.float32)


def stats(x):
    x = x.double()
    return np.array([x.mean().item(), x.std().item(), x.abs().max().item(), x.norm().item()], dtype=np.float64)


def main():
    torch.set_grad_enabled(False)
    os.makedirs(OUT, exist_ok=True)
    w = O.make_weights(0)
    m = rl.build_reference("video")
    m.load_state_dict(w, strict=True)          # proves names/shapes == reference state_dict
    g = {}
    g["weights_checksum"] = np.array([checksum(w["image_encoder.trunk.blocks.9.attn.qkv.weight"]),
                                      checksum(w["memory_attention.layers.2.linear1.weight"])])

    # ---- E2: SAM2Base.forward_image (ref sam2_base_official.py:548-582) -------------------
    img = synth_image_norm(1, 1)
    g["e2_input_checksum"] = np.array([checksum(img)])
    # per-block probes through the reference trunk (hieradet.py:283-299)
    trunk = m.image_encoder.trunk
    x = trunk.patch_embed(img)
    x = x + trunk._get_pos_embed(x.shape[1:3])
    probes, probe_stats = [sl(x, 0, slice(None, None, 64), slice(None, None, 64), slice(0, 8))], [stats(x)]
    for blk in trunk.blocks:
        x = blk(x)
        s = max(1, x.shape[1] // 4)
        probes.append(sl(x, 0, slice(None, None, s), slice(None, None, s), slice(0, 8)))
        probe_stats.append(stats(x))
    for i, p in enumerate(probes):
        g[f"e2_block_probe_{i}"] = p
    g["e2_block_stats"] = np.stack(probe_stats)
    ref = m.forward_image(img)
    outs = [ref["vision_features"], *ref["vision_pos_enc"], *ref["backbone_fpn"]]
    mine = O.forward_image(w, img)
    for i, (a, b) in enumerate(zip(outs, mine)):
        assert (a - b).abs().max().item() <= 1e-5 * max(1.0, a.abs().max().item()), i
        s = a.shape[-1] // 16
        g[f"e2_out{i}"] = sl(a, 0, slice(0, 16), slice(None, None, s), slice(None, None, s))
        g[f"e2_out{i}_stats"] = stats(a)

    # ---- E1: SAM2ImagePredictor.set_image_e2e + config 1 predict --------------------------
    from sam2.sam2_image_predictor import SAM2ImagePredictor
    img_u8 = synth_image_u8(0)
    g["e1_input_checksum"] = np.array([checksum(torch.as_tensor(img_u8).double())])
    pred = SAM2ImagePredictor(m)
    pred.set_image(img_u8)
    feats = [*pred._features["high_res_feats"], pred._features["image_embed"]]
    mine = O.set_image_e2e(w, torch.as_tensor(img_u8).permute(2, 0, 1)[None].float() / 255.0)
    for i, (a, b) in enumerate(zip(feats, mine)):
        assert (a - b).abs().max().item() <= 1e-5 * max(1.0, a.abs().max().item()), i
        s = a.shape[-1] // 16
        g[f"e1_out{i}"] = sl(a, 0, slice(0, 16), slice(None, None, s), slice(None, None, s))
        g[f"e1_out{i}_stats"] = stats(a)
    masks, iou, low = pred.predict(point_coords=np.array([[512, 512]]), point_labels=np.array([1]),
                                   multimask_output=True, return_logits=True)
    mm, mi, ml = O.image_predict(w, img_u8, [[512.0, 512.0]], [1], True)
    assert np.abs(low - ml.numpy()).max() < 1e-5 and np.abs(iou - mi.numpy()).max() < 1e-5
    g["c1_low_res_bits"] = np.packbits(low > 0)
    g["c1_low_res_slice"] = low[:, ::16, ::16].astype(np.float32)
    g["c1_iou"] = iou.astype(np.float32)
    g["c1_mask_area"] = np.array([(masks > 0).sum(axis=(1, 2))]).astype(np.int64)

    # ---- E3: MemoryAttention (ref memory_attention.py:263-349) ----------------------------
    for tag, (L, P, N) in {"a": (2, 8, 1), "b": (1, 0, 1), "c": (3, 64, 2)}.items():
        ins = synth_memattn_inputs(L, P, N, seed=3)
        a = m.memory_attention.inference_memory_attention_torch(*ins)
        b = O.memory_attention(w, *ins)
        assert (a - b).abs().max().item() < 1e-4
        g[f"e3{tag}_input_checksum"] = np.array([checksum(t) for t in ins])
        g[f"e3{tag}_out"] = sl(a, slice(None, None, 64), slice(None), slice(None, None, 8))
        g[f"e3{tag}_stats"] = stats(a)

    # ---- E4: MemoryEncoder (ref memory_encoder.py:233-241) --------------------------------
    pix, msk = synth_memenc_inputs(2, seed=4)
    a, apos = m.memory_encoder.inference_memory_torch(pix, msk)
    b, bpos = O.memory_encoder(w, pix, msk)
    assert (a - b).abs().max().item() < 1e-5 and (apos - bpos).abs().max().item() < 1e-6
    g["e4_input_checksum"] = np.array([checksum(pix), checksum(msk)])
    g["e4_out"] = sl(a, slice(None), slice(None, None, 4), slice(None, None, 8), slice(None, None, 8))
    g["e4_pos"] = sl(apos, 0, slice(None, None, 4), slice(None, None, 8), slice(None, None, 8))
    g["e4_stats"] = stats(a)

    # ---- config 3 (short): SAM2VideoPredictor.propagate_in_video --------------------------
    import sam2.sam2_video_predictor_official as vp
    T = 4
    frames = synth_video(T, seed=1)
    g["c3_input_checksum"] = np.array([checksum(frames)])
    vp.load_video_frames = lambda **k: (frames, 1024, 1024)
    st = m.init_state("synthetic")
    m.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
    ref_low = {}
    for fi, _, _ in m.propagate_in_video(st):
        d = st["output_dict_per_obj"][0]
        ref_low[fi] = d["cond_frame_outputs" if fi == 0 else "non_cond_frame_outputs"][fi]["pred_masks"].clone()
    for fi, low_t in O.video_propagate(w, frames):
        assert (low_t - ref_low[fi]).abs().max().item() < 1e-4, fi
        assert O.mask_iou(low_t, ref_low[fi]) > 0.9999
    g["c3_low_res_bits"] = np.stack([np.packbits((ref_low[t][0, 0] > 0).numpy()) for t in range(T)])
    g["c3_low_res_slice"] = np.stack([ref_low[t][0, 0, ::16, ::16].numpy() for t in range(T)]).astype(np.float32)

    path = os.path.join(OUT, "hiera_l_seed0.npz")
    np.savez_compressed(path, **g)
    print("wrote", path, os.path.getsize(path) / 1e3, "KB;", len(g), "arrays")


if __name__ == "__main__":
    main()
