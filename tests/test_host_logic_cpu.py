"""CPU: host-side logic of the video predictor that runs in plain torch (no engine involved) against the oracle's
restatement of the reference (sam2_base_official.py:825-950): memory-bank / object-pointer assembly, with and without
the per-video table of projected temporal encodings that replaces the reference's per-call host-to-device copy."""
import pytest
import torch

from oracle import sam2_oracle as O
from sam2_opt_b200 import sam_heads as H

torch.set_grad_enabled(False)


def _outputs(g, t_list):
    return {t: {"maskmem_features": torch.randn(1, 64, 64, 64, generator=g).bfloat16(),
                "maskmem_pos": torch.randn(1, 64, 64, 64, generator=g),
                "obj_ptr": torch.randn(1, 256, generator=g)} for t in t_list}


@pytest.fixture(scope="module")
def w():
    keys = ("maskmem_tpos_enc", "obj_ptr_tpos_proj.weight", "obj_ptr_tpos_proj.bias")
    full = O.make_weights(0)
    return {k: full[k].float() for k in keys}


@pytest.mark.parametrize("frame_idx,num_frames", [(1, 64), (5, 64), (20, 64), (40, 64), (3, 4), (9, 10)])
def test_assemble_memory_matches_oracle(w, frame_idx, num_frames):
    g = torch.Generator().manual_seed(frame_idx)
    cond = _outputs(g, [0])
    non_cond = _outputs(g, range(1, frame_idx))
    ref = O.assemble_memory(w, frame_idx, cond, non_cond, num_frames)
    cache = {}
    for kw in ({}, {"cache": cache}, {"cache": cache}):          # plain, table built, table reused
        got = H.assemble_memory(w, frame_idx, cond, non_cond, num_frames, **kw)
        assert [tuple(x.shape) for x in got] == [tuple(x.shape) for x in ref]
        for a, b in zip(got, ref):
            torch.testing.assert_close(a.float(), b.float(), rtol=1e-5, atol=1e-6)
    assert list(cache) == [(num_frames, min(num_frames, 16))]
    # expected token counts: L = min(frame_idx, 7) memory frames, 4 tokens per pointer, at most 16 pointers
    assert ref[0].shape[0] == min(frame_idx, 7) and ref[2].shape[0] == 4 * min(frame_idx, 16)


def test_assemble_memory_two_cond_frames_and_gap(w):
    """Two conditioning frames (one in the future is ignored for pointers) and a missing non-conditioning frame."""
    g = torch.Generator().manual_seed(7)
    cond = _outputs(g, [0, 30])
    non_cond = _outputs(g, [t for t in range(1, 12) if t != 9])
    ref = O.assemble_memory(w, 12, cond, non_cond, 64)
    got = H.assemble_memory(w, 12, cond, non_cond, 64, cache={})
    for a, b in zip(got, ref):
        torch.testing.assert_close(a.float(), b.float(), rtol=1e-5, atol=1e-6)


def test_heads_hook_is_declined_for_other_head_configurations():
    """speedup() rebinds `_forward_sam_heads` only for the SAM2.1 head configuration the library implements, and never for the
    fork's DAM4SAM track, whose heads return an eighth value (ref: modeling/sam2_base_dam4sam.py:493-503)."""
    import types
    from sam2_opt_b200.speedup import _heads_config_supported
    md = types.SimpleNamespace(use_multimask_token_for_obj_ptr=True, dynamic_multimask_via_stability=True, pred_obj_scores=True,
                               use_high_res_features=True, dynamic_multimask_stability_delta=0.05, dynamic_multimask_stability_thresh=0.98)

    def make(module):
        fn = lambda self, *a, **k: None
        fn.__module__ = module
        cls = type("Model", (), {"_forward_sam_heads": fn})
        m = cls()
        m.sam_mask_decoder = md
        for k, v in dict(use_high_res_features_in_sam=True, pred_obj_scores=True, fixed_no_obj_ptr=True, soft_no_obj_ptr=False,
                         use_obj_ptrs_in_encoder=True, training=False).items():
            setattr(m, k, v)
        return m

    assert _heads_config_supported(make("sam2.modeling.sam2_base_official"))
    assert not _heads_config_supported(make("sam2.modeling.sam2_base_dam4sam"))
    soft = make("sam2.modeling.sam2_base_official")
    soft.soft_no_obj_ptr = True
    assert not _heads_config_supported(soft)
    assert not _heads_config_supported(types.SimpleNamespace())


def test_other_hiera_sizes_are_refused_by_name():
    """weights.derive runs before anything reaches the device: a trunk that is not Hiera-L is refused with its dimensions in the
    message (ref: configs/sam2.1/sam2.1_hiera_b+.yaml: embed_dim 112, stages (2, 3, 16, 3))."""
    from sam2_opt_b200 import weights
    w = O.make_weights(0)
    weights.check_architecture(w)
    small = {k: v for k, v in w.items() if not k.startswith("image_encoder.trunk.blocks.") or int(k.split(".")[3]) < 24}
    small["image_encoder.trunk.patch_embed.proj.weight"] = torch.zeros(112, 3, 7, 7)
    with pytest.raises(ValueError, match="embed_dim 112, 24 blocks"):
        weights.derive(small)
    with pytest.raises(ValueError, match="not a SAM2 model"):
        weights.check_architecture({"foo": torch.zeros(1)})
