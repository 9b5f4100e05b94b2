"""CPU: `speedup(predictor, "b200")` wires the REFERENCE's own predictor objects to B200Executor sub-graphs correctly.
The engine is replaced by a stub that answers each sub-graph with the oracle (no GPU here), so what is checked is the
drop-in boundary: which callables are rebound, argument order / normalisation / no_mem_embed conventions, executor
bookkeeping and release — the predictions must equal the reference's own (golden) results."""
import numpy as np
import pytest
import torch

from oracle import reference_loader as rl
from oracle import sam2_oracle as O
from oracle.synth import synth_image_u8, synth_video

pytestmark = pytest.mark.skipif(not rl.available(), reason="/root/reference not present (GPU box)")
torch.set_grad_enabled(False)


class StubEngine:
    device = torch.device("cpu")

    def __init__(self, sd, device=None):
        self.w = {k: v.float() for k, v in sd.items()}
        self.calls = []

    def image_encode(self, img, normalised, add_no_mem_embed):
        self.calls.append(("image_encode", normalised, add_no_mem_embed))
        o = O.forward_image(self.w, img if normalised else O.normalize_image(img))
        f2 = o[6] + self.w["no_mem_embed"].view(1, -1, 1, 1) if add_no_mem_embed else o[6]
        return o[4], o[5], f2

    def position_encoding(self, c, size, n):
        return O.position_sine(c, size, size)[None].expand(n, -1, -1, -1)

    def memory_attention(self, curr, memory, curr_pos, memory_pos, ex=None, pex=None):
        self.calls.append(("memory_attention", memory.shape[0], 0 if ex is None else ex.shape[0]))
        if ex is None:
            ex, pex = torch.zeros(0, curr.shape[1], 64), torch.zeros(0, curr.shape[1], 64)
        return O.memory_attention(self.w, curr, memory, curr_pos, memory_pos, ex, pex)

    def memory_encode(self, pix, masks):
        self.calls.append(("memory_encode",))
        return O.memory_encoder(self.w, pix, masks)

    def release(self):
        pass


@pytest.fixture(scope="module")
def weights():
    return O.make_weights(0)


def test_image_predictor_speedup_dropin(weights, golden):
    from sam2_opt_b200 import speedup as S
    from sam2_opt_b200.executor import B200Executor, ModelExectuor
    m = rl.build_reference("base")
    m.load_state_dict(weights, strict=True)
    from sam2.sam2_image_predictor import SAM2ImagePredictor
    S.install(SAM2ImagePredictor)
    pred = SAM2ImagePredictor(m)
    pred.speedup("b200", engine_factory=StubEngine, warmup=False)
    assert len(pred.backend_contexts) == 1 and isinstance(pred.backend_contexts[0], (B200Executor, ModelExectuor))
    assert [d.name for d in pred.backend_contexts[0].GetModelInputDesc()] == ["image"]
    assert [d.name for d in pred.backend_contexts[0].GetModelOutputDesc()] == ["feature0", "feature1", "feature2"]
    pred.set_image(synth_image_u8(0))
    eng = pred._sam2b_engine
    assert eng.calls == [("image_encode", False, True)]       # E1: un-normalised input, no_mem_embed folded in
    masks, iou, low = pred.predict(point_coords=np.array([[512, 512]]), point_labels=np.array([1]), multimask_output=True, return_logits=True)
    np.testing.assert_allclose(iou, golden["c1_iou"], atol=1e-5)
    np.testing.assert_allclose(low[:, ::16, ::16], golden["c1_low_res_slice"], atol=1e-4)
    # wrong input count -> AssertionError like the reference executor; bad output type -> Exception
    with pytest.raises(AssertionError):
        pred.backend_contexts[0].Inference([], output_type="torch")
    with pytest.raises(Exception, match="unsupported output_type"):
        pred.backend_contexts[0].Inference([torch.zeros(1, 3, 1024, 1024)], output_type="tf")
    ctx = pred.backend_contexts[0]
    pred.speedup("torch")                                    # reference path restored, executor released
    assert pred.backend_contexts == [] and ctx._released
    assert pred.set_image_e2e == pred.set_image_e2e_torch
    with pytest.raises(RuntimeError, match="after Release"):
        ctx.Inference([torch.zeros(1, 3, 1024, 1024)])


def test_video_predictor_speedup_dropin(weights, golden):
    from sam2_opt_b200 import speedup as S
    m = rl.build_reference("video")
    m.load_state_dict(weights, strict=True)
    import sam2.sam2_video_predictor_official as vp
    T = 3
    frames = synth_video(4, seed=1)[:T]
    vp.load_video_frames = lambda **k: (frames, 1024, 1024)
    eng = S.speedup(m, "b200", engine_factory=StubEngine, warmup=False)
    assert [c.kind for c in m.memory_attention.backend_contexts] == ["memory_attention_none", "memory_attention_exclude"]
    assert len(m.memory_attention.backend_contexts[0].GetModelInputDesc()) == 4
    st = m.init_state("synthetic")
    m.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
    for fi, _, _ in m.propagate_in_video(st):
        d = st["output_dict_per_obj"][0]
        low = d["cond_frame_outputs" if fi == 0 else "non_cond_frame_outputs"][fi]["pred_masks"]
        np.testing.assert_allclose(low[0, 0, ::16, ::16].numpy(), golden["c3_low_res_slice"][fi], atol=2e-4)
    kinds = [c[0] for c in eng.calls]
    # the frames are encoded in ONE block (speedup(encode_ahead=16) fills the reference's per-frame feature cache ahead)
    assert kinds.count("image_encode") == 1 and kinds.count("memory_attention") == T - 1 and kinds.count("memory_encode") == T
    assert all(c[1:] == (True, False) for c in eng.calls if c[0] == "image_encode")   # E2: normalised in, no no_mem_embed
    assert [c[1:] for c in eng.calls if c[0] == "memory_attention"] == [(1, 4), (2, 8)]
    S.speedup(m, "torch")
    assert m.inference_image == m.inference_image_torch and m.memory_attention.backend_contexts == []


def test_install_postprocess_rebinds_reference_labelling():
    """The reference's fill_holes_in_mask_scores looks `get_connected_components` up in its own module at call time
    (misc.py:320), so rebinding the module attribute is enough; on this CPU-only box our function then raises (no CPU
    path) and the reference falls back to returning the mask unchanged with a warning (misc.py:325-336)."""
    import warnings
    import torch
    rl.build_reference("image")            # puts the reference package (with its stubs) on sys.path
    import sam2.utils.misc as misc
    from sam2_opt_b200 import postprocess
    from sam2_opt_b200.speedup import install_postprocess
    install_postprocess(misc)
    assert misc.get_connected_components is postprocess.get_connected_components
    m = torch.randn(1, 1, 16, 16)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        out = misc.fill_holes_in_mask_scores(m, 8)
    assert torch.equal(out, m) and any("CUDA tensor" in str(x.message) for x in w)
