"""GPU: the REFERENCE's own predictors (unmodified sources: /root/reference, or the oracle/_ref copy on the GPU box) driven
through `predictor.speedup("b200")` with the REAL sm_100a engine, against the same predictors on the reference torch path
in strict fp32 on the same device (SURVEY.md §8 rows a19 / b).

Gates (BASELINE.json north_star): predicted-mask IoU >= 0.99, image embeddings / high-res features relative L2 <= 2e-2;
`speedup("torch")` must restore the reference path bit for bit.  The reference's own per-object batch-1 loop
(sam2_video_predictor_official.py:689-725), its `.to(device)` hand-offs, the trimmed `none`-variant input list
(memory_attention.py:376-378) and its executor bookkeeping all run against the real kernels here."""
import numpy as np
import pytest
import torch

from oracle import reference_loader as rl

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not rl.available(), reason="reference sources not present (run `python -m oracle.build_ref` in the build container)")]


def rel_l2(a, b):
    return ((a.double() - b.double()).norm() / b.double().norm()).item()


@pytest.fixture(scope="module")
def weights():
    from oracle import sam2_oracle as O
    return O.make_weights(0)


def test_vendored_reference_is_unmodified():
    """The copy under oracle/_ref (when that is what is being used) still has the sha256 recorded at copy time."""
    from oracle import build_ref
    import os
    if os.path.realpath(rl.REF_ROOT) == os.path.realpath(build_ref.DST_ROOT):
        assert build_ref.verify()


def test_reference_image_predictor_on_the_real_engine(weights, golden):
    """Config 1 through the reference's SAM2ImagePredictor: set_image + predict, torch path vs speedup("b200")."""
    import ref_gpu
    from oracle.synth import synth_image_u8
    from sam2_opt_b200 import speedup as S
    from sam2_opt_b200.engine import Engine, launch_count
    torch.set_grad_enabled(False)
    m = ref_gpu.build("base", weights)
    from sam2.sam2_image_predictor import SAM2ImagePredictor
    S.install(SAM2ImagePredictor)
    pred = SAM2ImagePredictor(m)
    img = synth_image_u8(0)
    kw = dict(point_coords=np.array([[512, 512]]), point_labels=np.array([1]), multimask_output=True, return_logits=True)

    pred.set_image(img)
    ref_feats = [pred._features["high_res_feats"][0].clone(), pred._features["high_res_feats"][1].clone(), pred._features["image_embed"].clone()]
    ref_masks, ref_iou, ref_low = pred.predict(**kw)
    # the reference on this GPU in strict fp32 agrees with the committed CPU fixtures (made by the same sources on CPU)
    cpu_bits = np.unpackbits(golden["c1_low_res_bits"]).reshape(3, 256, 256).astype(bool)
    assert min(ref_gpu.iou(ref_low[i] > 0, cpu_bits[i]) for i in range(3)) >= 0.999
    np.testing.assert_allclose(ref_iou, golden["c1_iou"], atol=1e-4)

    l0 = launch_count()
    pred.speedup("b200")
    assert isinstance(pred._sam2b_engine, Engine) and len(pred.backend_contexts) == 1
    pred.set_image(img)
    assert launch_count() > l0, "speedup('b200') did not launch a single sam2b kernel"
    got_feats = [pred._features["high_res_feats"][0], pred._features["high_res_feats"][1], pred._features["image_embed"]]
    rels = [rel_l2(g, r) for g, r in zip(got_feats, ref_feats)]
    masks, iou, low = pred.predict(**kw)
    ious = [ref_gpu.iou(low[i] > 0, ref_low[i] > 0) for i in range(3)]
    ious_full = [ref_gpu.iou(masks[i] > 0, ref_masks[i] > 0) for i in range(3)]
    print("reference SAM2ImagePredictor + b200 engine: feature rel-L2", [f"{r:.4%}" for r in rels],
          "low-res IoU", [f"{v:.4f}" for v in ious], "1024^2 IoU", [f"{v:.4f}" for v in ious_full])
    assert max(rels) <= 2e-2
    assert min(ious) >= 0.99 and min(ious_full) >= 0.99
    assert np.abs(iou - ref_iou).max() < 2e-2

    ctx = pred.backend_contexts[0]
    pred.speedup("torch")                       # reference path restored, bit for bit
    assert pred.backend_contexts == [] and ctx._released
    pred.set_image(img)
    for a, b in zip([pred._features["high_res_feats"][0], pred._features["high_res_feats"][1], pred._features["image_embed"]], ref_feats):
        assert torch.equal(a, b)
    again = pred.predict(**kw)
    assert np.array_equal(again[2], ref_low)
    pred._sam2b_engine.release()


def test_reference_video_predictor_on_the_real_engine(weights):
    """Config 3 (first 16 frames: the 7-frame bank fills at frame 7, pointers keep growing to 15) through the reference's
    SAM2VideoPredictor and its own per-object loop: two objects, torch path vs speedup("b200") on the same device."""
    import ref_gpu
    from oracle.synth import synth_video
    from sam2_opt_b200 import speedup as S
    from sam2_opt_b200.engine import launch_count
    torch.set_grad_enabled(False)
    T = 16
    frames = synth_video(T, seed=1)
    clicks = {1: (512.0, 512.0), 2: (300.0, 700.0)}
    m = ref_gpu.build("video", weights)
    S.install(type(m))
    ref = ref_gpu.propagate_low_res(m, frames, clicks)

    l0 = launch_count()
    m.speedup("b200")
    assert [c.kind for c in m.memory_attention.backend_contexts] == ["memory_attention_none", "memory_attention_exclude"]
    got = ref_gpu.propagate_low_res(m, frames, clicks)
    n_launch = launch_count() - l0
    assert n_launch > 1000, n_launch
    assert sorted(got) == list(range(T))
    for oid in clicks:
        ious = [ref_gpu.iou(got[t][oid] > 0, ref[t][oid] > 0) for t in range(T)]
        print(f"reference SAM2VideoPredictor + b200 engine, object {oid}: per-frame IoU min {min(ious):.4f}", [f"{v:.4f}" for v in ious])
        assert min(ious) >= 0.99

    engine = m._sam2b_engine
    m.speedup("torch")
    assert m.inference_image == m.inference_image_torch and m.memory_attention.backend_contexts == []
    again = ref_gpu.propagate_low_res(m, frames, {1: clicks[1]}, max_track=2)
    for t in range(3):
        # frames 0-2 of a one-object run on the restored torch path equal the first torch run's object 1 bit for bit
        # (the reference tracks every object separately with batch 1)
        assert torch.equal(again[t][1], ref[t][1])
    engine.release()


def test_reference_vos_flow_mask_prompts_on_the_real_engine(weights):
    """the flow of the reference's tools/vos_inference.py:175-205 (`init_state`, one `add_new_mask` per object
    on the first frame, `propagate_in_video`) through the reference's SAM2VideoPredictor, torch path vs speedup("b200") on the
    same device.  `_use_mask_as_output` (sam2_base_official.py:496-546) keeps the mask itself as the frame's output and calls the
    rebound `_forward_sam_heads` with the down-sampled mask as prompt for the object pointer."""
    import ref_gpu
    from oracle.synth import synth_video
    from sam2_opt_b200 import speedup as S
    from sam2_opt_b200.engine import launch_count
    torch.set_grad_enabled(False)
    T = 9
    frames = synth_video(T, seed=2)
    yy, xx = np.mgrid[0:1024, 0:1024]
    masks = {1: (xx - 420) ** 2 + (yy - 500) ** 2 < 170 ** 2, 2: (np.abs(xx - 760) < 90) & (np.abs(yy - 300) < 140)}
    m = ref_gpu.build("video", weights)
    S.install(type(m))
    ref, ref_ptr = ref_gpu.propagate_low_res_from_masks(m, frames, masks)
    l0 = launch_count()
    m.speedup("b200")
    got, got_ptr = ref_gpu.propagate_low_res_from_masks(m, frames, masks)
    assert launch_count() - l0 > 1000 and sorted(got) == list(range(T))
    for oid in masks:
        assert torch.allclose(got[0][oid], ref[0][oid], atol=1e-4), "the prompted frame is the mask itself"
        r = rel_l2(got_ptr[oid], ref_ptr[oid])
        ious = [ref_gpu.iou((got[t][oid] > 0).numpy(), (ref[t][oid] > 0).numpy()) for t in range(1, T)]
        print(f"reference VOS flow + b200 engine, object {oid}: frame-0 pointer rel-L2 {r:.3e}, propagated IoU min {min(ious):.4f}")
        assert r <= 2e-2 and min(ious) >= 0.99
    engine = m._sam2b_engine
    m.speedup("torch")
    engine.release()


def test_reference_automatic_mask_generator_on_the_real_engine(weights):
    """SURVEY section 8 f-4 (other consumers of the same kernels): the reference's SAM2AutomaticMaskGenerator
    (automatic_mask_generator.py:170-294) owns a SAM2ImagePredictor, so `generator.predictor.speedup("b200")` is all it takes.
    A 4 x 4 grid of point prompts with every filter that depends on the (random-init) quality heads opened up, torch path vs
    engine on the same device: the same prompts must survive and every mask must match (IoU >= 0.99)."""
    import ref_gpu
    from oracle.synth import synth_image_u8
    from sam2_opt_b200 import speedup as S
    torch.set_grad_enabled(False)
    m = ref_gpu.build("base", weights)
    from sam2.automatic_mask_generator import SAM2AutomaticMaskGenerator
    from sam2.sam2_image_predictor import SAM2ImagePredictor
    S.install(SAM2ImagePredictor)
    gen = SAM2AutomaticMaskGenerator(m, points_per_side=4, points_per_batch=16, pred_iou_thresh=0.0, stability_score_thresh=0.0,
                                     box_nms_thresh=1.0, crop_n_layers=0, min_mask_region_area=0)
    img = synth_image_u8(0)
    ref = gen.generate(img)
    gen.predictor.speedup("b200")
    got = gen.generate(img)
    gen.predictor.speedup("torch")
    assert len(ref) >= 16 and len(got) == len(ref), (len(ref), len(got))

    def by_prompt(records):
        out = {}
        for r in records:
            out.setdefault(tuple(np.round(np.asarray(r["point_coords"][0]), 1)), []).append(r)
        return out

    a, b = by_prompt(ref), by_prompt(got)
    assert sorted(a) == sorted(b)
    worst = 1.0
    for k in a:
        assert len(a[k]) == len(b[k])
        # the masks of one prompt come out ordered by the quality heads, whose scores are near-tied with random weights:
        # match every reference mask with its best counterpart
        for r in a[k]:
            worst = min(worst, max(ref_gpu.iou(r["segmentation"], g["segmentation"]) for g in b[k]))
    print(f"reference SAM2AutomaticMaskGenerator + b200 engine: {len(ref)} masks from 16 prompts, worst matched IoU {worst:.4f}")
    assert worst >= 0.99
    gen.predictor._sam2b_engine.release()
