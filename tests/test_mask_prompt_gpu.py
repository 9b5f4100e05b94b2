"""GPU: mask prompts (`add_new_mask`, the entry the reference's VOS tool uses: tools/vos_inference.py:175-190) on the standalone
video predictor against the UNMODIFIED reference's torch path in strict fp32 on the same device.  The prompted frame's output
is the mask itself (interpolation only: allclose), its object pointer comes from the SAM heads with the mask as dense prompt
(rel-L2 <= 2e-2: it sees the bf16 trunk features), and every propagated frame must reach mask IoU >= 0.99."""
import numpy as np
import pytest
import torch

from oracle import reference_loader as rl

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not rl.available(), reason="reference sources not present")
def test_mask_prompt_and_propagation_vs_reference_fp32():
    import ref_gpu
    from oracle import sam2_oracle as O
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    torch.set_grad_enabled(False)
    weights = O.make_weights(0)
    T = 10
    frames = synth_video(T, seed=2)
    yy, xx = np.mgrid[0:1024, 0:1024]
    masks = {1: (xx - 420) ** 2 + (yy - 500) ** 2 < 170 ** 2,                     # a disk
             2: (np.abs(xx[::2, ::2] - 760) < 90) & (np.abs(yy[::2, ::2] - 300) < 140)}    # a rectangle given at 512 x 512: resized
    m = ref_gpu.build("video", weights)
    ref, ref_ptr = ref_gpu.propagate_low_res_from_masks(m, frames, masks)
    del m
    torch.cuda.empty_cache()
    vp = SAM2VideoPredictor(weights)
    st = vp.init_state(frames)
    for oid, mk in masks.items():
        vp.add_new_mask(st, 0, oid, mk)
    assert [t for t, _, _ in vp.propagate_in_video(st)] == list(range(T))
    for oid in masks:
        got0 = st["cond"][0][oid]
        assert torch.allclose(got0["pred_masks"][0, 0].cpu(), ref[0][oid], atol=1e-4), "the prompted frame is the mask itself"
        r = ((got0["obj_ptr"][0].cpu() - ref_ptr[oid]).norm() / ref_ptr[oid].norm()).item()
        ious = [ref_gpu.iou((st["non_cond"][oid][t]["pred_masks"][0, 0] > 0).cpu().numpy(), (ref[t][oid] > 0).numpy()) for t in range(1, T)]
        print(f"object {oid}: frame-0 pointer rel-L2 {r:.3e}, propagated IoU min {min(ious):.4f}", [f"{v:.4f}" for v in ious])
        assert r <= 2e-2 and min(ious) >= 0.99
    vp.release()
