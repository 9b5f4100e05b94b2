"""CPU: the oracle's restatement of the mask-prompt path (oracle.sam2_oracle.use_mask_as_output) pinned against the UNMODIFIED
reference's SAM2Base._use_mask_as_output (sam2_base_official.py:496-545) on the same synthetic weights and inputs — outputs to
fp32 round-off.  (The GPU test tests/test_mask_prompt_gpu.py then drives the whole add_new_mask + propagation flow on the
engine against the reference predictor.)"""
import pytest
import torch

from oracle import reference_loader as rl

pytestmark = pytest.mark.skipif(not rl.available(), reason="reference sources not present")


def test_use_mask_as_output_matches_the_reference():
    from oracle import sam2_oracle as O
    torch.set_grad_enabled(False)
    w = O.make_weights(0)
    m = rl.build_reference("video")
    m.load_state_dict(w, strict=True)
    m.eval()
    g = torch.Generator().manual_seed(3)
    pix = torch.randn(2, 256, 64, 64, generator=g)
    hr = [0.5 * torch.randn(2, 32, 256, 256, generator=g), 0.5 * torch.randn(2, 64, 128, 128, generator=g)]
    yy, xx = torch.meshgrid(torch.arange(1024), torch.arange(1024), indexing="ij")
    masks = torch.stack([((xx - 400) ** 2 + (yy - 600) ** 2 < 150 ** 2), torch.zeros(1024, 1024, dtype=torch.bool)])[:, None]   # one empty mask
    want = m._use_mask_as_output(pix, hr, masks)
    got = O.use_mask_as_output(w, pix, hr, masks)
    for name, a, b in zip(("low_res_multimasks", "high_res_multimasks", "ious", "low_res_masks", "high_res_masks", "obj_ptr", "object_score_logits"), got, want):
        assert a.shape == b.shape, (name, a.shape, b.shape)
        err = (a - b).abs().max().item() / max(1.0, b.abs().max().item())
        assert err < 1e-5, (name, err)
    assert got[6][0].item() == 10.0 and got[6][1].item() == -10.0          # the empty mask says "no object"
