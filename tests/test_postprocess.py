"""Hole filling / connected components: known-answer cases for the oracle (CPU) and CUDA-vs-oracle parity (GPU).
Integer work: areas and the filled masks are compared bit for bit; labels only up to renaming."""
import numpy as np
import pytest
import torch

from oracle import postprocess_oracle as PO


def _known_answer_cases():
    m = np.ones((1, 1, 8, 8), np.float32)
    m[0, 0, 2, 2] = -1.0                      # 1-pixel hole
    m[0, 0, 4, 4] = -1.0                      # two pixels touching only diagonally: ONE 8-connected hole of area 2
    m[0, 0, 5, 5] = -1.0
    m[0, 0, 0, 5:8] = 0.0                     # score exactly 0 counts as background; area 3
    exp_area = np.zeros((8, 8), np.int32)
    exp_area[2, 2] = 1
    exp_area[4, 4] = exp_area[5, 5] = 2
    exp_area[0, 5:8] = 3
    return m, exp_area


def test_oracle_known_answers():
    m, exp_area = _known_answer_cases()
    labels, counts = PO.connected_components(m <= 0)
    assert np.array_equal(counts[0, 0], exp_area)
    assert labels[0, 0, 4, 4] == labels[0, 0, 5, 5] != 0          # diagonal neighbours share a label (8-connectivity)
    assert len({labels[0, 0, 2, 2], labels[0, 0, 4, 4], labels[0, 0, 0, 5]}) == 3
    f2 = PO.fill_holes_in_mask_scores(m, 2)
    assert f2[0, 0, 2, 2] == np.float32(0.1) and f2[0, 0, 4, 4] == np.float32(0.1) and f2[0, 0, 5, 5] == np.float32(0.1)
    assert np.array_equal(f2[0, 0, 0, 5:8], m[0, 0, 0, 5:8])      # area 3 > max_area 2: left alone
    f1 = PO.fill_holes_in_mask_scores(m, 1)
    assert f1[0, 0, 2, 2] == np.float32(0.1) and f1[0, 0, 4, 4] == -1.0
    assert np.array_equal(PO.fill_holes_in_mask_scores(m, 3) > 0, np.ones((1, 1, 8, 8), bool))


def _same_partition(a, b):
    """labels equal up to renaming (per image)"""
    for i in range(a.shape[0]):
        x, y = a[i].ravel().astype(np.int64), b[i].ravel().astype(np.int64)
        if not np.array_equal(x > 0, y > 0):
            return False
        pairs = np.unique(np.stack([x, y], 1), axis=0)
        if len(np.unique(pairs[:, 0])) != len(pairs) or len(np.unique(pairs[:, 1])) != len(pairs):
            return False
    return True


def _cases():
    rng = np.random.default_rng(0)
    yield "known", _known_answer_cases()[0]
    yield "random50", rng.standard_normal((4, 1, 256, 256)).astype(np.float32)
    yield "sparse_holes", np.where(rng.random((3, 1, 256, 256)) < 0.02, -1.0, 1.0).astype(np.float32)
    yield "all_background", -np.ones((2, 1, 64, 64), np.float32)
    yield "all_foreground", np.ones((2, 1, 64, 64), np.float32)
    cb = np.indices((128, 128)).sum(0) % 2
    yield "checkerboard", np.where(cb, -1.0, 1.0).astype(np.float32)[None, None]       # one big diagonal-connected component
    sp = np.ones((256, 256), np.float32)                                                # spiral: one long thin component
    for k in range(0, 120, 4):
        sp[k, k:256 - k] = -1; sp[k:256 - k, 255 - k] = -1; sp[255 - k, k:256 - k] = -1; sp[k + 4:256 - k, k] = -1
    yield "spiral", sp[None, None]
    yield "ragged", rng.standard_normal((2, 1, 37, 91)).astype(np.float32)


@pytest.mark.gpu
@pytest.mark.parametrize("name,mask", list(_cases()), ids=[n for n, _ in _cases()])
def test_connected_components_and_fill_vs_oracle(name, mask):
    from sam2_opt_b200 import postprocess as PP
    m = torch.from_numpy(mask).cuda()
    labels, counts = PP.get_connected_components(m <= 0)
    torch.cuda.synchronize()
    ref_l, ref_c = PO.connected_components(mask <= 0)
    assert np.array_equal(counts.cpu().numpy(), ref_c)
    assert _same_partition(labels.cpu().numpy(), ref_l)
    for max_area in (1, 8, 100000):
        got = PP.fill_holes_in_mask_scores(m, max_area).cpu().numpy()
        assert np.array_equal(got, PO.fill_holes_in_mask_scores(mask, max_area)), (name, max_area)


@pytest.mark.gpu
def test_fill_is_idempotent_and_only_raises_scores():
    from sam2_opt_b200 import postprocess as PP
    g = torch.Generator(device="cuda").manual_seed(3)
    m = torch.randn(16, 1, 256, 256, device="cuda", generator=g)
    f = PP.fill_holes_in_mask_scores(m, 8)
    assert torch.equal(PP.fill_holes_in_mask_scores(f, 8), f)          # filled holes are foreground now
    assert bool((f >= m).all()) and bool(((f != m) <= (m <= 0)).all())


def test_rejects_cpu_tensor():
    from sam2_opt_b200 import postprocess as PP
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        PP.get_connected_components(torch.zeros(1, 1, 8, 8, dtype=torch.bool))


# ---- mask up-sampling (+ threshold): sam2b_op_upsample_masks vs oracle/postprocess_oracle.py:upsample_bilinear ----
def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _logits(n, c, h, w, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((n, c, h, w)) * 8).astype(np.float32)


UPSAMPLE_CASES = [  # (n, c, h, w, H, W): SAM2's 256 -> 1024, video resolutions (W % 4 != 0 takes the scalar route), down-sampling, identity
    (2, 3, 256, 256, 1024, 1024), (1, 1, 256, 256, 480, 854), (3, 1, 256, 256, 720, 1280), (1, 2, 64, 64, 37, 53),
    (1, 1, 256, 256, 256, 256), (1, 1, 256, 256, 100, 100), (2, 1, 1, 1, 8, 8), (1, 1, 5, 7, 1, 1)]


@pytest.mark.parametrize("case", UPSAMPLE_CASES[:5] + UPSAMPLE_CASES[6:], ids=str)
def test_upsample_oracle_pinned_against_torch_interpolate(case):
    """The operator the reference calls (transforms.py:119) on CPU; torch's CPU kernel may contract multiply-adds, so the
    pin is 2e-6 relative to the logit scale, and the thresholded masks may differ only where |logit| is below that."""
    n, c, h, w, H, W = case
    x = _logits(n, c, h, w, 1)
    want = torch.nn.functional.interpolate(torch.from_numpy(x), (H, W), mode="bilinear", align_corners=False).numpy()
    got = PO.upsample_bilinear(x, (H, W))
    assert got.shape == want.shape and got.dtype == np.float32
    tol = 2e-6 * float(np.abs(x).max())
    assert float(np.abs(got - want).max()) <= tol
    diff = (got > 0) != (want > 0)
    assert not diff.any() or float(np.abs(want[diff]).max()) <= tol


def test_upsample_oracle_known_answers():
    x = np.array([[[[0.0, 4.0], [8.0, 12.0]]]], np.float32)
    y = PO.upsample_bilinear(x, (4, 4))
    assert np.array_equal(y[0, 0, 0], np.array([0.0, 1.0, 3.0, 4.0], np.float32))        # corners clamp, 1/4-3/4 weights inside
    assert np.array_equal(y[0, 0, :, 0], np.array([0.0, 2.0, 6.0, 8.0], np.float32))
    assert np.array_equal(PO.upsample_bilinear(x, (2, 2)), x)                            # identity at the same size
    assert np.array_equal(PO.upsample_bilinear(np.full((1, 1, 3, 3), 2.5, np.float32), (7, 5)), np.full((1, 1, 7, 5), 2.5, np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("case", UPSAMPLE_CASES, ids=str)
def test_upsample_masks_bit_exact_vs_oracle(case):
    from sam2_opt_b200 import postprocess as PP
    n, c, h, w, H, W = case
    x = _logits(n, c, h, w, 2)
    want = PO.upsample_bilinear(x, (H, W))
    xd = torch.from_numpy(x).cuda()
    out, mask = PP.upsample_masks(xd, (H, W), threshold=0.0)
    assert out.shape == (n, c, H, W) and mask.dtype == torch.bool
    assert np.array_equal(_bits(out.cpu().numpy()), _bits(want))
    assert np.array_equal(mask.cpu().numpy(), want > 0)
    only = PP.upsample_masks(xd, (H, W), threshold=1.5, return_logits=False)
    assert np.array_equal(only.cpu().numpy(), want > 1.5)
    assert np.array_equal(_bits(PP.upsample_masks(xd, (H, W)).cpu().numpy()), _bits(want))


@pytest.mark.gpu
def test_upsample_masks_full_size_properties():
    """64 object masks 256^2 -> 1024^2 (the video predictor's per-frame load): constants are preserved, the result is bounded
    by the source's range, and it matches torch's CUDA operator within fp32 rounding."""
    from sam2_opt_b200 import postprocess as PP
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(64, 1, 256, 256, device="cuda", generator=g) * 8
    y, m = PP.upsample_masks(x, (1024, 1024), threshold=0.0)
    assert float(y.max()) <= float(x.max()) and float(y.min()) >= float(x.min())
    assert torch.equal(m, y > 0)
    ref = torch.nn.functional.interpolate(x, (1024, 1024), mode="bilinear", align_corners=False)
    assert float((y - ref).abs().max()) <= 2e-6 * float(x.abs().max())
    c = torch.full((2, 1, 256, 256), -3.25, device="cuda")
    assert torch.equal(PP.upsample_masks(c, (480, 854)), torch.full((2, 1, 480, 854), -3.25, device="cuda"))


def test_upsample_masks_rejects_cpu_tensor_and_bad_arguments():
    from sam2_opt_b200 import postprocess as PP
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        PP.upsample_masks(torch.zeros(1, 1, 8, 8), (16, 16))
