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
