"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's hole-filling post-processor.

Reference: sam2/utils/misc.py:47-63 (get_connected_components) and :311-338 (fill_holes_in_mask_scores), whose labelling
is done by sam2/csrc/connected_components.cu — a block-based union-find that merges a 2x2 block with its left, top-left,
top and top-right neighbour blocks whenever any pair of their pixels touches, including diagonally (:75-118), i.e.
8-connectivity.  The reference kernel needs a GPU and a torch extension build and cannot run in the build container, so
this oracle is pinned by construction instead: scipy.ndimage.label with the full 3x3 structuring element is the textbook
definition of 8-connected components, and hand-made known-answer cases (tests/test_postprocess.py) pin the
connectivity, the area rule (<= max_area) and the fill value 0.1.  Label VALUES are not part of the contract (the only
consumer tests labels > 0); areas and the filled mask are compared bit for bit."""
import numpy as np
from scipy import ndimage

_EIGHT = np.ones((3, 3), dtype=np.int32)


def connected_components(mask: np.ndarray):
    """mask bool [N,1,H,W] -> (labels int32, counts int32): per-image 8-connected labels (0 = background) and areas."""
    mask = np.asarray(mask).astype(bool)
    labels = np.zeros(mask.shape, np.int32)
    counts = np.zeros(mask.shape, np.int32)
    for i in range(mask.shape[0]):
        lab, k = ndimage.label(mask[i, 0], structure=_EIGHT)
        area = np.bincount(lab.ravel(), minlength=k + 1)
        area[0] = 0
        labels[i, 0] = lab
        counts[i, 0] = area[lab]
    return labels, counts


def fill_holes_in_mask_scores(mask: np.ndarray, max_area: int) -> np.ndarray:
    """ref misc.py:311-338: holes = components of (mask <= 0) with area <= max_area; filled with the score 0.1."""
    assert max_area > 0
    mask = np.asarray(mask, np.float32)
    labels, areas = connected_components(mask <= 0)
    return np.where((labels > 0) & (areas <= max_area), np.float32(0.1), mask)


def _bilinear_coord(out_size: int, in_size: int):
    """torch's area_pixel_compute_source_index (align_corners=False) and index / weight computation of
    upsample_bilinear2d (aten/src/ATen/native/UpSample.h), in fp32: src = max(fma(scale, dst + 0.5, -0.5), 0),
    scale = float(in) / out; i0 = int(src); i1 = i0 + (i0 < in - 1); l1 = src - i0; l0 = 1 - l1.
    torch's builds contract `scale * (dst + 0.5) - 0.5` into one fused multiply-add (checked against the CPU operator in
    tests/test_postprocess.py: with two roundings the source coordinate is off by an ulp for ~40 % of the columns of a
    256 -> 854 resize); the float64 evaluation below is exact before the single rounding to fp32, i.e. the fma."""
    scale = np.float32(in_size) / np.float32(out_size)
    dst = np.arange(out_size, dtype=np.float32)
    src = (np.float64(scale) * (dst.astype(np.float64) + 0.5) - 0.5).astype(np.float32)
    src = np.maximum(src, np.float32(0.0)).astype(np.float32)
    i0 = np.minimum(src.astype(np.int64), in_size - 1)
    i1 = i0 + (i0 < in_size - 1)
    l1 = np.clip(src - i0.astype(np.float32), np.float32(0), np.float32(1)).astype(np.float32)
    l0 = (np.float32(1.0) - l1).astype(np.float32)
    return i0, i1, l0, l1


def upsample_bilinear(low: np.ndarray, size) -> np.ndarray:
    """fp32 [N,C,h,w] -> [N,C,H,W]: F.interpolate(mode="bilinear", align_corners=False) as called by
    SAM2Transforms.postprocess_masks (sam2/utils/transforms.py:119), every operation rounded to fp32 separately:
    out = hl0 * (wl0 * v00 + wl1 * v01) + hl1 * (wl0 * v10 + wl1 * v11)."""
    x = np.asarray(low, np.float32)
    H, W = int(size[0]), int(size[1])
    y0, y1, hl0, hl1 = _bilinear_coord(H, x.shape[-2])
    x0, x1, wl0, wl1 = _bilinear_coord(W, x.shape[-1])
    r0, r1 = x[..., y0, :], x[..., y1, :]
    top = (wl0 * r0[..., x0]).astype(np.float32) + (wl1 * r0[..., x1]).astype(np.float32)
    bot = (wl0 * r1[..., x0]).astype(np.float32) + (wl1 * r1[..., x1]).astype(np.float32)
    hl0, hl1 = hl0[:, None], hl1[:, None]
    return ((hl0 * top).astype(np.float32) + (hl1 * bot).astype(np.float32)).astype(np.float32)
