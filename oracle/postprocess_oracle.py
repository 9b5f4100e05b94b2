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
