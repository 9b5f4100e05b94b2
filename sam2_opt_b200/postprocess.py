"""Mask post-processing of the video predictor on the device: connected components and small-hole filling.

Mirrors ``get_connected_components`` / ``fill_holes_in_mask_scores`` of the reference (sam2/utils/misc.py:47-63, 311-338),
whose native kernel (sam2/csrc/connected_components.cu) is replaced by ``sam2b_op_connected_components`` — batched over all
masks in four launches instead of a host loop of six launches per mask."""
import torch

from . import _lib
from .engine import _ptr, _stream


def get_connected_components(mask: torch.Tensor):
    """mask: bool / uint8 CUDA tensor [N,1,H,W] -> (labels int32 [N,1,H,W], counts int32 [N,1,H,W]).
    Background (False) pixels get label 0 and count 0; every foreground pixel gets a positive label shared by its
    8-connected component and that component's area (ref: misc.py:47-63)."""
    if not mask.is_cuda:
        raise RuntimeError("get_connected_components: expected a CUDA tensor (sam2_opt_b200 has no CPU path)")
    if mask.dim() != 4 or mask.shape[1] != 1:
        raise ValueError(f"get_connected_components expects [N,1,H,W], got {tuple(mask.shape)}")
    m = mask.to(torch.uint8).contiguous()
    n, _, h, w = m.shape
    labels = torch.empty((n, 1, h, w), device=m.device, dtype=torch.int32)
    counts = torch.empty((n, 1, h, w), device=m.device, dtype=torch.int32)
    if n:
        _lib.check(_lib.load().sam2b_op_connected_components(_ptr(m), _ptr(labels), _ptr(counts), n, h, w, _stream()),
                   "op_connected_components")
    return labels, counts


def fill_holes_in_mask_scores(mask: torch.Tensor, max_area: int) -> torch.Tensor:
    """Fill holes (8-connected background regions, score <= 0, of at most ``max_area`` pixels) with the score 0.1
    (ref: misc.py:311-338).  Unlike the reference there is no try/except fallback: a failing kernel raises."""
    assert max_area > 0, "max_area must be positive"
    labels, areas = get_connected_components(mask <= 0)
    is_hole = (labels > 0) & (areas <= max_area)
    return torch.where(is_hole, 0.1, mask)
