"""Mask post-processing of the video predictor on the device: connected components and small-hole filling.

Mirrors ``get_connected_components`` / ``fill_holes_in_mask_scores`` of the reference (sam2/utils/misc.py:47-63, 311-338),
whose native kernel (sam2/csrc/connected_components.cu) is replaced by ``sam2b_op_connected_components`` — batched over all
masks in four launches instead of a host loop of six launches per mask."""
import torch

from . import _lib
from .engine import _ptr, _stream, on_tensor_device


@on_tensor_device
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


@on_tensor_device
def upsample_masks(low: torch.Tensor, size, threshold=None, return_logits: bool = True):
    """Mask logits [N,C,h,w] (fp32, CUDA) -> output resolution ``size`` = (H, W), bilinear with align_corners=False, in one
    pass; mirrors the ``F.interpolate(masks, orig_hw, mode="bilinear", align_corners=False)`` of
    ``SAM2Transforms.postprocess_masks`` (ref: sam2/utils/transforms.py:119) and, with ``threshold`` given, the
    ``masks > mask_threshold`` that follows it (ref: sam2_image_predictor.py:300-302) without re-reading the logits.
    Returns logits, (logits, bool mask) or — with ``return_logits=False`` — the bool mask only (1 B instead of 5 B written
    per pixel)."""
    if not low.is_cuda:
        raise RuntimeError("upsample_masks: expected a CUDA tensor (sam2_opt_b200 has no CPU path)")
    if low.dim() != 4:
        raise ValueError(f"upsample_masks expects [N,C,h,w], got {tuple(low.shape)}")
    if threshold is None and not return_logits:
        raise ValueError("upsample_masks: nothing to return (no threshold and return_logits=False)")
    x = low.float().contiguous()
    n, c, h, w = x.shape
    H, W = int(size[0]), int(size[1])
    out = torch.empty((n, c, H, W), device=x.device, dtype=torch.float32) if return_logits else None
    mask = torch.empty((n, c, H, W), device=x.device, dtype=torch.uint8) if threshold is not None else None
    if n * c:
        _lib.check(_lib.load().sam2b_op_upsample_masks(_ptr(x), _ptr(out) if out is not None else None,
                                                       _ptr(mask) if mask is not None else None, n * c, h, w, H, W,
                                                       float(threshold if threshold is not None else 0.0), _stream()),
                   "op_upsample_masks")
    if mask is None:
        return out
    mask = mask.view(torch.bool)
    return (out, mask) if return_logits else mask
