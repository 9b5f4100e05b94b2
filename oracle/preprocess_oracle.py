"""TEST INFRASTRUCTURE ONLY (never imported by the product): CPU restatement of the reference's image preprocessing
for frames that are already at the model resolution.

Reference: sam2/utils/transforms.py:27-46 (SAM2Transforms: ToTensor -> Resize -> Normalize; ToTensor is
uint8 HWC -> fp32 CHW `.div(255)`, Resize to (1024, 1024) is the identity for a 1024x1024 frame, Normalize is
`.sub(mean).div(std)` with fp32 mean/std) and the video loader sam2/utils/misc.py:96 (`img_np / 255.0`, a float64
division that is stored into an fp32 tensor) and :161-162 (`img -= mean; img /= std`).

Both are chains of correctly rounded IEEE-754 operations, so the result is defined bit for bit: the oracle repeats them
in numpy fp32.  The float64 route of the video loader rounds to the same fp32 value as a direct fp32 division (a
quotient computed with >= 2*24+2 significand bits and rounded again to 24 bits equals the directly rounded quotient);
`tests/test_preprocess.py` checks that claim over all 256 byte values, and pins this file against torch's own CPU
`div`/`sub` on the same inputs (the operators the reference calls)."""
import numpy as np

IMG_MEAN = (0.485, 0.456, 0.406)
IMG_STD = (0.229, 0.224, 0.225)


def to_tensor(hwc_u8: np.ndarray) -> np.ndarray:
    """uint8 [N,H,W,3] -> fp32 [N,3,H,W] in [0,1]  (transforms.py:29 ToTensor)."""
    x = np.asarray(hwc_u8)
    assert x.dtype == np.uint8 and x.ndim == 4 and x.shape[-1] == 3
    return (x.astype(np.float32) / np.float32(255.0)).transpose(0, 3, 1, 2).copy()


def normalize(x: np.ndarray) -> np.ndarray:
    """fp32 [N,3,H,W] -> (x - mean) / std per colour plane  (transforms.py:31 Normalize; misc.py:161-162)."""
    mean = np.asarray(IMG_MEAN, np.float32)[None, :, None, None]
    std = np.asarray(IMG_STD, np.float32)[None, :, None, None]
    return ((x.astype(np.float32) - mean) / std).astype(np.float32)


def preprocess_u8(hwc_u8: np.ndarray, normalise: bool) -> np.ndarray:
    x = to_tensor(hwc_u8)
    return normalize(x) if normalise else x


def video_loader_route(hwc_u8: np.ndarray) -> np.ndarray:
    """misc.py:96 + :161-162: float64 division stored as fp32, then in-place fp32 normalisation."""
    x = (np.asarray(hwc_u8).astype(np.float64) / 255.0).astype(np.float32).transpose(0, 3, 1, 2).copy()
    return normalize(x)


# ---------------------------------------------------------------------------------------------------------------------
# Antialiased bilinear resize (frames that are not at the model resolution)
# ---------------------------------------------------------------------------------------------------------------------
# Reference: SAM2ImagePredictor._set_image_ (sam2_image_predictor.py:173-196) resizes `images / 255.0` with
# torchvision's Resize((1024, 1024)) (sam2/utils/transforms.py:30), which for tensors is
# F.interpolate(mode="bilinear", antialias=True, align_corners=False).  That operator is a separable triangle filter whose
# support grows with the down-scaling factor (third-party algorithm, PyTorch 2.x `_upsample_bilinear2d_aa`; restated here from
# its published definition, per output index i along one axis, all in fp32):
#     scale   = in_size / out_size                     support = scale if scale >= 1 else 1
#     center  = scale * (i + 0.5)                      invscale = 1 / scale if scale >= 1 else 1
#     xmin    = max(0, int(center - support + 0.5))    xsize = min(in_size, int(center + support + 0.5)) - xmin
#     w_j     = max(0, 1 - |(j + xmin - center + 0.5) * invscale|),  j = 0 .. xsize-1,   then w_j /= sum_j w_j
#     out[i]  = sum_j w_j * in[xmin + j]               (accumulated left to right)
# The width axis is filtered first (fp32 intermediate of shape [in_h, out_w]), then the height axis.
# The accumulation is out = fma(in_j, w_j, out) after the plain product of the first tap (what torch's vectorised CPU kernel
# does: its `t += x * w` loop is contracted to fused multiply-adds); with that order this restatement reproduces torch's CPU
# operator BIT FOR BIT on every case of tests/test_preprocess.py, and the CUDA kernel is bit-exact against this file.
# (The fused multiply-add is emulated as round32(float64(x) * float64(w) + float64(acc)): the product of two fp32 values is
# exact in fp64 and the double rounding of the sum could differ from a true fma only in a halfway case of measure ~2^-29.)
def _aa_axis_weights(in_size: int, out_size: int):
    f = np.float32
    scale = f(in_size) / f(out_size)
    support = scale if scale >= f(1.0) else f(1.0)
    invscale = f(1.0) / scale if scale >= f(1.0) else f(1.0)
    kmax = int(np.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, np.int32)
    xsize = np.zeros(out_size, np.int32)
    w = np.zeros((out_size, kmax), np.float32)
    for i in range(out_size):
        center = f(scale * f(f(i) + f(0.5)))
        lo = max(int(f(f(center - support) + f(0.5))), 0)
        n = min(int(f(f(center + support) + f(0.5))), in_size) - lo
        n = min(max(n, 0), kmax)
        total = f(0.0)
        for j in range(n):
            t = f(f(f(f(j + lo) - center) + f(0.5)) * invscale)
            wj = f(max(f(0.0), f(f(1.0) - abs(t))))
            w[i, j] = wj
            total = f(total + wj)
        if total != 0:
            for j in range(n):
                w[i, j] = f(w[i, j] / total)
        xmin[i], xsize[i] = lo, n
    return xmin, xsize, w


def _aa_filter_axis(x: np.ndarray, axis: int, out_size: int) -> np.ndarray:
    """x fp32 [..., in_size, ...] -> [..., out_size, ...] along `axis`; first tap a plain product, then fused multiply-adds,
    left to right."""
    x = np.moveaxis(np.asarray(x, np.float32), axis, -1)
    xmin, xsize, w = _aa_axis_weights(x.shape[-1], out_size)
    out = np.empty(x.shape[:-1] + (out_size,), np.float32)
    for i in range(out_size):
        lo, n = int(xmin[i]), int(xsize[i])
        acc = (x[..., lo] * w[i, 0]).astype(np.float32)
        for j in range(1, n):
            acc = (x[..., lo + j].astype(np.float64) * np.float64(w[i, j]) + acc.astype(np.float64)).astype(np.float32)
        out[..., i] = acc
    return np.moveaxis(out, -1, axis)


def resize_bilinear_aa(x: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """fp32 [N,C,H,W] -> [N,C,out_h,out_w]: F.interpolate(mode="bilinear", antialias=True, align_corners=False)."""
    x = np.asarray(x, np.float32)
    assert x.ndim == 4
    y = _aa_filter_axis(x, 3, out_w) if x.shape[3] != out_w else x
    return _aa_filter_axis(y, 2, out_h) if x.shape[2] != out_h else y


def resize_frames_u8(hwc_u8: np.ndarray, out_h: int, out_w: int, normalise: bool) -> np.ndarray:
    """uint8 [N,H,W,3] frames -> fp32 [N,3,out_h,out_w]: /255, antialiased resize, optional mean/std normalisation
    (the image predictor's order, sam2_image_predictor.py:193: `self._transforms.resize(images / 255.0)`)."""
    y = resize_bilinear_aa(to_tensor(hwc_u8), out_h, out_w)
    return normalize(y) if normalise else y
