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
