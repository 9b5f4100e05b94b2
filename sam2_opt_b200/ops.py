"""torch-tensor front ends of the library's building-block entry points (``sam2b_op_*``): used by the parity tests
and by callers that want a single kernel.  Every function enqueues on the current stream of the tensors' device."""
import ctypes as C
from typing import Optional

import torch

from . import _lib
from .engine import _ptr, _stream, on_tensor_device

ACT = {None: 0, "none": 0, "gelu": 1, "relu": 2}


@on_tensor_device
def gemm(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None, act=None, alpha: float = 1.0,
         gamma: Optional[torch.Tensor] = None, residual: Optional[torch.Tensor] = None, res_row_mod: int = 0,
         out_dtype=torch.float32, bias_along_m: bool = False, block_n: int = 0) -> torch.Tensor:
    """out[b] = act(alpha * a[b] @ b[b]^T + bias) * gamma + residual.   a: [M,K] or [B,M,K] bf16; b: [N,K] or [B,N,K]."""
    lib = _lib.load()
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda
    a3 = a if a.dim() == 3 else a[None]
    b3 = b if b.dim() == 3 else b[None]
    assert a3.stride(2) == 1 and b3.stride(2) == 1
    batch, M, K = a3.shape
    N = b3.shape[1]
    out = torch.empty((batch, M, N), device=a.device, dtype=out_dtype)
    res3 = None
    if residual is not None:
        res3 = residual if residual.dim() == 3 else residual[None]
        assert res3.dtype == torch.float32 and res3.stride(2) == 1
    _lib.check(lib.sam2b_op_gemm(
        _ptr(a3), a3.stride(1), a3.stride(0) if batch > 1 else 0,
        _ptr(b3), b3.stride(1), b3.stride(0) if b3.shape[0] > 1 else 0,
        M, N, K, batch, float(alpha), _ptr(bias), int(bias_along_m), ACT[act], _ptr(gamma),
        _ptr(res3), 0 if res3 is None else res3.stride(1),
        0 if res3 is None or res3.shape[0] == 1 else res3.stride(0), int(res_row_mod),
        _ptr(out if out_dtype == torch.float32 else None), _ptr(out if out_dtype == torch.bfloat16 else None),
        N, M * N, int(block_n), _stream()), "op_gemm")
    return out if a.dim() == 3 else out[0]


@on_tensor_device
def layernorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, out_dtype=torch.bfloat16):
    lib = _lib.load()
    x = x.contiguous()
    Cdim = x.shape[-1]
    rows = x.numel() // Cdim
    y = torch.empty(x.shape, device=x.device, dtype=out_dtype)
    _lib.check(lib.sam2b_op_layernorm(_ptr(x), _ptr(gamma), _ptr(beta), float(eps),
                                      _ptr(y if out_dtype == torch.bfloat16 else None),
                                      _ptr(y if out_dtype == torch.float32 else None), rows, Cdim, _stream()), "op_layernorm")
    return y


@on_tensor_device
def cast_bf16(x: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    x = x.contiguous()
    y = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    _lib.check(lib.sam2b_op_cast_bf16(_ptr(x), _ptr(y), x.numel(), _stream()), "op_cast_bf16")
    return y


@on_tensor_device
def preprocess_u8(hwc: torch.Tensor, normalise: bool = False) -> torch.Tensor:
    """uint8 [N,H,W,3] (or [H,W,3]) on the device -> fp32 [N,3,H,W]: /255 and, if `normalise`, (x - mean) / std with the
    ImageNet constants, bit-identical to the reference's ToTensor -> Normalize (sam2/utils/transforms.py:27-46)."""
    lib = _lib.load()
    if hwc.dim() == 3:
        hwc = hwc[None]
    if hwc.dtype != torch.uint8 or hwc.dim() != 4 or hwc.shape[-1] != 3 or not hwc.is_cuda:
        raise ValueError(f"preprocess_u8: expected a CUDA uint8 [N,H,W,3] tensor, got {hwc.dtype} {tuple(hwc.shape)} on {hwc.device}")
    hwc = hwc.contiguous()
    n, h, w, _ = hwc.shape
    y = torch.empty((n, 3, h, w), device=hwc.device, dtype=torch.float32)
    _lib.check(lib.sam2b_op_preprocess_u8(_ptr(hwc), _ptr(y), n, h, w, int(bool(normalise)), _stream()), "op_preprocess_u8")
    return y


@on_tensor_device
def resize_frames(src: torch.Tensor, size=(1024, 1024), normalise: bool = False) -> torch.Tensor:
    """Frames at any resolution -> fp32 [N,3,H_out,W_out] by antialiased bilinear interpolation, the operator behind the
    reference's Resize((1024, 1024)) (sam2/utils/transforms.py:30; sam2_image_predictor.py:193).  src: CUDA uint8 [N,H,W,3]
    (raw frames: x / 255 first, as the reference does) or CUDA fp32 [N,3,H,W]; with `normalise`, (x - mean) / std follows."""
    lib = _lib.load()
    if not src.is_cuda or src.dim() != 4:
        raise ValueError(f"resize_frames: expected a 4-D CUDA tensor, got {tuple(src.shape)} on {src.device}")
    u8 = src.dtype == torch.uint8
    if u8 and src.shape[-1] != 3:
        raise ValueError(f"resize_frames: uint8 frames must be [N,H,W,3], got {tuple(src.shape)}")
    if not u8 and (src.dtype != torch.float32 or src.shape[1] != 3):
        raise ValueError(f"resize_frames: float frames must be fp32 [N,3,H,W], got {src.dtype} {tuple(src.shape)}")
    src = src.contiguous()
    n = src.shape[0]
    h, w = (src.shape[1], src.shape[2]) if u8 else (src.shape[2], src.shape[3])
    oh, ow = int(size[0]), int(size[1])
    out = torch.empty((n, 3, oh, ow), device=src.device, dtype=torch.float32)
    ws = torch.empty(int(lib.sam2b_op_resize_frames_workspace(n, h, w, oh, ow)), device=src.device, dtype=torch.uint8)
    _lib.check(lib.sam2b_op_resize_frames(_ptr(src), int(u8), n, h, w, _ptr(out), oh, ow, int(bool(normalise)), _ptr(ws), _stream()),
               "op_resize_frames")
    return out


@on_tensor_device
def maxpool2x2_nhwc(x: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    B, H, W, Cd = x.shape
    y = torch.empty((B, H // 2, W // 2, Cd), device=x.device, dtype=torch.float32)
    _lib.check(lib.sam2b_op_maxpool2x2_nhwc(_ptr(x.contiguous()), _ptr(y), B, H, W, Cd, _stream()), "op_maxpool")
    return y


@on_tensor_device
def window_attention(qkv: torch.Tensor, heads: int, ws: int, q_pool: bool) -> torch.Tensor:
    """qkv [B,H,W,3*heads*72] bf16 -> [B,Ho,Wo,heads*72] bf16  (ws == H: global attention)."""
    lib = _lib.load()
    B, H, W, C3 = qkv.shape
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous() and C3 == 3 * heads * 72
    Ho, Wo = (H // 2, W // 2) if q_pool else (H, W)
    out = torch.empty((B, Ho, Wo, heads * 72), device=qkv.device, dtype=torch.bfloat16)
    _lib.check(lib.sam2b_op_window_attention(_ptr(qkv), _ptr(out), B, H, W, heads, ws, int(q_pool), _stream()),
               "op_window_attention")
    return out


@on_tensor_device
def softmax_rows(s: torch.Tensor, ldp: Optional[int] = None) -> torch.Tensor:
    lib = _lib.load()
    rows, cols = s.shape
    ldp = ldp or ((cols + 63) // 64 * 64)
    p = torch.empty((rows, ldp), device=s.device, dtype=torch.bfloat16)
    _lib.check(lib.sam2b_op_softmax_rows(_ptr(s), s.stride(0), _ptr(p), ldp, rows, cols, _stream()), "op_softmax_rows")
    return p


@on_tensor_device
def attention_hd256(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """softmax(q k^T / 16) v for q [n,Tq,256], k / v [n,Lk,256] bf16 (single head; the MemoryAttention core)."""
    lib = _lib.load()
    assert q.dtype == k.dtype == v.dtype == torch.bfloat16 and q.is_cuda
    q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
    n, tq, d = q.shape
    lk = k.shape[1]
    assert d == 256 and k.shape == v.shape == (n, lk, 256)
    out = torch.empty_like(q)
    ws = torch.empty(int(lib.sam2b_op_attention_hd256_workspace(n, tq, lk)), device=q.device, dtype=torch.uint8)
    _lib.check(lib.sam2b_op_attention_hd256(_ptr(q), _ptr(k), _ptr(v), n, tq, lk, _ptr(out), _ptr(ws), _stream()),
               "op_attention_hd256")
    return out


@on_tensor_device
def attention_hd256_v64(q: torch.Tensor, k: torch.Tensor, v64: torch.Tensor, lk_rows: Optional[torch.Tensor] = None) -> torch.Tensor:
    """softmax(q k^T / 16) v64 with 64-wide values (cross-attention with v_proj folded behind it): q [n,tq,256], k [n,lk,256],
    v64 [n,lk,64] bf16 -> [n,tq,64] bf16.  lk_rows: optional int32 CUDA tensor [n] of per-object key counts."""
    lib = _lib.load()
    assert q.dtype == k.dtype == v64.dtype == torch.bfloat16 and q.is_cuda
    q, k, v64 = q.contiguous(), k.contiguous(), v64.contiguous()
    n, tq, _ = q.shape
    lk = k.shape[1]
    assert k.shape == (n, lk, 256) and v64.shape == (n, lk, 64) and q.shape[2] == 256
    if lk_rows is not None:
        assert lk_rows.dtype == torch.int32 and lk_rows.is_cuda and lk_rows.numel() == n
    out = torch.empty((n, tq, 64), device=q.device, dtype=torch.bfloat16)
    ws = torch.empty(int(lib.sam2b_op_attention_hd256_v64_workspace(n, tq, lk)), device=q.device, dtype=torch.uint8)
    _lib.check(lib.sam2b_op_attention_hd256_v64(_ptr(q), _ptr(k), _ptr(v64), n, tq, lk, _ptr(lk_rows), _ptr(out), _ptr(ws), _stream()),
               "op_attention_hd256_v64")
    return out


@on_tensor_device
def gemm_ln_producer(a: torch.Tensor, b: torch.Tensor, bias: torch.Tensor, residual: torch.Tensor, block_n: int = 0):
    """x = a @ b^T + bias + residual (fp32) -> (x, bf16(x), stats [M, parts, 2] partial (sum, sum x^2) per row)."""
    lib = _lib.load()
    M, K = a.shape
    N = b.shape[0]
    x = torch.empty((M, N), device=a.device, dtype=torch.float32)
    xb = torch.empty((M, N), device=a.device, dtype=torch.bfloat16)
    cap = 4 * ((N + 31) // 32)
    stats = torch.zeros((M, cap, 2), device=a.device, dtype=torch.float32)
    parts = C.c_int(0)
    _lib.check(lib.sam2b_op_gemm_ln_producer(_ptr(a), _ptr(b), M, N, K, _ptr(bias), _ptr(residual), _ptr(x), _ptr(xb),
                                             _ptr(stats), cap, C.byref(parts), int(block_n), _stream()), "op_gemm_ln_producer")
    # the kernel wrote a dense [M, parts, 2] array at the start of the buffer
    return x, xb, stats.flatten()[: M * parts.value * 2].view(M, parts.value, 2)


@on_tensor_device
def gemm_ln_consumer(xb: torch.Tensor, w_scaled: torch.Tensor, bias: torch.Tensor, stats: torch.Tensor,
                     eps: float, act=None, out_dtype=torch.bfloat16, block_n: int = 0) -> torch.Tensor:
    """act(LayerNorm(x) W^T + b) from raw bf16 rows xb: w_scaled = bf16(W diag(gamma) with zero-mean rows), bias = b + W beta."""
    lib = _lib.load()
    M, K = xb.shape
    N = w_scaled.shape[0]
    assert stats.is_contiguous() and stats.shape[0] == M
    out = torch.empty((M, N), device=xb.device, dtype=out_dtype)
    _lib.check(lib.sam2b_op_gemm_ln_consumer(_ptr(xb), _ptr(w_scaled), M, N, K, _ptr(bias), _ptr(stats),
                                             stats.shape[1], float(eps), ACT[act],
                                             _ptr(out if out_dtype == torch.float32 else None),
                                             _ptr(out if out_dtype == torch.bfloat16 else None), int(block_n), _stream()),
               "op_gemm_ln_consumer")
    return out
