"""Python owner of one ``sam2b`` engine handle: uploads a state_dict (+ derived constants) and exposes the three
sub-graph calls on torch CUDA tensors.  torch is plumbing here (device memory + the caller's current stream);
all arithmetic happens inside libsam2b.so."""
import ctypes as C
import threading
from typing import Dict, Optional

import torch

from . import _lib, weights


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(0 if t is None else t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (sam2_opt_b200 has no CPU path)")
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.float().contiguous()
    return t


class Engine:
    """One engine per (model, device).  Thread-safe: calls are serialised with a lock like the reference executor
    (ref: sam2/ytools/tensorrt/tensorrt_executor.py:24,248,291) and enqueue on the caller's current stream."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("sam2_opt_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        self._lock = threading.Lock()
        h = C.c_void_p()
        _lib.check(self.lib.sam2b_create(self.device.index or 0, C.byref(h)), "create")
        self._h = h
        self._upload(state_dict)
        self._pos_cache = {}

    def _stream(self):
        """the caller's current stream ON THE ENGINE'S DEVICE (which need not be the calling thread's current device)"""
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- weights ---------------------------------------------------------------------------------------------
    def _upload(self, sd: Dict[str, torch.Tensor]):
        with torch.no_grad():
            sd = {k: v.detach().to(self.device, torch.float32).contiguous() for k, v in sd.items()
                  if torch.is_tensor(v) and v.is_floating_point()}
            sd.update(weights.derive(sd))
            for name, t in sd.items():
                shape = (C.c_int64 * max(1, t.dim()))(*(t.shape if t.dim() else (1,)))
                _lib.check(self.lib.sam2b_set_tensor(self._h, name.encode(), _ptr(t), shape, max(1, t.dim())),
                           f"set_tensor({name})")
            torch.cuda.synchronize(self.device)
            _lib.check(self.lib.sam2b_finalize(self._h), "finalize")
        self.no_obj_embed_spatial = sd.get("no_obj_embed_spatial")

    # -- constants -------------------------------------------------------------------------------------------
    def position_encoding(self, channels: int, size: int, n: int) -> torch.Tensor:
        """Sine position encoding of a (size x size) level, expanded (not copied) to batch n."""
        key = (channels, size)
        if key not in self._pos_cache:
            self._pos_cache[key] = weights.position_sine(channels, size, size, self.device)
        return self._pos_cache[key][None].expand(n, -1, -1, -1)

    # -- E1 / E2 ---------------------------------------------------------------------------------------------
    def image_encode(self, img: torch.Tensor, normalised: bool, add_no_mem_embed: bool):
        """img [n,3,1024,1024] fp32 CUDA -> (fpn0 [n,32,256,256], fpn1 [n,64,128,128], fpn2 [n,256,64,64])."""
        img = _f32c(img, "img")
        if img.dim() != 4 or tuple(img.shape[1:]) != (3, 1024, 1024):
            raise ValueError(f"image_encode expects [n,3,1024,1024], got {tuple(img.shape)}")
        n = img.shape[0]
        f0 = torch.empty((n, 32, 256, 256), device=self.device, dtype=torch.float32)
        f1 = torch.empty((n, 64, 128, 128), device=self.device, dtype=torch.float32)
        f2 = torch.empty((n, 256, 64, 64), device=self.device, dtype=torch.float32)
        with self._lock:
            _lib.check(self.lib.sam2b_image_encode(self._h, _ptr(img), n, int(normalised), int(add_no_mem_embed),
                                                   _ptr(f0), _ptr(f1), _ptr(f2), self._stream()), "image_encode")
        return f0, f1, f2

    def image_encode_host(self, img: torch.Tensor, normalised: bool, add_no_mem_embed: bool, out=None):
        """Same through HOST tensors (pinned recommended): copies overlap compute inside the library."""
        assert not img.is_cuda and img.dtype == torch.float32 and img.is_contiguous()
        n = img.shape[0]
        if out is None:
            out = (torch.empty((n, 32, 256, 256), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 64, 128, 128), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 256, 64, 64), dtype=torch.float32).pin_memory())
        with self._lock:
            _lib.check(self.lib.sam2b_image_encode_host(self._h, _ptr(img), n, int(normalised), int(add_no_mem_embed),
                                                        _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), self._stream()),
                       "image_encode_host")
        return out

    def image_encode_host_u8(self, frames: torch.Tensor, add_no_mem_embed: bool, out=None):
        """Raw HOST frames uint8 [n,1024,1024,3] (pinned recommended) -> the three HOST feature levels.  The frames are
        uploaded as bytes and ToTensor -> Normalize (ref: sam2/utils/transforms.py:27-46) runs on the device."""
        if frames.is_cuda or frames.dtype != torch.uint8 or not frames.is_contiguous() or frames.dim() != 4 \
                or tuple(frames.shape[1:]) != (1024, 1024, 3):
            raise ValueError(f"image_encode_host_u8 expects a contiguous host uint8 [n,1024,1024,3] tensor, got {frames.dtype} {tuple(frames.shape)}")
        n = frames.shape[0]
        if out is None:
            out = (torch.empty((n, 32, 256, 256), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 64, 128, 128), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 256, 64, 64), dtype=torch.float32).pin_memory())
        with self._lock:
            _lib.check(self.lib.sam2b_image_encode_host_u8(self._h, _ptr(frames), n, int(add_no_mem_embed),
                                                           _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), self._stream()),
                       "image_encode_host_u8")
        return out

    # -- E3 --------------------------------------------------------------------------------------------------
    def memory_attention(self, curr, memory, curr_pos, memory_pos, memory_exclude=None, memory_pos_exclude=None):
        curr, curr_pos = _f32c(curr, "curr"), _f32c(curr_pos, "curr_pos")
        memory, memory_pos = _f32c(memory, "memory"), _f32c(memory_pos, "memory_pos")
        L, n = memory.shape[0], curr.shape[1]
        P = 0 if memory_exclude is None else memory_exclude.shape[0]
        if P:
            memory_exclude, memory_pos_exclude = _f32c(memory_exclude, "memory_exclude"), _f32c(memory_pos_exclude, "memory_pos_exclude")
        out = torch.empty_like(curr)
        with self._lock:
            _lib.check(self.lib.sam2b_memory_attention(
                self._h, _ptr(curr), _ptr(curr_pos), _ptr(memory), _ptr(memory_pos), L,
                _ptr(memory_exclude if P else None), _ptr(memory_pos_exclude if P else None), P, n, _ptr(out), self._stream()),
                "memory_attention")
        return out

    # -- E4 --------------------------------------------------------------------------------------------------
    def memory_encode(self, pix_feat, masks):
        pix_feat, masks = _f32c(pix_feat, "pix_feat"), _f32c(masks, "masks")
        n = pix_feat.shape[0]
        x = torch.empty((n, 64, 64, 64), device=self.device, dtype=torch.float32)
        with self._lock:
            _lib.check(self.lib.sam2b_memory_encode(self._h, _ptr(pix_feat), _ptr(masks), n, _ptr(x), self._stream()),
                       "memory_encode")
        return x, self.position_encoding(64, 64, n)

    # -- life cycle ------------------------------------------------------------------------------------------
    def release(self):
        """Idempotent (ref: TensorRTExecutor.Release also runs from __del__, tensorrt_executor.py:311-334)."""
        h, self._h = getattr(self, "_h", None), None
        if h:
            self.lib.sam2b_destroy(h)

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def launch_count() -> int:
    return int(_lib.load().sam2b_launch_count())


def profile_begin():
    _lib.check(_lib.load().sam2b_profile_begin(), "profile_begin")


def profile_end():
    """-> list of (name, ms, flops, bytes), one per kernel launch since profile_begin()."""
    lib = _lib.load()
    need = C.c_int64(0)
    buf = C.create_string_buffer(1 << 24)
    _lib.check(lib.sam2b_profile_end(buf, len(buf), C.byref(need)), "profile_end")
    rows = []
    for line in buf.value.decode().splitlines()[1:]:
        name, ms, fl, by = line.rsplit(",", 3)
        rows.append((name, float(ms), float(fl), float(by)))
    return rows
