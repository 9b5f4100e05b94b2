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


def _stream(t: Optional[torch.Tensor] = None):
    """the current stream of `t`'s device (of the current device without a tensor)"""
    return C.c_void_p(torch.cuda.current_stream(None if t is None else t.device).cuda_stream)


def on_tensor_device(fn):
    """Decorator for the single-kernel front ends (ops.py, postprocess.py): run with the device of the first tensor
    argument current, so that the launch and `_stream()` both mean that tensor's GPU whatever the caller's current device
    is (the sam2b_op_* entry points launch on the calling thread's current device)."""
    import functools

    @functools.wraps(fn)
    def wrapper(t, *a, **k):
        if torch.is_tensor(t) and t.is_cuda:
            with torch.cuda.device(t.device):
                return fn(t, *a, **k)
        return fn(t, *a, **k)
    return wrapper


def normalise_device(device) -> torch.device:
    """torch.device("cuda") has no index: bind it to the CURRENT device once, so that the engine, its output tensors and
    the streams it enqueues on all mean the same GPU (under torchrun the current device is cuda:LOCAL_RANK, not cuda:0)."""
    d = torch.device(device if device is not None else "cuda")
    if d.type != "cuda":
        raise RuntimeError(f"sam2_opt_b200 needs a CUDA (sm_100a) device, got {d}; there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device()) if d.index is None else d


def _f32c(t: torch.Tensor, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"{name}: expected a CUDA tensor (sam2_opt_b200 has no CPU path)")
    if t.dtype != torch.float32 or not t.is_contiguous():
        t = t.float().contiguous()
    return t


class Engine:
    """One engine per (model, device).  Thread-safe: calls enqueue on the caller's current stream and are serialised like
    the reference executor's (ref: sam2/ytools/tensorrt/tensorrt_executor.py:24,248-252,288-291) — on the host by a lock
    and on the device by an event chain inside the library (engine.cu `EngineCall`), so two threads with private streams
    can share one engine (video_multi_thread.py:36-60) without their kernels overlapping on the shared workspaces."""

    def __init__(self, state_dict: Dict[str, torch.Tensor], device: Optional[torch.device] = None):
        if not torch.cuda.is_available():
            raise RuntimeError("sam2_opt_b200 needs a CUDA (sm_100a) device; there is no CPU fallback")
        self.lib = _lib.load()
        self.device = normalise_device(device)
        self._lock = threading.Lock()
        h = C.c_void_p()
        _lib.check(self.lib.sam2b_create(self.device.index, C.byref(h)), "create")
        self._h = h
        self._upload(state_dict)
        self._pos_cache = {}

    def _stream(self):
        """the caller's current stream ON THE ENGINE'S DEVICE (which need not be the calling thread's current device)"""
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    # -- weights ---------------------------------------------------------------------------------------------
    def _upload(self, sd: Dict[str, torch.Tensor]):
        """Snapshot of the live state_dict (ref: the reference bakes the weights into its engine at build time).  The
        one-time folds (weights.derive) run on the HOST in fp64 and everything crosses to the device as ONE flat buffer:
        engine start-up launches no torch kernels, only the library's own bf16 repack."""
        with torch.no_grad():
            sd = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in sd.items()
                  if torch.is_tensor(v) and v.is_floating_point()}
            sd.update(weights.derive(sd))
            names = list(sd)
            offs, total = [], 0
            for k in names:
                offs.append(total)
                total += (sd[k].numel() + 63) // 64 * 64            # 256-byte aligned slices
            flat = torch.empty(total, dtype=torch.float32).pin_memory()
            for k, o in zip(names, offs):
                flat[o:o + sd[k].numel()].copy_(sd[k].reshape(-1))
            dev_flat = flat.to(self.device)                          # one H2D copy
            base = dev_flat.data_ptr()
            for k, o in zip(names, offs):
                t = sd[k]
                shape = (C.c_int64 * max(1, t.dim()))(*(t.shape if t.dim() else (1,)))
                _lib.check(self.lib.sam2b_set_tensor(self._h, k.encode(), C.c_void_p(base + 4 * o), shape, max(1, t.dim())),
                           f"set_tensor({k})")
            torch.cuda.synchronize(self.device)
            _lib.check(self.lib.sam2b_finalize(self._h), "finalize")
            del dev_flat
        nos = sd.get("no_obj_embed_spatial")
        self.no_obj_embed_spatial = None if nos is None else nos.to(self.device)

    # -- constants -------------------------------------------------------------------------------------------
    def position_encoding(self, channels: int, size: int, n: int) -> torch.Tensor:
        """Sine position encoding of a (size x size) level, expanded (not copied) to batch n."""
        key = (channels, size)
        if key not in self._pos_cache:
            pe = torch.empty((channels, size, size), device=self.device, dtype=torch.float32)
            with self._lock:
                _lib.check(self.lib.sam2b_position_encoding(channels, size, size, _ptr(pe), self._stream()), "position_encoding")
            self._pos_cache[key] = pe
        return self._pos_cache[key][None].expand(n, -1, -1, -1)

    # -- E1 / E2 ---------------------------------------------------------------------------------------------
    def image_encode(self, img: torch.Tensor, normalised: bool, add_no_mem_embed: bool, out=None):
        """img [n,3,1024,1024] fp32 CUDA -> (fpn0 [n,32,256,256], fpn1 [n,64,128,128], fpn2 [n,256,64,64]).
        out: optional caller-owned contiguous fp32 CUDA tensors of those shapes (e.g. this rank's slices of an all-gather
        buffer: the neck then writes its results straight where the collective reads them)."""
        img = _f32c(img, "img")
        if img.dim() != 4 or tuple(img.shape[1:]) != (3, 1024, 1024):
            raise ValueError(f"image_encode expects [n,3,1024,1024], got {tuple(img.shape)}")
        n = img.shape[0]
        shapes = ((n, 32, 256, 256), (n, 64, 128, 128), (n, 256, 64, 64))
        if n == 0:
            return tuple(torch.empty(s, device=self.device, dtype=torch.float32) for s in shapes)
        if out is None:
            f0, f1, f2 = (torch.empty(s, device=self.device, dtype=torch.float32) for s in shapes)
        else:
            f0, f1, f2 = out
            for t, s in zip(out, shapes):
                if tuple(t.shape) != s or t.dtype != torch.float32 or not t.is_contiguous() or t.device != self.device:
                    raise ValueError(f"image_encode: out tensors must be contiguous fp32 {shapes} on {self.device}")
        with self._lock:
            _lib.check(self.lib.sam2b_image_encode(self._h, _ptr(img), n, int(normalised), int(add_no_mem_embed),
                                                   _ptr(f0), _ptr(f1), _ptr(f2), self._stream()), "image_encode")
        return f0, f1, f2

    def image_encode_host(self, img: torch.Tensor, normalised: bool, add_no_mem_embed: bool, out=None, wait=True):
        """Same through HOST tensors (pinned recommended): copies overlap compute inside the library.
        wait=False (serving loops): return once everything is enqueued, so that the next call's uploads overlap this call's
        compute; `img` and `out` must then stay alive and untouched until `host_sync()`."""
        assert not img.is_cuda and img.dtype == torch.float32 and img.is_contiguous()
        n = img.shape[0]
        if out is None:
            out = (torch.empty((n, 32, 256, 256), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 64, 128, 128), dtype=torch.float32).pin_memory(),
                   torch.empty((n, 256, 64, 64), dtype=torch.float32).pin_memory())
        fn = self.lib.sam2b_image_encode_host if wait else self.lib.sam2b_image_encode_host_async
        with self._lock:
            _lib.check(fn(self._h, _ptr(img), n, int(normalised), int(add_no_mem_embed),
                          _ptr(out[0]), _ptr(out[1]), _ptr(out[2]), self._stream()), "image_encode_host")
        return out

    def host_sync(self):
        """Wait for every upload, kernel and download the host-buffer entry points have issued so far."""
        _lib.check(self.lib.sam2b_host_sync(self._h, self._stream()), "host_sync")

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
    def memory_attention(self, curr, memory, curr_pos, memory_pos, memory_exclude=None, memory_pos_exclude=None,
                         L_rows=None, P_rows=None, counts_dev=None):
        """E3.  L_rows / P_rows (optional sequences of n ints): objects whose banks differ share one call — the tensors are
        padded to the largest bank and object b uses memory[:L_rows[b]] / memory_exclude[:P_rows[b]] only.
        counts_dev: the same counts as a DEVICE int32 tensor [2,n] (frames, pointer tokens) — no host data, capturable."""
        curr, curr_pos = _f32c(curr, "curr"), _f32c(curr_pos, "curr_pos")
        memory, memory_pos = _f32c(memory, "memory"), _f32c(memory_pos, "memory_pos")
        L, n = memory.shape[0], curr.shape[1]
        P = 0 if memory_exclude is None else memory_exclude.shape[0]
        if P:
            memory_exclude, memory_pos_exclude = _f32c(memory_exclude, "memory_exclude"), _f32c(memory_pos_exclude, "memory_pos_exclude")
        out = torch.empty_like(curr)
        if counts_dev is not None:
            if counts_dev.dtype != torch.int32 or not counts_dev.is_cuda or tuple(counts_dev.shape) != (2, n) or not counts_dev.is_contiguous():
                raise ValueError(f"memory_attention: counts_dev must be a contiguous CUDA int32 tensor [2,{n}]")
            with self._lock:
                _lib.check(self.lib.sam2b_memory_attention_varlen_dev(
                    self._h, _ptr(curr), _ptr(curr_pos), _ptr(memory), _ptr(memory_pos), L,
                    _ptr(memory_exclude if P else None), _ptr(memory_pos_exclude if P else None), P, n, _ptr(counts_dev),
                    _ptr(out), self._stream()), "memory_attention_varlen_dev")
            return out
        if L_rows is not None or P_rows is not None:
            arr = lambda v: None if v is None else (C.c_int32 * n)(*[int(x) for x in v])
            if (L_rows is not None and len(L_rows) != n) or (P_rows is not None and len(P_rows) != n):
                raise ValueError(f"memory_attention: L_rows / P_rows must have one entry per object ({n})")
            with self._lock:
                _lib.check(self.lib.sam2b_memory_attention_varlen(
                    self._h, _ptr(curr), _ptr(curr_pos), _ptr(memory), _ptr(memory_pos), L,
                    _ptr(memory_exclude if P else None), _ptr(memory_pos_exclude if P else None), P, n, arr(L_rows), arr(P_rows),
                    _ptr(out), self._stream()), "memory_attention_varlen")
            return out
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

    def mask_downscale(self, mask_prompt):
        """PromptEncoder._embed_masks (ref: sam/prompt_encoder.py:152-155): mask prompt logits [n,1,256,256] -> the dense
        embedding [n,256,64,64] that `sam_heads(dense_embedding=...)` takes."""
        m = _f32c(mask_prompt, "mask_prompt")
        if m.dim() != 4 or tuple(m.shape[1:]) != (1, 256, 256):
            raise ValueError(f"mask_downscale expects [n,1,256,256], got {tuple(m.shape)}")
        out = torch.empty((m.shape[0], 256, 64, 64), device=self.device, dtype=torch.float32)
        _lib.check(self.lib.sam2b_mask_downscale(self._h, _ptr(m), m.shape[0], _ptr(out), self._stream()), "mask_downscale")
        return out

    def memory_encode_low_res(self, pix_feat, low_res_masks, object_score_logits=None, binarize=False):
        """E4 straight from the SAM heads' low-resolution logits [n,1,256,256] (ref: _encode_new_memory,
        sam2_base_official.py:978-1026): sigmoid(bilinear x4) * 20 - 10 (binarize: the (mask > 0) form of prompted frames) is
        evaluated inside the down-sampler's first convolution; with object_score_logits [n,1] the no-object embedding is added
        too.  pix_feat [1 or n,256,64,64].  -> (maskmem_features fp32 [n,64,64,64], maskmem_pos_enc)"""
        pix_feat, low = _f32c(pix_feat, "pix_feat"), _f32c(low_res_masks, "low_res_masks")
        n = low.shape[0]
        if tuple(low.shape[1:]) != (1, 256, 256) or pix_feat.shape[0] not in (1, n) or tuple(pix_feat.shape[1:]) != (256, 64, 64):
            raise ValueError(f"memory_encode_low_res: pix_feat {tuple(pix_feat.shape)} / low_res_masks {tuple(low.shape)}")
        obj = None if object_score_logits is None else _f32c(object_score_logits, "object_score_logits").reshape(n)
        x = torch.empty((n, 64, 64, 64), device=self.device, dtype=torch.float32)
        with self._lock:
            _lib.check(self.lib.sam2b_memory_encode_low_res(self._h, _ptr(pix_feat), 0 if pix_feat.shape[0] == 1 and n > 1 else 256 * 4096,
                                                            _ptr(low), int(bool(binarize)), _ptr(obj), n, _ptr(x), self._stream()),
                       "memory_encode_low_res")
        return x, self.position_encoding(64, 64, n)

    # -- SAM heads --------------------------------------------------------------------------
    def sam_heads(self, backbone_features, high_res_features, point_coords=None, point_labels=None, dense_embedding=None,
                  multimask_output=False, decoder_only=False, high_res=True, high_res_multimasks=False, sparse_embeddings=None,
                  mask_tokens=False):
        """Prompt encoder (points) + mask decoder + object-pointer head for n prompts / objects in one library call
        (ref: sam2_base_official.py:338-494).  backbone_features [n,256,64,64] may be any view whose channel or pixel
        dimension is contiguous (the permuted E3 output, an expanded single image ...); high_res_features = [fpn0, fpn1]
        with batch 1 (shared) or n.  point_coords [n,P,2] / point_labels [n,P] or None (a tracked frame).
        sparse_embeddings [n,S,256]: the sparse prompt embeddings themselves instead of points (no padding point is added).
        -> (low_res_multimasks [n,M,256,256], high_res_multimasks or None, ious [n,M], low_res_masks [n,1,256,256],
            high_res_masks [n,1,1024,1024] or None, obj_ptr [n,256], object_score_logits [n,1])
           + mask tokens [n,4,256] as an eighth element with mask_tokens=True"""
        bf = backbone_features
        if not bf.is_cuda or bf.dtype != torch.float32:
            raise RuntimeError("sam_heads: expected a CUDA fp32 tensor (sam2_opt_b200 has no CPU path)")
        if bf.dim() != 4 or tuple(bf.shape[1:]) != (256, 64, 64):
            raise ValueError(f"sam_heads expects backbone_features [n,256,64,64], got {tuple(bf.shape)}")
        n = bf.shape[0]
        sn, sc, sy, sx = bf.stride()
        if sy != 64 * sx or (sx != 1 and sc != 1):
            bf = bf.contiguous()
            sn, sc, sy, sx = bf.stride()
        if n == 1:
            sn = 0
        hr = []
        for f, shp in zip(high_res_features, ((32, 256, 256), (64, 128, 128))):
            if tuple(f.shape[1:]) != shp or f.shape[0] not in (1, n):
                raise ValueError(f"sam_heads: high-res feature {tuple(f.shape)} does not match [1|{n},{shp}]")
            if f.dtype != torch.float32 or f.stride()[1:] != (shp[1] * shp[2], shp[2], 1):
                f = f.float().contiguous()
            hr.append((f, 0 if f.shape[0] == 1 else f.stride(0)))
        P = 0
        if point_coords is not None:
            point_coords = _f32c(point_coords, "point_coords").reshape(n, -1, 2)
            P = point_coords.shape[1]
            point_labels = point_labels.to(torch.int32).contiguous().reshape(n, P)
        S = 0
        if sparse_embeddings is not None:
            sparse_embeddings = _f32c(sparse_embeddings, "sparse_embeddings")
            if sparse_embeddings.dim() != 3 or sparse_embeddings.shape[0] != n or sparse_embeddings.shape[2] != 256:
                raise ValueError(f"sam_heads: sparse_embeddings {tuple(sparse_embeddings.shape)} != [{n},S,256]")
            S, P = sparse_embeddings.shape[1], 0
            if S == 0:          # an empty tensor has no usable address: the library still needs "given, with 0 rows"
                sparse_embeddings = torch.zeros((n, 1, 256), device=self.device)
        if dense_embedding is not None:
            dense_embedding = _f32c(dense_embedding, "dense_embedding")
            if tuple(dense_embedding.shape) != (n, 256, 64, 64):
                raise ValueError(f"sam_heads: dense_embedding {tuple(dense_embedding.shape)} != {(n, 256, 64, 64)}")
        M = 3 if multimask_output else 1
        new = lambda *shape: torch.empty(shape, device=self.device, dtype=torch.float32)
        low_multi, ious, low, ptr, obj = new(n, M, 256, 256), new(n, M), new(n, 1, 256, 256), new(n, 256), new(n, 1)
        high = new(n, 1, 1024, 1024) if high_res else None
        high_multi = new(n, M, 1024, 1024) if high_res_multimasks else None
        toks = new(n, 4, 256) if mask_tokens else None
        with self._lock:
            _lib.check(self.lib.sam2b_sam_heads(
                self._h, _ptr(bf), sn, sc, sx, _ptr(hr[0][0]), hr[0][1], _ptr(hr[1][0]), hr[1][1],
                _ptr(point_coords if P else None), _ptr(point_labels if P else None), P, _ptr(sparse_embeddings), S,
                _ptr(dense_embedding), n, int(bool(multimask_output)), int(bool(decoder_only)), _ptr(low_multi), _ptr(high_multi),
                _ptr(ious), _ptr(low), _ptr(high), _ptr(ptr), _ptr(obj), _ptr(toks), self._stream()), "sam_heads")
        out = (low_multi, high_multi, ious, low, high, ptr, obj)
        return out + (toks,) if mask_tokens else out

    # -- CUDA graphs over engine calls -----------------------------------------------------------------------------
    def workspace_generation(self) -> int:
        """Changes whenever the library re-allocated a workspace buffer: graphs captured before hold stale addresses."""
        return int(self.lib.sam2b_workspace_generation(self._h))

    def replay(self, graph: "torch.cuda.CUDAGraph"):
        """Replay a graph captured over engine calls on the caller's current stream, serialised against every other use
        of the engine's shared workspaces like a plain call (host lock + the library's cross-stream event chain)."""
        st = self._stream()
        with self._lock:
            _lib.check(self.lib.sam2b_stream_acquire(self._h, st), "stream_acquire")
            graph.replay()
            _lib.check(self.lib.sam2b_stream_release(self._h, st), "stream_release")

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
