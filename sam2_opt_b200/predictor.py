"""Standalone predictors with the reference's public API, running the hot path on the sm_100a engine.

``SAM2ImagePredictor`` / ``SAM2VideoPredictor`` mirror the reference classes' user-facing calls
(ref: sam2/sam2/sam2_image_predictor.py:141-202,388-589; sam2/sam2/sam2_video_predictor_official.py:148-205,267-399,
586-736,843-945) over a plain ``state_dict`` with the reference's tensor names, so a user without the reference package
installed (e.g. the GPU box) still has ``set_image`` / ``predict`` / ``init_state`` / ``add_new_points_or_box`` /
``propagate_in_video`` and the ``speedup()`` / ``release()`` switch.  The image encoder, memory attention and memory
encoder run inside libsam2b.so, and so do the prompt encoder / mask decoder / pointer head (``Engine.sam_heads``; the fp32 torch
restatement in ``sam_heads.py`` remains as ``native_heads=False`` and as what the GPU tests compare against); the memory-bank
assembly is device-resident rings plus, in steady state, one CUDA graph per tracked frame.

Differences from the reference that are deliberate: all objects of a video are batched through the engine in one call
per frame (the reference loops objects with batch 1, sam2_video_predictor_official.py:689-725), and frame embeddings are
produced a few frames ahead in one batched encoder call.
"""
import itertools
import threading
from collections import OrderedDict
from typing import Dict, List, Optional

import numpy as np
import torch
import torch.nn.functional as F

from . import ops
from . import sam_heads as H
from .postprocess import upsample_masks
from .executor import B200Executor

IMG_MEAN = (0.485, 0.456, 0.406)
IMG_STD = (0.229, 0.224, 0.225)


class _Base:
    def __init__(self, state_dict: Dict[str, torch.Tensor], device="cuda", backend="b200", native_heads=True):
        from .engine import normalise_device
        self.device = normalise_device(device)     # "cuda" -> the current device, once (torchrun: cuda:LOCAL_RANK)
        self.w = {k: v.detach().to(self.device, torch.float32).contiguous() for k, v in state_dict.items() if torch.is_tensor(v)}
        self.engine = None
        self.backend_contexts: List[B200Executor] = []
        self._dense_pe = H.dense_pe(self.w)
        # prompt encoder / mask decoder / pointer head inside libsam2b (Engine.sam_heads); False keeps the
        # fp32 torch restatement (sam_heads.py), which the GPU tests compare against
        self.native_heads = bool(native_heads)
        self.speedup(backend)

    def speedup(self, backend="b200", use_cache=True, model_root_path=None):
        """Same switch as the reference (ref: sam2_image_predictor.py:94-138).  "b200" builds the engine from the
        current weights (re-snapshotting them if called again).  This package ships no torch implementation of the hot
        path, so "torch" / "tensorrt" / "onnxruntime" raise instead of silently running something else."""
        for ctx in self.backend_contexts:
            ctx.Release()
        self.backend_contexts = []
        getattr(self, "_heads_graphs", {}).clear()      # captured SAM-heads graphs (video predictor)
        getattr(self, "_track_graphs", {}).clear()      # captured tracked-frame graphs
        if self.engine is not None:
            self.engine.release()
            self.engine = None
        if backend is None:
            return
        if backend.lower() not in ("b200", "sm100", "sm_100a", "cuda"):
            raise Exception(f"unsupported runtime backend={backend} (the reference torch/TensorRT paths live in the reference package)")
        from .engine import Engine
        self.engine = Engine(self.w, self.device)
        self.backend_contexts = [B200Executor(self.engine, k) for k in
                                 ("set_image_e2e", "inference_image", "memory_attention_none", "memory_attention_exclude", "memory_encoder")]
        self._ex = {c.kind: c for c in self.backend_contexts}

    def release(self):
        self.speedup(None)

    def _prep_image(self, image, normalise: bool = False) -> torch.Tensor:
        """HWC image (uint8, or float in [0,255]) -> [1,3,1024,1024] fp32 in [0,1] (ref: sam2_image_predictor.py:173-196: /255,
        then Resize((1024, 1024)) = antialiased bilinear).  Both steps run in libsam2b: `preprocess_u8` at the model
        resolution (the resize is the identity there), `resize_frames` otherwise; `normalise` appends (x - mean) / std."""
        img = torch.as_tensor(np.ascontiguousarray(image)).to(self.device)
        if img.dim() != 3 or img.shape[-1] != 3:
            raise ValueError(f"expected an HWC image with 3 channels, got {tuple(img.shape)}")
        if img.dtype == torch.uint8:
            if tuple(img.shape[:2]) == (1024, 1024):
                return ops.preprocess_u8(img, normalise=normalise)         # one pass: bytes in, fp32 planes out
            return ops.resize_frames(img[None], (1024, 1024), normalise=normalise)
        chw = (img.permute(2, 0, 1)[None].float() / 255.0).contiguous()
        return ops.resize_frames(chw, (1024, 1024), normalise=normalise)


class SAM2ImagePredictor(_Base):
    def __init__(self, state_dict, device="cuda", mask_threshold=0.0, backend="b200", native_heads=True):
        super().__init__(state_dict, device, backend, native_heads)
        self.mask_threshold = mask_threshold
        self.reset_predictor()

    def reset_predictor(self):
        self._is_image_set, self._features, self._orig_hw, self._is_batch = False, None, None, False

    @torch.no_grad()
    def set_image(self, image: np.ndarray):
        self.reset_predictor()
        self._orig_hw = [image.shape[:2]]
        f0, f1, f2 = self._ex["set_image_e2e"].Inference([self._prep_image(image)], output_type="torch")
        self._features = {"image_embed": f2, "high_res_feats": [f0, f1]}
        self._is_image_set = True

    @torch.no_grad()
    def set_image_batch(self, image_list: List[np.ndarray]):
        self.reset_predictor()
        assert isinstance(image_list, list)
        self._orig_hw = [im.shape[:2] for im in image_list]
        batch = torch.cat([self._prep_image(im) for im in image_list], 0)
        f0, f1, f2 = self._ex["set_image_e2e"].Inference([batch], output_type="torch")
        self._features = {"image_embed": f2, "high_res_feats": [f0, f1]}
        self._is_image_set, self._is_batch = True, True

    def get_image_embedding(self) -> torch.Tensor:
        if not self._is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) to generate an embedding.")
        return self._features["image_embed"]

    @torch.no_grad()
    def predict(self, point_coords=None, point_labels=None, box=None, mask_input=None, multimask_output=True,
                return_logits=False, normalize_coords=True, img_idx=0):
        """ref: sam2_image_predictor.py:388-589.  Returns (masks [C,H,W], iou_predictions [C], low_res_masks [C,256,256])."""
        if not self._is_image_set:
            raise RuntimeError("An image must be set with .set_image(...) before mask prediction.")
        oh, ow = self._orig_hw[img_idx]
        coords = labels = None
        if point_coords is not None:
            assert point_labels is not None, "point_labels must be supplied if point_coords is supplied."
            coords = torch.as_tensor(np.asarray(point_coords), dtype=torch.float32, device=self.device).clone().reshape(-1, 2)
            if normalize_coords:
                coords[:, 0] = coords[:, 0] / ow * 1024
                coords[:, 1] = coords[:, 1] / oh * 1024
            coords = coords[None]
            labels = torch.as_tensor(np.asarray(point_labels), dtype=torch.int32, device=self.device).reshape(1, -1)
        if box is not None:
            b = torch.as_tensor(np.asarray(box), dtype=torch.float32, device=self.device).clone().reshape(-1, 2, 2)
            if normalize_coords:
                b[..., 0] = b[..., 0] / ow * 1024
                b[..., 1] = b[..., 1] / oh * 1024
            bl = torch.tensor([[2, 3]], dtype=torch.int32, device=self.device).repeat(b.shape[0], 1)
            coords = b if coords is None else torch.cat([b, coords], 1)
            labels = bl if labels is None else torch.cat([bl, labels], 1)
        mask_t = None
        if mask_input is not None:
            mask_t = torch.as_tensor(mask_input, dtype=torch.float32, device=self.device)
            if mask_t.dim() == 3:
                mask_t = mask_t[None]
        hr = [f[img_idx:img_idx + 1] for f in self._features["high_res_feats"]]
        emb = self._features["image_embed"][img_idx:img_idx + 1]
        if self.native_heads:
            # prompt encoder (points) + mask decoder inside libsam2b (sam2b_sam_heads, decoder-only semantics); a mask prompt
            # enters as the dense embedding of the reference's mask_downscaling convolutions; without any point the decoder sees
            # no sparse token at all (prompt_encoder.py:197-211), which the library takes as an empty set of sparse embeddings
            nb = 1 if coords is None else coords.shape[0]
            dense = None if mask_t is None else self.engine.mask_downscale(mask_t).expand(nb, -1, -1, -1)
            sparse = torch.empty((1, 0, H.D_MODEL), device=self.device) if coords is None else None
            low, _, iou, _, _, _, _ = self.engine.sam_heads(emb.expand(nb, -1, -1, -1), hr, coords, labels, dense,
                                                            multimask_output=multimask_output, decoder_only=True, high_res=False,
                                                            sparse_embeddings=sparse)
        else:
            sparse, dense = H.prompt_encoder(self.w, coords, labels, mask_t)
            batched = coords is not None and coords.shape[0] > 1
            low, iou, _, _ = H.mask_decoder(self.w, emb, sparse, dense, multimask_output, hr, repeat_image=batched, pe=self._dense_pe)
        # postprocess_masks (ref: utils/transforms.py:119) and the threshold in one pass: sam2b_op_upsample_masks
        if return_logits:
            masks = upsample_masks(low.float(), (oh, ow))
        else:
            masks = upsample_masks(low.float(), (oh, ow), threshold=self.mask_threshold, return_logits=False)
        low = torch.clamp(low, -32.0, 32.0)
        return masks[0].float().cpu().numpy(), iou[0].float().cpu().numpy(), low[0].float().cpu().numpy()


class _GraphedHeads:
    """the SAM heads of a tracked (un-prompted) frame — prompt encoder with the padding point, two-way
    transformer, upscaling, hyper-network MLPs, best-IoU select, x4 bilinear, obj_ptr_proj (sam_heads.forward_sam_heads) — are
    ~130 small fp32 kernels whose launch cost, not their arithmetic, bounds a frame with many objects.  The shapes depend on
    the number of objects only, so the whole sequence is captured once per object count into a CUDA graph and replayed:
    identical kernels and arithmetic, one launch.  Inputs are copied into the graph's static buffers; the outputs that
    outlive the frame (low-res masks, object pointers, object scores) are cloned out."""

    def __init__(self, w, n, pe, device):
        self.pix = torch.zeros(n, H.D_MODEL, 64, 64, device=device)
        self.hr = [torch.zeros(1, 32, 256, 256, device=device), torch.zeros(1, 64, 128, 128, device=device)]
        run = lambda: H.forward_sam_heads(w, self.pix, high_res_features=[h.expand(n, -1, -1, -1) for h in self.hr],
                                          multimask_output=True, pe=pe)
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(side):
            for _ in range(2):
                run()                       # warm-up outside capture (cuBLAS workspaces, lazy module loads)
        torch.cuda.current_stream(device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        # thread_local: other host threads (the reference's video_multi_thread.py drives one predictor from two) may keep
        # issuing CUDA work while this thread captures
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.out = run()

    def __call__(self, pix, hr):
        self.pix.copy_(pix)
        self.hr[0].copy_(hr[0])
        self.hr[1].copy_(hr[1])
        self.graph.replay()
        _, _, _, low, high, ptr, obj = self.out
        return low.clone(), high, ptr.clone(), obj.clone()      # `high` is consumed (memory encoder) before the next replay


class _BankRing:
    """the spatial memory bank of one group of n objects as a device-resident ring in the E3 input layout.

    The reference re-gathers the bank on every frame for every object (sam2_base_official.py:839-884: .to(device), flatten,
    permute, `torch.cat` over up to 7 frames of [4096, n, 64] features and as many position tensors).  Between consecutive frames
    the bank changes by ONE frame: the newest tracked frame enters, the oldest leaves.  Attention is invariant to the order of
    the keys as long as each key keeps its own position encoding, so the slots need not stay in temporal order: the new frame
    overwrites the slot of the frame that left (one transposing copy of [n,64,64,64] bf16 into [4096, n, 64] fp32) and the
    position tensor is rebuilt in one pass from constants: the sine encoding (identical for every frame and object) plus the
    temporal encoding of each slot's CURRENT age (`maskmem_tpos_enc`, :881-883) - the only thing that changes for the frames
    that stay.  Values are identical to `sam_heads.assemble_memory`; only the slot order (hence the fp32 summation order inside
    the attention) differs."""

    def __init__(self, w, n, device, capacity=H.NUM_MASKMEM + 1):
        self.w, self.n, self.device = w, n, device
        self.mem = torch.empty((capacity, 4096, n, H.MEM_DIM), device=device, dtype=torch.float32)
        self.pos = torch.empty_like(self.mem)
        self.frames = []                        # (frame tag, record stamp) held by slot i; conditioning frames are tagged ("c", t)
        self.pe = None                          # [4096, 1, 64] sine encoding, taken from the first stored frame
        self._idx = {}                          # tuple of slot ages -> device index tensor (no host-to-device copy per frame)

    def _write(self, slot, rec):
        self.mem[slot].copy_(rec["maskmem_features"].flatten(2).permute(2, 0, 1))     # bf16 [n,64,4096] -> fp32 [4096,n,64]
        if self.pe is None:
            self.pe = rec["maskmem_pos"][:1].flatten(2).permute(2, 0, 1).contiguous()

    def update(self, t, cond_b, non_cond_b):
        """Bring the ring to the bank of frame t (ref: the selection rule of sam2_base_official.py:829-868: every conditioning
        frame, then the tracked frames t-6 .. t-1 that exist) and return (memory [L,4096,n,64], memory_pos [L,4096,n,64])."""
        # a slot is identified by (frame, stamp of the stored record): a frame that was re-tracked or corrected since is new
        want = {(("c", tt), rec["stamp"]): (0, rec) for tt, rec in cond_b.items()}
        for t_pos in range(1, H.NUM_MASKMEM):
            tt = t - (H.NUM_MASKMEM - t_pos)
            rec = non_cond_b.get(tt)
            if rec is not None:
                want[(tt, rec["stamp"])] = (t_pos, rec)
        if len(want) > self.mem.shape[0]:       # many conditioning frames: grow once
            extra = len(want) - self.mem.shape[0]
            self.mem = torch.cat([self.mem, torch.empty((extra,) + tuple(self.mem.shape[1:]), device=self.device)], 0)
            self.pos = torch.empty_like(self.mem)
        # slots of frames that left are re-used by the newcomers; the surviving frames keep their slots (steady state: exactly
        # one frame leaves and one enters)
        free = [i for i, f in enumerate(self.frames) if f not in want]
        new = [f for f in want if f not in self.frames]
        for f in new:
            if free:
                slot = free.pop(0)
                self.frames[slot] = f
            else:
                slot = len(self.frames)
                self.frames.append(f)
            self._write(slot, want[f][1])
        if free:                                # the bank shrank (re-prompting): drop the holes, moving the tail down
            for slot in sorted(free, reverse=True):
                last = len(self.frames) - 1
                if slot != last:
                    self.mem[slot].copy_(self.mem[last])
                    self.frames[slot] = self.frames[last]
                self.frames.pop()
        L = len(self.frames)
        # memory_pos = sine encoding + temporal encoding of each slot's current age (tpos index NUM_MASKMEM - t_pos - 1)
        ages = tuple(H.NUM_MASKMEM - want[f][0] - 1 for f in self.frames)
        idx = self._idx.get(ages)
        if idx is None:
            idx = self._idx[ages] = torch.tensor(ages, device=self.device)
        tpos = self.w["maskmem_tpos_enc"].reshape(H.NUM_MASKMEM, 1, 1, H.MEM_DIM).index_select(0, idx)
        torch.add(self.pe[None].expand(L, 4096, self.n, H.MEM_DIM), tpos.expand(L, 4096, self.n, H.MEM_DIM), out=self.pos[:L])
        return self.mem[:L], self.pos[:L]


class _GraphedTrack:
    """the tracked frame of one group of n objects as ONE CUDA graph.

    Every tracked frame after a single conditioning frame runs the same ~150 kernels: memory_pos assembly, E3, the SAM heads,
    E4 and the bank update.  At one object the GPU needs 1.2 ms for them while the host needs 1.9 ms to enqueue them (4.5 ms
    while the bank is still filling and every frame has a new shape; tools/frame_timeline.py), so the sequence is captured once
    and replayed.  What differs from frame to frame is DATA on the device: the bank lives in a ring `mem [7,4096,n,64]` (slot 0
    the conditioning frame t_c, frame t in slot 1 + (t - t_c - 1) % 6), the object pointers in `ptrs [16,n,256]` (slot 0 the
    conditioning frame, frame t in slot 1 + (t - t_c - 1) % 15), and a small control vector (one pinned H2D copy per frame)
    carries each slot's current age / temporal distance, the slots the new frame overwrites and HOW MANY frames / pointers the
    bank holds: E3 always runs at L = 7, P = 64 with the per-object counts read on the device
    (sam2b_memory_attention_varlen_dev), so the frames right after the prompt (1 .. 6 bank frames, 1 .. 15 pointers: the
    filled slots are a prefix of the rings) replay the same graph.  Attention does not depend on the order of its keys as long
    as each keeps its own position term, so the values equal `sam_heads.assemble_memory` up to fp32 summation order
    (ref: sam2_base_official.py:825-950)."""

    TABLE_ROWS = 2049          # projected temporal encodings of the pointer distances 0 .. 2048 (longer videos: a larger table)

    def __init__(self, vp, n):
        dev, w = vp.device, vp.w
        self.vp, self.n = vp, n
        self.mem = torch.empty((H.NUM_MASKMEM, 4096, n, H.MEM_DIM), device=dev)
        self.pos = torch.empty_like(self.mem)
        self.ptrs = torch.empty((16, n, H.D_MODEL), device=dev)
        self.pe = torch.empty((4096, 1, H.MEM_DIM), device=dev)           # sine encoding of the bank frames
        self.table = torch.zeros((self.TABLE_ROWS, H.MEM_DIM), device=dev)
        self.tpos = w["maskmem_tpos_enc"].reshape(H.NUM_MASKMEM, 1, 1, H.MEM_DIM)
        self.ctl = torch.zeros(27, dtype=torch.int64, device=dev)          # ages[7] | distances[16] | mem slot | ptr slot | L | P
        # pinned staging ring for the control vector: slot i is rewritten only after the copy that last read it has completed
        # (an event per slot; a caller that does not synchronise per frame lets the host run frames ahead of the device)
        self.ctl_host = [torch.zeros(27, dtype=torch.int64).pin_memory() for _ in range(4)]
        self.ctl_done = [None] * len(self.ctl_host)
        self.counts = torch.zeros((2, n), dtype=torch.int32, device=dev)   # E3's per-object (frames, pointer tokens)
        self.f2 = torch.empty((1, 256, 64, 64), device=dev)
        self.hr = [torch.empty((1, 32, 256, 256), device=dev), torch.empty((1, 64, 128, 128), device=dev)]
        self.pos2 = vp.engine.position_encoding(256, 64, 1).flatten(2).permute(2, 0, 1).expand(-1, n, -1).contiguous()
        self.graph, self.out, self.gen, self.warm = None, None, -1, False
        self.owner, self.next_t, self.t_cond, self.served = None, -1, 0, 0

    def bind(self, owner, num_frames, t_cond, cond_rec, recs, t):
        """Load the state of one video at frame t (the captured graph and the static buffers are kept across videos)."""
        rows = lambda m: m.flatten(2).permute(2, 0, 1)                    # [n,64,64,64] -> [4096,n,64]
        self.owner, self.next_t, self.t_cond, self.served = owner, t, t_cond, 0
        self.mem[0].copy_(rows(cond_rec["maskmem_features"]))
        self.ptrs[0].copy_(cond_rec["obj_ptr"])
        self.pe.copy_(rows(cond_rec["maskmem_pos"][:1]))
        for tt in range(max(t_cond + 1, t - 6), t):
            self.mem[1 + (tt - t_cond - 1) % 6].copy_(rows(recs[tt]["maskmem_features"]))
        for tt in range(max(t_cond + 1, t - 15), t):
            self.ptrs[1 + (tt - t_cond - 1) % 15].copy_(recs[tt]["obj_ptr"])
        table = H._ptr_tpos_table(self.vp.w, num_frames, 16, self.vp._tpos_cache)  # [num_frames + 1, 64]
        if table.shape[0] > self.table.shape[0]:
            self.table = torch.zeros((table.shape[0], H.MEM_DIM), device=self.vp.device)
            self.graph = None                                              # the captured address is stale
        self.table[:table.shape[0]].copy_(table)

    def _control(self, t):
        k0 = t - self.t_cond - 1                                           # tracked frames since the conditioning frame
        ages = [H.NUM_MASKMEM - 1]                                         # conditioning frame: t_pos 0
        for s in range(1, H.NUM_MASKMEM):
            # slot s holds frame tt = t - 1 - r, the latest one with (tt - t_c - 1) = s - 1 (mod 6); t_pos = 7 - (t - tt) -> index r
            ages.append((k0 - s) % 6 if (s <= k0 or k0 >= 6) else 0)       # (unfilled slots: any valid index, never read)
        dists = [t - self.t_cond]
        for p in range(1, 16):
            dists.append(1 + (k0 - p) % 15 if (p <= k0 or k0 >= 15) else 0)
        return ages + dists + [1 + k0 % 6, 1 + k0 % 15, 1 + min(k0, 6), 4 * (1 + min(k0, 15))]

    def _body(self):
        vp, n = self.vp, self.n
        torch.add(self.pe[None].expand(H.NUM_MASKMEM, 4096, n, H.MEM_DIM),
                  self.tpos.index_select(0, self.ctl[0:7]).expand(H.NUM_MASKMEM, 4096, n, H.MEM_DIM), out=self.pos)
        ptr_tok = self.ptrs.reshape(16, n, 4, H.MEM_DIM).permute(0, 2, 1, 3).flatten(0, 1)                    # [64,n,64]
        ppos = self.table.index_select(0, self.ctl[7:23]).unsqueeze(1).expand(-1, n, H.MEM_DIM).repeat_interleave(4, dim=0)
        curr = self.f2.flatten(2).permute(2, 0, 1).expand(-1, n, -1)
        self.counts.copy_(self.ctl[25:27, None].expand(2, n))              # int64 -> int32, one row of L and one of P
        pix = vp.engine.memory_attention(curr, self.mem, self.pos2, self.pos, ptr_tok, ppos, counts_dev=self.counts)
        pix = pix.permute(1, 2, 0).reshape(n, H.D_MODEL, 64, 64)
        _, _, _, low, _, ptr, obj = vp.engine.sam_heads(pix, self.hr, multimask_output=True, high_res=False)
        mf, mp = vp._encode_memory_low(self.f2, low, obj, False)
        # the new frame replaces the oldest one in both rings (after everything that read them)
        self.mem.index_copy_(0, self.ctl[23:24], mf.float().flatten(2).permute(2, 0, 1)[None])
        self.ptrs.index_copy_(0, self.ctl[24:25], ptr[None])
        return low, ptr, obj, mf, mp

    def step(self, t, f2, hr):
        assert t == self.next_t
        vp = self.vp
        i = t % len(self.ctl_host)
        if self.ctl_done[i] is not None:
            self.ctl_done[i].synchronize()
        h = self.ctl_host[i]
        h.copy_(torch.tensor(self._control(t), dtype=torch.int64))
        self.ctl.copy_(h, non_blocking=True)
        self.ctl_done[i] = torch.cuda.Event()
        self.ctl_done[i].record(torch.cuda.current_stream(vp.device))
        self.f2.copy_(f2)
        self.hr[0].copy_(hr[0])
        self.hr[1].copy_(hr[1])
        if not self.warm:
            out = self._body()                        # eager: the real frame, and the warm-up every capture needs
            self.warm = True
        else:
            if self.graph is None or self.gen != vp.engine.workspace_generation():
                # (re-)capture: no kernel runs here; a workspace that another call re-allocated invalidates the old graph
                gen = vp.engine.workspace_generation()       # read BEFORE the capture: a re-allocation by another host thread
                side = torch.cuda.Stream(device=vp.device)    # while this one captures must invalidate the new graph as well
                side.wait_stream(torch.cuda.current_stream(vp.device))
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph, stream=side, capture_error_mode="thread_local"):
                    self.out = self._body()
                torch.cuda.current_stream(vp.device).wait_stream(side)
                self.gen = gen
            vp.engine.replay(self.graph)
            out = self.out
        self.served += 1
        self.next_t = t + 1
        low, ptr, obj, mf, mp = out
        return low.clone(), ptr.clone(), obj.clone(), mf.clone(), mp


class SAM2VideoPredictor(_Base):
    """Forward video propagation with click / box prompts (BASELINE.json configs 3 and 4)."""

    def __init__(self, state_dict, device="cuda", backend="b200", encode_ahead=16, fill_hole_area=0, graph_heads=True,
                 add_all_frames_to_correct_as_cond=False, bank_ring=True, native_heads=True, graph_track=True):
        """fill_hole_area: as the reference's constructor argument (sam2_video_predictor_official.py:25,889-894; the
        reference's builder sets 8 when its native extension is available, build_sam.py:128-129): holes of at most that
        many pixels in the low-res mask scores are filled before the masks are stored / returned.  0 = off, which is what
        the reference does without `sam2._C` and what the golden fixtures use.
        encode_ahead: frames pushed through the image encoder per engine call (one micro-batch of 16 runs at 3.0 ms per
        frame against 3.3 ms at 8; per-frame results do not depend on it — the encoder is batch-invariant, tested).
        graph_heads: replay the SAM heads of tracked frames from a CUDA graph (see _GraphedHeads)."""
        super().__init__(state_dict, device, backend, native_heads)
        self.encode_ahead = max(1, int(encode_ahead))
        self.fill_hole_area = int(fill_hole_area)
        self.graph_heads = bool(graph_heads)
        # bank_ring: keep the spatial memory bank of a group of objects in a device-resident ring in the E3 input layout
        # (see _BankRing) instead of re-assembling it with stack / cat on every frame
        self.bank_ring = bool(bank_ring)
        self._stamps = itertools.count(1)          # identity of stored bank records (see _BankRing.update)
        # ref: sam2_video_predictor_official.py:31-33 - whether a correction click on a tracked frame makes it a conditioning frame
        self.add_all_frames_to_correct_as_cond = bool(add_all_frames_to_correct_as_cond)
        # graph_track: replay the steady-state tracked frame (full bank, 16 pointers) from one CUDA graph (see _GraphedTrack)
        self.graph_track = bool(graph_track)
        self.graph_track_early_objects = 4     # up to this many objects the graph also serves the frames where the bank is filling
        self._heads_graphs = {}            # (object count, host thread) -> _GraphedHeads
        self._track_graphs = {}            # (object count, host thread) -> _GraphedTrack
        self._copy_stream = None           # host -> device copies of the next block of frames (see _upload)
        self._tpos_cache = {}              # (num_frames, max_ptrs) -> projected temporal encodings of the object pointers

    @torch.no_grad()
    def init_state(self, frames=None, video_height=None, video_width=None, offload_video_to_cpu=True):
        """frames: float tensor [T,3,1024,1024] already normalised (what the reference's loader returns,
        utils/misc.py:213-277) or a list of uint8 HWC numpy frames."""
        if isinstance(frames, (list, tuple)):
            video_height, video_width = frames[0].shape[:2]
            mean = torch.tensor(IMG_MEAN, device=self.device)[None, :, None, None]
            std = torch.tensor(IMG_STD, device=self.device)[None, :, None, None]
            frames = torch.cat([self._prep_frame(f, mean, std).cpu() for f in frames], 0)
        assert frames.dim() == 4 and tuple(frames.shape[1:]) == (3, 1024, 1024)
        st = {"images": frames if offload_video_to_cpu else frames.to(self.device), "num_frames": frames.shape[0],
              "video_height": video_height or 1024, "video_width": video_width or 1024,
              "obj_ids": [], "prompts": OrderedDict(), "cond": OrderedDict(), "non_cond": OrderedDict(),
              "feat_cache": {}, "frames_tracked": {}, "non_cond_b": {}, "cond_b": {}, "pending_non_cond": set(), "rings": {},
              "gtrack": {}}
        return st

    def reset_state(self, st):
        for k in ("prompts", "cond", "non_cond", "non_cond_b", "cond_b"):
            st[k].clear()
        st["obj_ids"], st["feat_cache"], st["frames_tracked"] = [], {}, {}
        st["pending_non_cond"].clear()
        st["rings"].clear()
        st["gtrack"].clear()
        st.pop("prefetch", None)

    def _prep_frame(self, f, mean, std):
        """one uint8 HWC frame -> normalised [1,3,1024,1024] fp32 (ref: utils/misc.py:96,161-162)"""
        return self._prep_image(f, normalise=True)

    # -- per-frame image features, produced `encode_ahead` frames at a time --------------------------------------
    def _upload(self, st, t):
        """Frames [t, t + encode_ahead) on the device + the block's end.  Host-resident frames (offload_video_to_cpu) cross PCIe
        on a side stream one block AHEAD: while block k is encoded and tracked, block k + 1 is already being copied, so the
        encoder never waits for its 200 MB of input (0.5 ms per frame at 25 GB/s when the copy sat on the compute stream)."""
        t1 = min(st["num_frames"], t + self.encode_ahead)
        if st["images"].is_cuda:
            return st["images"][t:t1], t1
        cur = torch.cuda.current_stream(self.device)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        pre = st.pop("prefetch", None)
        if pre is not None and pre[0] == (t, t1):
            cur.wait_event(pre[2])
            imgs = pre[1]
        else:
            imgs = st["images"][t:t1].to(self.device, non_blocking=True)
        if t1 < st["num_frames"]:
            t2 = min(st["num_frames"], t1 + self.encode_ahead)
            with torch.cuda.stream(self._copy_stream):
                nxt = st["images"][t1:t2].to(self.device, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            nxt.record_stream(cur)                   # allocated on the copy stream, consumed on the compute stream
            st["prefetch"] = ((t1, t2), nxt, ev)
        return imgs, t1

    def _features(self, st, t):
        if t not in st["feat_cache"]:
            imgs, t1 = self._upload(st, t)
            outs = self._ex["inference_image"].Inference([imgs], output_type="torch")
            st["feat_cache"] = {t + i: (outs[6][i:i + 1], outs[3][i:i + 1], [outs[4][i:i + 1], outs[5][i:i + 1]]) for i in range(t1 - t)}
        return st["feat_cache"][t]      # fpn2 [1,256,64,64], pos2 [1,256,64,64], [fpn0, fpn1]

    def _encode_memory(self, f2, high_res_masks, obj_logits, from_points):
        """ref: sam2_base_official.py:978-1026 + bf16 storage (sam2_video_predictor_official.py:887,939)"""
        n = high_res_masks.shape[0]
        m = (high_res_masks > 0).float() if from_points else torch.sigmoid(high_res_masks)
        x, pos = self._ex["memory_encoder"].Inference([f2.expand(n, -1, -1, -1), m * 20.0 - 10.0], output_type="torch")
        is_obj = (obj_logits > 0).float()
        x = x + (1 - is_obj[..., None, None]) * self.w["no_obj_embed_spatial"][..., None, None].expand(*x.shape)
        return x.to(torch.bfloat16), pos

    def _encode_memory_low(self, f2, low, obj_logits, from_points):
        """The same from the low-resolution logits (sam2b_memory_encode_low_res): the x4 bilinear up-sampling, the sigmoid /
        binarisation, the * 20 - 10 scaling and the no-object embedding happen inside the library; the 1024^2 mask is never
        written.  f2 [1,256,64,64] is shared by the n objects."""
        x, pos = self.engine.memory_encode_low_res(f2, low, obj_logits, binarize=from_points)
        return x.to(torch.bfloat16), pos

    def _video_res(self, st, low):
        hw = (st["video_height"], st["video_width"])
        return upsample_masks(low, hw)         # ref: _get_orig_video_res_output (sam2_video_predictor_official.py:843-862)

    @torch.no_grad()
    def add_new_points_or_box(self, st, frame_idx, obj_id, points=None, labels=None, clear_old_points=True,
                              normalize_coords=True, box=None):
        """ref: sam2_video_predictor_official.py:267-399.  Returns (frame_idx, obj_ids, video_res_masks)."""
        if obj_id not in st["obj_ids"]:
            if st["frames_tracked"]:
                raise RuntimeError(f"Cannot add new object id {obj_id} after tracking starts; call reset_state first.")
            st["obj_ids"].append(obj_id)
        pts = torch.zeros(0, 2) if points is None else torch.as_tensor(np.asarray(points), dtype=torch.float32).reshape(-1, 2)
        lab = torch.zeros(0, dtype=torch.int32) if labels is None else torch.as_tensor(np.asarray(labels), dtype=torch.int32).reshape(-1)
        if box is not None:
            if not clear_old_points:
                raise ValueError("cannot add box without clearing old points")
            pts = torch.cat([torch.as_tensor(np.asarray(box), dtype=torch.float32).reshape(2, 2), pts], 0)
            lab = torch.cat([torch.tensor([2, 3], dtype=torch.int32), lab], 0)
        pts = pts.clone()
        if normalize_coords:
            pts[:, 0] = pts[:, 0] / st["video_width"] * 1024
            pts[:, 1] = pts[:, 1] / st["video_height"] * 1024
        st["gtrack"].clear()                       # the banks change: captured steady-state trackers start over
        key = (frame_idx, obj_id)
        if not clear_old_points and key in st["prompts"]:
            pts = torch.cat([st["prompts"][key][0].cpu(), pts], 0)
            lab = torch.cat([st["prompts"][key][1].cpu(), lab], 0)
        st["prompts"][key] = (pts.to(self.device), lab.to(self.device))
        # A frame that has not been tracked yet is an initial conditioning frame (SAM on the plain image features, no memory);
        # a click on an already tracked frame corrects the tracked mask: memory-conditioned features, stored as a
        # non-conditioning output unless add_all_frames_to_correct_as_cond (ref: :328-349).
        is_init = frame_idx not in st["frames_tracked"]
        is_cond = is_init or self.add_all_frames_to_correct_as_cond
        # the previous mask logits of this object on this frame (conditioning output first, then tracked output), clamped,
        # go back into the SAM mask decoder as its mask prompt together with the new clicks (ref: :351-366)
        prev = st["cond"].get(frame_idx, {}).get(obj_id) or st["non_cond"].get(obj_id, {}).get(frame_idx)
        prev_logits = None if prev is None else torch.clamp(prev["pred_masks"], -32.0, 32.0)
        out = self._run_click_frame(st, frame_idx, obj_id, is_init, prev_logits)
        if is_cond:
            st["cond"].setdefault(frame_idx, {})[obj_id] = out
            st["cond_b"].pop(frame_idx, None)          # stacked copies of this frame's conditioning outputs are stale
        else:
            st["non_cond"].setdefault(obj_id, {})[frame_idx] = out
            st["pending_non_cond"].add((frame_idx, obj_id))
        st["non_cond_b"].pop(frame_idx, None)          # the batched record of this frame no longer matches this object
        low = torch.cat([self._latest_low(st, frame_idx, o) for o in st["obj_ids"]], 0)
        return frame_idx, list(st["obj_ids"]), self._video_res(st, low)

    def add_new_points(self, *args, **kwargs):
        """Deprecated alias kept by the reference (sam2_video_predictor_official.py:401-404)."""
        return self.add_new_points_or_box(*args, **kwargs)

    @torch.no_grad()
    def add_new_mask(self, st, frame_idx, obj_id, mask):
        """A mask prompt (what the reference's VOS tool feeds, tools/vos_inference.py:175-190).  ref:
        sam2_video_predictor_official.py:406-486 and SAM2Base._use_mask_as_output (sam2_base_official.py:496-545): the mask itself
        becomes the frame's output (+10 / -10 logits), only the object pointer comes from the SAM heads, which see the mask
        through `mask_downsample` as their dense prompt.  Returns (frame_idx, obj_ids, video_res_masks)."""
        if obj_id not in st["obj_ids"]:
            if st["frames_tracked"]:
                raise RuntimeError(f"Cannot add new object id {obj_id} after tracking starts; call reset_state first.")
            st["obj_ids"].append(obj_id)
        m = torch.as_tensor(np.asarray(mask))
        assert m.dim() == 2, "add_new_mask expects a 2-D mask"
        m = m.bool()[None, None].float().to(self.device)
        if tuple(m.shape[-2:]) != (1024, 1024):
            m = (F.interpolate(m, size=(1024, 1024), align_corners=False, mode="bilinear", antialias=True) >= 0.5).float()
        st["gtrack"].clear()
        st["prompts"].pop((frame_idx, obj_id), None)          # a mask prompt replaces the clicks of that frame (ref: :437)
        is_init = frame_idx not in st["frames_tracked"]
        is_cond = is_init or self.add_all_frames_to_correct_as_cond
        f2, _, hr = self._features(st, frame_idx)
        high = m * 20.0 - 10.0
        low = F.interpolate(high, size=(256, 256), align_corners=False, mode="bilinear", antialias=True)
        prompt = F.conv2d(m, self.w["mask_downsample.weight"], self.w["mask_downsample.bias"], stride=4)     # [1,1,256,256]
        if self.native_heads:
            dense = self.engine.mask_downscale(H.mask_prompt_256(prompt))
            ptr = self.engine.sam_heads(f2, hr, None, None, dense, multimask_output=False, high_res=False)[5]
        else:
            ptr = H.forward_sam_heads(self.w, f2, mask_inputs=prompt, high_res_features=hr, multimask_output=False, pe=self._dense_pe)[5]
        lam = (m.flatten(1) > 0).any(dim=1, keepdim=True).float()             # the mask, not the decoder, says whether the object is there
        out = {"pred_masks": self._fill_holes(low), "obj_ptr": lam * ptr + (1 - lam) * self.w["no_obj_ptr"],
               "object_score_logits": 20.0 * lam - 10.0, "maskmem_features": None, "maskmem_pos": None}
        if is_cond:
            st["cond"].setdefault(frame_idx, {})[obj_id] = out
            st["cond_b"].pop(frame_idx, None)
        else:
            st["non_cond"].setdefault(obj_id, {})[frame_idx] = out
            st["pending_non_cond"].add((frame_idx, obj_id))
        st["non_cond_b"].pop(frame_idx, None)
        low_all = torch.cat([self._latest_low(st, frame_idx, o) for o in st["obj_ids"]], 0)
        return frame_idx, list(st["obj_ids"]), self._video_res(st, low_all)

    def _latest_low(self, st, t, o):
        rec = st["cond"].get(t, {}).get(o) or st["non_cond"].get(o, {}).get(t)
        if rec is None:
            return torch.full((1, 1, 256, 256), H.NO_OBJ_SCORE, device=self.device)
        return rec["pred_masks"]

    def _run_click_frame(self, st, t, o, is_init, prev_logits):
        """One prompted frame of one object (ref: _run_single_frame_inference with batch 1 and run_mem_encoder=False,
        sam2_video_predictor_official.py:367-383; track_step, sam2_base_official.py:1028-1087)."""
        f2, pos2, hr = self._features(st, t)
        pts, lab = st["prompts"][(t, o)]
        pin = {"point_coords": pts[None], "point_labels": lab[None]}
        if is_init:
            pix = f2 + self.w["no_mem_embed"].view(1, -1, 1, 1)
        else:
            (cond_keys, hist), = self._history_groups(st, t, [o]).keys()
            pix = self._memory_conditioned(st, t, [(cond_keys, hist, [o])], f2, pos2, per_object=True)
        # ref: _use_multimask, sam2_base_official.py:1181-1189 (multimask_output_for_tracking, 0 <= points <= 1)
        multimask = pts.shape[0] <= 1
        if self.native_heads and pts.shape[0] > 0:
            dense = None if prev_logits is None else self.engine.mask_downscale(H.mask_prompt_256(prev_logits))
            _, _, _, low, _, ptr, obj = self.engine.sam_heads(pix, hr, pts[None], lab[None], dense, multimask_output=multimask,
                                                              high_res=False)
        else:
            _, _, _, low, high, ptr, obj = H.forward_sam_heads(self.w, pix, point_inputs=pin, mask_inputs=prev_logits,
                                                               high_res_features=hr, multimask_output=multimask, pe=self._dense_pe)
        return {"pred_masks": self._fill_holes(low), "obj_ptr": ptr, "object_score_logits": obj, "maskmem_features": None,
                "maskmem_pos": None}

    def _fill_holes(self, low):
        if self.fill_hole_area > 0:
            from .postprocess import fill_holes_in_mask_scores
            low = fill_holes_in_mask_scores(low, self.fill_hole_area)
        return low

    @torch.no_grad()
    def propagate_in_video_preflight(self, st):
        """Encode the memory of every prompted frame (ref: sam2_video_predictor_official.py:586-649)."""
        for t, per_obj in st["cond"].items():
            # a frame that became a conditioning frame is no longer a tracked one (ref: :640-643)
            for o in per_obj:
                if st["non_cond"].get(o, {}).pop(t, None) is not None:
                    st["non_cond_b"].pop(t, None)
            todo = [o for o, out in per_obj.items() if out["maskmem_features"] is None]
            if not todo:
                continue
            f2 = self._features(st, t)[0]
            low = torch.cat([per_obj[o]["pred_masks"] for o in todo], 0)
            obj = torch.cat([per_obj[o]["object_score_logits"] for o in todo], 0)
            if self.native_heads:
                mf, mp = self._encode_memory_low(f2, low, obj, True)
            else:
                high = F.interpolate(low, size=(1024, 1024), mode="bilinear", align_corners=False)
                mf, mp = self._encode_memory(f2, high, obj, from_points=True)
            for i, o in enumerate(todo):
                per_obj[o]["maskmem_features"], per_obj[o]["maskmem_pos"] = mf[i:i + 1], mp[i:i + 1]
            st["cond_b"].pop(t, None)
        # correction clicks on tracked frames (stored as non-conditioning outputs): their memory is encoded here as well,
        # from the binarised mask like every frame the user interacted with (ref: :608-629, is_mask_from_pts=True)
        for t, o in sorted(st["pending_non_cond"]):
            out = st["non_cond"].get(o, {}).get(t)
            if out is None or out["maskmem_features"] is not None:
                continue
            if self.native_heads:
                mf, mp = self._encode_memory_low(self._features(st, t)[0], out["pred_masks"], out["object_score_logits"], True)
            else:
                high = F.interpolate(out["pred_masks"], size=(1024, 1024), mode="bilinear", align_corners=False)
                mf, mp = self._encode_memory(self._features(st, t)[0], high, out["object_score_logits"], from_points=True)
            out["maskmem_features"], out["maskmem_pos"] = mf, mp
        st["pending_non_cond"].clear()

    @torch.no_grad()
    def propagate_in_video(self, st, start_frame_idx=None, max_frame_num_to_track=None):
        """Generator of (frame_idx, obj_ids, video_res_masks [n,1,H,W]) — ref: sam2_video_predictor_official.py:652-736."""
        self.propagate_in_video_preflight(st)
        if not st["cond"]:
            raise RuntimeError("No input points or masks are provided for any object; please add inputs first.")
        T, objs = st["num_frames"], list(st["obj_ids"])
        t0 = min(st["cond"]) if start_frame_idx is None else start_frame_idx
        t1 = T - 1 if max_frame_num_to_track is None else min(T - 1, t0 + max_frame_num_to_track)
        # One frame of look-ahead: frame t + 1 is enqueued on the GPU before frame t is handed to the caller, so the device
        # keeps working while the caller synchronises on / copies out frame t's masks (results and order are unchanged).
        pending = None
        for t in range(t0, t1 + 1):
            per_obj_low = {}
            tracked = [o for o in objs if not (t in st["cond"] and o in st["cond"][t])]
            for o in objs:
                if o not in tracked:
                    per_obj_low[o] = st["cond"][t][o]["pred_masks"]
            if tracked:
                outs = self._track_frame(st, t, tracked)
                for o, out in zip(tracked, outs):
                    st["non_cond"].setdefault(o, {})[t] = out
                    per_obj_low[o] = out["pred_masks"]
            st["frames_tracked"][t] = True
            low = torch.cat([per_obj_low[o] for o in objs], 0)
            cur = (t, objs, self._video_res(st, low))
            if pending is not None:
                yield pending
            pending = cur
        if pending is not None:
            yield pending

    def _history_groups(self, st, t, objs):
        """Objects that share their conditioning frames and their tracked history (the usual case: everything prompted on
        the same frame) form one group whose memory bank is assembled once from the per-frame BATCHED outputs
        (SURVEY section 8 f-2): no per-object float / add / stack / cat kernels.  Anything else takes the per-object path."""
        groups = OrderedDict()
        for o in objs:
            cond_keys = tuple(tt for tt, per in st["cond"].items() if o in per and per[o]["maskmem_features"] is not None)
            nc = st["non_cond"].get(o, {})
            hist = tuple(tt for tt in range(max(0, t - 16), t) if tt in nc and nc[tt]["maskmem_features"] is not None)
            groups.setdefault((cond_keys, hist), []).append(o)
        return groups

    def _group_memory(self, st, t, g_objs, per_object=False, cond_keys=None, hist=None):
        """Memory bank + pointer tokens of the objects `g_objs` (which share their history) for frame t
        (ref: _prepare_memory_conditioned_features, sam2_base_official.py:797-950) ->
        [memory [L,4096,n,64], memory_pos, ptr tokens [P,n,64], ptr pos]."""
        key = tuple(g_objs)
        if cond_keys is None:
            (cond_keys, hist), = self._history_groups(st, t, g_objs[:1]).keys()
        batched = {} if per_object else {tt: st["non_cond_b"].get(tt, {}).get(key) for tt in hist}
        if not per_object and all(v is not None for v in batched.values()):
            cond_b = {tt: self._stacked_cond(st, tt, key) for tt in cond_keys}
            if self.bank_ring:
                ring = st["rings"].get(key)
                if ring is None:
                    ring = st["rings"][key] = _BankRing(self.w, len(key), self.device)
                mem, mpos = ring.update(t, cond_b, batched)
                ptr, ppos = H.assemble_pointers(self.w, t, cond_b, batched, st["num_frames"], cache=self._tpos_cache)
                return [mem, mpos, ptr, ppos]
            return list(H.assemble_memory(self.w, t, cond_b, batched, st["num_frames"], cache=self._tpos_cache))
        per = []
        for o in g_objs:
            cond = {tt: st["cond"][tt][o] for tt in cond_keys}
            nc = {tt: r for tt, r in st["non_cond"].get(o, {}).items() if r["maskmem_features"] is not None}
            per.append(H.assemble_memory(self.w, t, cond, nc, st["num_frames"], cache=self._tpos_cache))
        return [torch.cat([m[i] for m in per], dim=2 if i < 2 else 1) for i in range(4)]

    def _memory_conditioned(self, st, t, groups, f2, pos2, per_object=False):
        """Current-frame features fused with the memory banks through boundary E3 -> [n_total,256,64,64] in the order of the
        concatenated groups.  `groups`: [(cond_keys, hist, [objects])].  Groups whose banks differ in size (objects added or
        re-prompted on different frames) still share ONE engine call: the banks are padded to the largest and the engine gets
        the per-object frame / pointer counts (sam2b_memory_attention_varlen; SURVEY section 8a "dynamic shapes")."""
        mems = [self._group_memory(st, t, objs, per_object, ck, hs) for ck, hs, objs in groups]
        sizes = [(m[0].shape[0], m[2].shape[0], m[0].shape[2]) for m in mems]          # (L, P, n) per group
        n = sum(sz[2] for sz in sizes)
        curr = f2.flatten(2).permute(2, 0, 1).expand(-1, n, -1)
        curr_pos = pos2.flatten(2).permute(2, 0, 1).expand(-1, n, -1)
        if len({sz[:2] for sz in sizes}) == 1:
            mem = mems[0] if len(mems) == 1 else [torch.cat([m[i] for m in mems], dim=2 if i < 2 else 1) for i in range(4)]
            if sizes[0][1]:
                pix = self._ex["memory_attention_exclude"].Inference([curr, mem[0], curr_pos, mem[1], mem[2], mem[3]], output_type="torch")[0]
            else:
                pix = self._ex["memory_attention_none"].Inference([curr, mem[0], curr_pos, mem[1]], output_type="torch")[0]
        else:
            Lm, Pm = max(sz[0] for sz in sizes), max(sz[1] for sz in sizes)
            mem = torch.zeros((Lm, 4096, n, H.MEM_DIM), device=self.device)
            mpos = torch.zeros_like(mem)
            ptr = torch.zeros((max(Pm, 4), n, H.MEM_DIM), device=self.device)
            ppos = torch.zeros_like(ptr)
            L_rows, P_rows, o0 = [], [], 0
            for m, (L, P, ng) in zip(mems, sizes):
                mem[:L, :, o0:o0 + ng], mpos[:L, :, o0:o0 + ng] = m[0], m[1]
                if P:
                    ptr[:P, o0:o0 + ng], ppos[:P, o0:o0 + ng] = m[2], m[3]
                L_rows += [L] * ng
                P_rows += [P] * ng
                o0 += ng
            pix = self.engine.memory_attention(curr, mem, curr_pos, mpos, ptr[:Pm] if Pm else None, ppos[:Pm] if Pm else None,
                                               L_rows=L_rows, P_rows=P_rows)
        return pix.permute(1, 2, 0).reshape(n, H.D_MODEL, 64, 64)

    def _track_frame(self, st, t, objs):
        """One non-prompted frame for a list of objects, batched through the engine (ref: track_step,
        sam2_base_official.py:1028-1179): ONE memory-attention call, ONE pass of the SAM heads and ONE memory-encoder call for
        all objects, whatever their histories."""
        f2, pos2, hr = self._features(st, t)
        groups = [(ck, hs, g) for (ck, hs), g in self._history_groups(st, t, objs).items()]
        order = [o for _, _, g in groups for o in g]
        n = len(order)
        if self.graph_track and self.native_heads and len(groups) == 1:
            outs = self._graph_track_frame(st, t, groups[0], f2, hr)
            if outs is not None:
                return [outs[o] for o in objs]
        pix = self._memory_conditioned(st, t, groups, f2, pos2)
        if self.native_heads:
            # ONE library call: ~50 launches of libsam2b's own kernels (split-bf16 tcgen05 GEMMs on the image side, fp32 token side)
            _, _, _, low, high, ptr, obj = self.engine.sam_heads(pix, hr, multimask_output=True, high_res=False)
        elif self.graph_heads:
            gk = (n, threading.get_ident())    # static buffers are per graph: one graph per host thread
            if gk not in self._heads_graphs:
                self._heads_graphs[gk] = _GraphedHeads(self.w, n, self._dense_pe, self.device)
            low, high, ptr, obj = self._heads_graphs[gk](pix, hr)
        else:
            hr_n = [h.expand(n, -1, -1, -1) for h in hr]
            _, _, _, low, high, ptr, obj = H.forward_sam_heads(self.w, pix, high_res_features=hr_n, multimask_output=True, pe=self._dense_pe)
        mf, mp = self._encode_memory_low(f2, low, obj, False) if self.native_heads else self._encode_memory(f2, high, obj, from_points=False)
        low = self._fill_holes(low)            # after the memory is encoded from the unfilled masks, as in the reference
        results, o0 = {}, 0
        for _, _, g_objs in groups:
            sl = slice(o0, o0 + len(g_objs))
            st["non_cond_b"].setdefault(t, {})[tuple(g_objs)] = {"maskmem_features": mf[sl], "maskmem_pos": mp[sl], "obj_ptr": ptr[sl],
                                                                 "stamp": next(self._stamps)}
            o0 += len(g_objs)
        for i, o in enumerate(order):
            results[o] = {"pred_masks": low[i:i + 1], "obj_ptr": ptr[i:i + 1], "object_score_logits": obj[i:i + 1],
                          "maskmem_features": mf[i:i + 1], "maskmem_pos": mp[i:i + 1]}
        return [results[o] for o in objs]

    def _graph_track_frame(self, st, t, group, f2, hr):
        """The steady-state tracked frame through _GraphedTrack, or None when the frame is not in steady state: one
        conditioning frame before t, tracked frames t-15 .. t-1 all present as batched records of this very group."""
        cond_keys, hist, g_objs = group
        key = tuple(g_objs)
        if len(cond_keys) != 1 or cond_keys[0] >= t or st["num_frames"] < 17 or len(st["cond"]) != 1:
            return None
        # while the bank is filling the graph still runs E3 at its full size (the counts only mask): worth it where the frame is
        # bound by the host (few objects), not where the GPU is the limit and a 2-frame bank is 4x cheaper than a 7-frame one
        if t - cond_keys[0] <= 15 and len(g_objs) > self.graph_track_early_objects:
            return None
        gk = (len(key), threading.get_ident())          # static buffers and graph: one per object count and host thread
        tr = self._track_graphs.get(gk)
        owner = (id(st), key)
        if tr is None or tr.owner != owner or tr.next_t != t or st["gtrack"].get(key) is not tr:
            recs = {tt: st["non_cond_b"].get(tt, {}).get(key) for tt in range(max(cond_keys[0] + 1, t - 15), t)}
            if any(r is None for r in recs.values()):
                return None
            if tr is None:
                tr = self._track_graphs[gk] = _GraphedTrack(self, len(key))
            tr.bind(owner, st["num_frames"], cond_keys[0], self._stacked_cond(st, cond_keys[0], key), recs, t)
            st["gtrack"][key] = tr
        low, ptr, obj, mf, mp = tr.step(t, f2, hr)
        st["non_cond_b"].setdefault(t, {})[key] = {"maskmem_features": mf, "maskmem_pos": mp, "obj_ptr": ptr, "stamp": next(self._stamps)}
        low = self._fill_holes(low)
        return {o: {"pred_masks": low[i:i + 1], "obj_ptr": ptr[i:i + 1], "object_score_logits": obj[i:i + 1],
                    "maskmem_features": mf[i:i + 1], "maskmem_pos": mp[i:i + 1]} for i, o in enumerate(g_objs)}

    def _stacked_cond(self, st, t, key):
        """Conditioning-frame outputs of the objects `key`, stacked once along the batch dimension."""
        c = st["cond_b"].setdefault(t, {})
        if key not in c:
            per = st["cond"][t]
            c[key] = {k: torch.cat([per[o][k] for o in key], 0) for k in ("maskmem_features", "maskmem_pos", "obj_ptr")}
            c[key]["stamp"] = next(self._stamps)
        return c[key]
