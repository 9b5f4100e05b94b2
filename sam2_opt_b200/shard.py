"""Host-side sharding of the hot path over the GPUs of one box: the path shards embarrassingly —
images of a batch, videos, and objects of a video are independent — so every rank runs the same kernels on its own
slice and the only collective is the gather of embeddings / masks (north_star: "NCCL over NVLink used only to gather
embeddings and masks").  One process per GPU under torchrun; `gloo` on CPU exercises exactly the same code.

Reference counterpart: the reference has no multi-GPU path at all (one process, one GPU,
`video_multi_thread.py:36-88` shares one predictor between two threads); this is the B200-box replacement for it.
"""
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def split_range(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) slice of `n` units for `rank`: the first n % world ranks get one extra unit."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} of world {world}")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_videos_objects(n_videos: int, n_objects: int, world: int, rank: int) -> List[Tuple[int, int, int]]:
    """BASELINE config 4 (8 videos x 16 objects over 1/2/4/8 GPUs): returns [(video, obj_lo, obj_hi)] for `rank`.
    Videos go to ranks first; when there are more ranks than videos, each video's objects are split over
    world // n_videos ranks (the frame embedding is then recomputed per rank: 1.8 TFLOP < an idle GPU)."""
    if world <= n_videos:
        lo, hi = split_range(n_videos, world, rank)
        return [(v, 0, n_objects) for v in range(lo, hi)]
    per_video = world // n_videos
    video = rank // per_video
    if video >= n_videos:          # left-over ranks when world is not a multiple of n_videos
        return []
    lo, hi = split_range(n_objects, per_video, rank % per_video)
    return [(video, lo, hi)] if hi > lo else []


def all_gather_batch(x: torch.Tensor, counts: Sequence[int], group=None) -> torch.Tensor:
    """Gathers per-rank batches `x` [n_r, ...] (n_r = counts[rank]) into [sum(counts), ...] on every rank.
    Equal counts use one `all_gather_into_tensor` (a single NCCL ring / NVLS op); ragged counts pad to the maximum."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if len(counts) != world or x.shape[0] != counts[rank]:
        raise ValueError(f"all_gather_batch: counts {list(counts)} do not match world {world} / local batch {x.shape[0]}")
    x = x.contiguous()
    if len(set(counts)) == 1:
        out = torch.empty((world * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=group)
        return out
    mx = max(counts)
    pad = torch.zeros((mx,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    pad[: x.shape[0]] = x
    out = torch.empty((world * mx,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * mx: r * mx + counts[r]] for r in range(world)], 0)


def encode_images_sharded(encode_fn, images: torch.Tensor, group=None):
    """Config 5: `images` [N,3,1024,1024] is the GLOBAL batch (same tensor on every rank, or at least this rank's slice
    valid); each rank encodes its contiguous slice with `encode_fn(slice) -> tuple of [n_r, ...] tensors` and every
    output level is all-gathered back to [N, ...] on all ranks."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    n = images.shape[0]
    counts = [split_range(n, world, r)[1] - split_range(n, world, r)[0] for r in range(world)]
    lo, hi = split_range(n, world, rank)
    outs = encode_fn(images[lo:hi]) if hi > lo else None
    if min(counts) == 0:
        # fewer images than ranks: the engine refuses an empty batch (sam2b_image_encode: n <= 0), so a rank without images
        # does not call it and contributes zero rows; it learns the levels' shapes from rank 0, which always has an image
        meta = [[(tuple(o.shape[1:]), o.dtype) for o in outs] if rank == 0 else None]
        dist.broadcast_object_list(meta, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        if outs is None:
            outs = tuple(torch.empty((0,) + shp, dtype=dt, device=images.device) for shp, dt in meta[0])
    return tuple(all_gather_batch(o, counts, group) for o in outs)


# ---------------------------------------------------------------------------------------------------------------------
# The embedding gather of the batch-sharded image encoder (BASELINE config 5 / bench.py --gpus N)
# ---------------------------------------------------------------------------------------------------------------------
LEVEL_SHAPES = ((32, 256, 256), (64, 128, 128), (256, 64, 64))     # feat0, feat1, feat2 per image (fp32 NCHW), 16.8 MB


class _DevArray:
    """Device memory owned by libsam2b (cudaMalloc), exposed to torch through the CUDA array interface."""

    def __init__(self, ptr: int, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class EmbeddingGather:
    """All three feature levels of every rank's batch, gathered onto every rank, overlapped with the next step.

    mode "p2p" (what "auto" picks when peer memory is available): `slots` gather buffers per level live in device memory
      allocated by the library and mapped into every peer through CUDA IPC.  `local_out(k)` hands the engine this rank's
      slice of slot k % slots, so the neck GEMMs write their results where the collective reads them (no staging copy);
      `publish(k)` pushes that slice into every peer's buffer on a side stream with copy-engine transfers over NVLink (no
      SM is taken from the next step's persistent kernels) and closes with a one-element NCCL all-reduce on the same stream:
      its completion on rank r proves every peer's pushes of step k have landed in r's buffer (`wait(k)` / `drain()` make
      the compute stream wait for it).  With three slots a slot is re-written by step k + 3, after the arrival barriers of
      steps k + 1 and k + 2, i.e. after every rank has finished the compute that followed its use of step k's result.
    mode "nccl": one in-place `all_gather_into_tensor` per level on the compute stream (exposed; the fallback).
    mode "none" (or a single process): local buffers only."""

    def __init__(self, batch: int, device, mode: str = "auto", group=None, slots: int = 3):
        if mode not in ("auto", "p2p", "nccl", "none"):
            raise ValueError(f"EmbeddingGather: unknown mode {mode!r}")
        self.batch, self.device = int(batch), torch.device(device)
        self.group = group
        dist_on = dist.is_available() and dist.is_initialized() and mode != "none"
        self.world = dist.get_world_size(group) if dist_on else 1
        self.rank = dist.get_rank(group) if dist_on else 0
        self._lib = None
        self._owned, self._mapped = [], []
        self._err = None
        self._pending = {}
        self.mode = "none" if self.world == 1 else mode
        if self.mode in ("auto", "p2p"):
            # every collective below is reached by every rank whether or not its own setup succeeded
            if self._setup_p2p(slots):
                self.mode = "p2p"
            elif mode == "p2p":
                raise RuntimeError(f"EmbeddingGather: peer memory unavailable ({self._err})")
            else:
                self.mode = "nccl"
        if self.mode != "p2p":
            n = self.world * self.batch if self.mode == "nccl" else self.batch
            self.slots = 1
            self.bufs = [[torch.empty((n,) + s, device=self.device, dtype=torch.float32) for s in LEVEL_SHAPES]]

    # -- p2p plumbing --------------------------------------------------------------------------------------------------
    def _setup_p2p(self, slots) -> bool:
        import ctypes as C
        from . import _lib
        idx = self.device.index
        n = self.world * self.batch
        handles, bufs, local_ptr = [], [], []
        ok = True
        try:
            self._lib = lib = _lib.load()
            for s in range(slots):
                row, ptrs = [], []
                for shp in LEVEL_SHAPES:
                    numel = n * shp[0] * shp[1] * shp[2]
                    p, h = C.c_void_p(), C.create_string_buffer(64)
                    _lib.check(lib.sam2b_peer_alloc(idx, numel * 4, C.byref(p), h), "peer_alloc")
                    self._owned.append(p.value)
                    handles.append(h.raw)
                    ptrs.append(p.value)
                    row.append(torch.as_tensor(_DevArray(p.value, (n,) + shp), device=self.device))
                bufs.append(row)
                local_ptr.append(ptrs)
        except Exception as ex:
            ok, self._err = False, f"{type(ex).__name__}: {ex}"[:200]
        everyone = [None] * self.world
        dist.all_gather_object(everyone, (ok, handles), group=self.group)
        ok = all(e[0] for e in everyone)
        peer_ptr = {}                          # rank -> [slot][level] mapped pointer
        if ok:
            try:
                for r, (_, hs) in enumerate(everyone):
                    if r == self.rank:
                        continue
                    flat = []
                    for h in hs:
                        p = C.c_void_p()
                        _lib.check(self._lib.sam2b_peer_open(idx, h, C.byref(p)), f"peer_open(rank {r})")
                        self._mapped.append(p.value)
                        flat.append(p.value)
                    peer_ptr[r] = [flat[s * len(LEVEL_SHAPES):(s + 1) * len(LEVEL_SHAPES)] for s in range(slots)]
            except Exception as ex:
                ok, self._err = False, f"{type(ex).__name__}: {ex}"[:200]
        elif self._err is None:
            self._err = "a peer could not allocate / export its gather buffer"
        agree = torch.tensor([1.0 if ok else 0.0], device=self.device)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
        if agree.item() < 1.0:
            if self._err is None:
                self._err = "a peer could not map the gather buffers"
            bufs = None
            self._teardown_p2p()
            return False
        self.slots, self.bufs, self._local_ptr, self._peer_ptr = slots, bufs, local_ptr, peer_ptr
        self._comm = torch.cuda.Stream(device=self.device)
        self._flag = torch.zeros(1, device=self.device)
        self._ready = [torch.cuda.Event() for _ in range(slots)]
        self._arrived = [torch.cuda.Event() for _ in range(slots)]
        return True

    def _teardown_p2p(self):
        if self._lib is None:
            return
        idx = self.device.index
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        self.bufs = []
        for p in self._mapped:
            self._lib.sam2b_peer_close(idx, p)
        self._mapped = []
        for p in self._owned:
            self._lib.sam2b_peer_free(idx, p)
        self._owned = []

    # -- per step ------------------------------------------------------------------------------------------------------
    def local_out(self, k: int):
        """The three tensors the engine writes step k's results into: this rank's slice of the gather buffers."""
        row = self.bufs[k % self.slots]
        if self.mode == "none":
            return tuple(row)
        if self.mode == "p2p" and k in self._pending:
            raise RuntimeError("EmbeddingGather: slot reused before its previous gather was waited for")
        lo, hi = self.rank * self.batch, (self.rank + 1) * self.batch
        return tuple(t[lo:hi] for t in row)

    def gathered(self, k: int):
        """The full [world * batch, ...] tensors of step k (valid on the compute stream after `wait(k)`)."""
        return tuple(self.bufs[k % self.slots])

    def publish(self, k: int):
        """Start the gather of step k (whose local slice was produced on the current stream)."""
        if self.mode == "none":
            return
        s = k % self.slots
        if self.mode == "nccl":
            for full, mine in zip(self.bufs[0], self.local_out(k)):
                dist.all_gather_into_tensor(full, mine, group=self.group)      # in place: `mine` is a slice of `full`
            return
        import ctypes as C
        from . import _lib
        cur = torch.cuda.current_stream(self.device)
        self._ready[s].record(cur)
        self._comm.wait_event(self._ready[s])
        cs = C.c_void_p(self._comm.cuda_stream)
        with torch.cuda.stream(self._comm):
            for d in range(1, self.world):                  # rotated peer order: at any time every rank targets another peer
                r = (self.rank + d) % self.world
                for lvl, shp in enumerate(LEVEL_SHAPES):
                    per_rank = self.batch * shp[0] * shp[1] * shp[2] * 4
                    off = self.rank * per_rank
                    _lib.check(self._lib.sam2b_peer_copy(C.c_void_p(self._peer_ptr[r][s][lvl] + off),
                                                         C.c_void_p(self._local_ptr[s][lvl] + off), per_rank, cs), "peer_copy")
            dist.all_reduce(self._flag, group=self.group)    # arrival barrier: all pushes of step k precede it on every rank
            self._arrived[s].record(self._comm)
        self._pending[k] = s
        for old in [x for x in self._pending if x <= k - self.slots + 1]:
            # the slot about to be handed out next must not still be in flight: make the compute stream wait for it
            self.wait(old)

    def wait(self, k: int):
        """Make the current stream wait until step k's gather is complete on this rank."""
        s = self._pending.pop(k, None)
        if s is not None:
            torch.cuda.current_stream(self.device).wait_event(self._arrived[s])

    def drain(self):
        for k in sorted(self._pending):
            self.wait(k)

    def describe(self) -> str:
        if self.mode == "none":
            return "none"
        per_rank = self.batch * sum(a * b * c for a, b, c in LEVEL_SHAPES) * 4
        what = f"all three levels (feat0, feat1, feat2; {per_rank / 1e6:.0f} MB per rank per step) gathered onto every rank in the timed region"
        if self.mode == "p2p":
            return (what + ": neck writes into the local gather buffer, slices pushed to the peers' IPC-mapped buffers by copy-engine "
                    f"transfers over NVLink on a side stream (no SM), {self.slots} slots, one-element NCCL all-reduce as the arrival barrier; overlaps the next step")
        return what + ": in-place nccl all_gather_into_tensor per level on the compute stream (exposed)" + (f" [p2p unavailable: {self._err}]" if self._err else "")

    def close(self):
        self.drain()
        if self.mode == "p2p":
            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)       # nobody unmaps while a peer may still push
            self._teardown_p2p()
        self.bufs = []
