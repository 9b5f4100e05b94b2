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
    outs = encode_fn(images[lo:hi])
    return tuple(all_gather_batch(o, counts, group) for o in outs)
