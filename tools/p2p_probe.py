#!/usr/bin/env python
"""Dev tool (torchrun, N >= 2): bandwidth of the copy-engine pushes of shard.EmbeddingGather in isolation, next to NCCL's
all_gather of the same bytes.  python -m torch.distributed.run --nproc-per-node 2 tools/p2p_probe.py"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200 import _lib, shard  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
dist.init_process_group("nccl", device_id=dev)
B = 32
g = shard.EmbeddingGather(B, dev, mode="p2p")
lib = _lib.load()
per_rank = B * sum(a * b * c for a, b, c in shard.LEVEL_SHAPES) * 4
print(rank, "can_access_peer", [torch.cuda.can_device_access_peer(local, j) for j in range(world) if j != local], flush=True)


def timeit(fn, reps=5):
    dist.barrier(); torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


def pushes():
    cs = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for d in range(1, world):
        r = (rank + d) % world
        for lvl, shp in enumerate(shard.LEVEL_SHAPES):
            nbytes = B * shp[0] * shp[1] * shp[2] * 4
            off = rank * nbytes
            lib.sam2b_peer_copy(C.c_void_p(g._peer_ptr[r][0][lvl] + off), C.c_void_p(g._local_ptr[0][lvl] + off), nbytes, cs)


def publish_wait(k=[0]):
    g.publish(k[0]); g.wait(k[0]); k[0] += 1


def local_copy():
    cs = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for lvl, shp in enumerate(shard.LEVEL_SHAPES):
        nbytes = B * shp[0] * shp[1] * shp[2] * 4
        lib.sam2b_peer_copy(C.c_void_p(g._local_ptr[1][lvl]), C.c_void_p(g._local_ptr[0][lvl]), nbytes, cs)


full = [torch.empty((world * B,) + s, device=dev) for s in shard.LEVEL_SHAPES]


def nccl():
    for f in full:
        dist.all_gather_into_tensor(f, f[rank * B:(rank + 1) * B])


flag = torch.zeros(1, device=dev)
for name, fn in (("peer pushes only", pushes), ("publish + wait (pushes + arrival all-reduce)", publish_wait), ("local D2D copy of the same bytes", local_copy),
                 ("nccl all_gather_into_tensor x3", nccl), ("1-element all_reduce", lambda: dist.all_reduce(flag))):
    fn(); torch.cuda.synchronize()
    ms = timeit(fn)
    if rank == 0:
        sent = per_rank * (world - 1)
        print(f"{name:50s} {ms:8.3f} ms   ({sent / 1e6:.0f} MB out per rank -> {sent / ms / 1e6:.1f} GB/s)", flush=True)
g.close()
dist.destroy_process_group()
