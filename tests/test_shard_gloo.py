"""N > 1 host logic on CPU: world_size-2 gloo processes run the same sharding / gather code the B200 ranks run
under torchrun + NCCL (sam2_opt_b200/shard.py).  No GPU, no compute kernels: the encode function is a stand-in."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sam2_opt_b200 import shard


def test_split_range_balanced_and_complete():
    for n in (0, 1, 7, 8, 32, 33, 255, 256):
        for world in (1, 2, 3, 4, 8):
            got = [shard.split_range(n, world, r) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == n
            assert all(got[i][1] == got[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in got]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.split_range(4, 2, 2)


def test_shard_videos_objects_config4():
    # BASELINE config 4: 8 videos x 16 objects over 1/2/4/8 GPUs, plus 16 GPUs (objects split)
    for world in (1, 2, 4, 8, 16):
        seen = set()
        for r in range(world):
            for v, lo, hi in shard.shard_videos_objects(8, 16, world, r):
                for o in range(lo, hi):
                    assert (v, o) not in seen
                    seen.add((v, o))
        assert len(seen) == 8 * 16
    assert shard.shard_videos_objects(8, 16, 16, 5) == [(2, 8, 16)]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_images, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        images = torch.randn(n_images, 3, 8, 8, generator=g)          # same global batch on every rank

        def encode(x):     # stand-in for Engine.image_encode: three "levels" that depend on the image content
            return x.mean((2, 3)), x[:, :, ::2, ::2] * 2.0, x.flatten(1)[:, :5]

        outs = shard.encode_images_sharded(encode, images)
        ref = encode(images)
        ok = all(torch.equal(a, b) for a, b in zip(outs, ref))
        # ragged gather on its own
        counts = [3, 1][:world] if world == 2 else [1] * world
        x = torch.full((counts[rank], 2), float(rank))
        gathered = shard.all_gather_batch(x, counts)
        ok = ok and gathered.shape[0] == sum(counts) and gathered[:counts[0]].eq(0).all().item()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_images", [4, 5])
def test_encode_images_sharded_gloo_world2(n_images):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_images, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def _gather_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # the collective ("nccl") mode of the embedding gather runs unchanged over gloo on host tensors: the "engine" writes
        # its three levels into local_out(k) (this rank's slice of the gather buffers), publish(k) completes the other slices
        B = 1
        g = shard.EmbeddingGather(B, "cpu", mode="nccl")
        ok = g.mode == "nccl" and g.world == world and "all three levels" in g.describe()
        for k in range(2):
            for lvl, t in enumerate(g.local_out(k)):
                t.fill_(100.0 * k + 10.0 * rank + lvl)
            g.publish(k)
            g.wait(k)
            for lvl, full in enumerate(g.gathered(k)):
                ok = ok and tuple(full.shape) == (world * B,) + shard.LEVEL_SHAPES[lvl]
                for r in range(world):
                    ok = ok and bool((full[r * B:(r + 1) * B] == 100.0 * k + 10.0 * r + lvl).all())
        g.drain()
        g.close()
        # a global batch smaller than the world: the rank with an empty slice contributes zero rows
        images = torch.arange(3 * 4 * 4, dtype=torch.float32).reshape(1, 3, 4, 4)
        def encode_nonempty(x):      # like the engine (sam2b_image_encode rejects n <= 0): never called on an empty slice
            assert x.shape[0] > 0
            return x * 2.0, x.flatten(1)

        outs = shard.encode_images_sharded(encode_nonempty, images)
        ok = ok and torch.equal(outs[0], images * 2.0) and torch.equal(outs[1], images.flatten(1))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def test_embedding_gather_collective_mode_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]


def test_embedding_gather_single_process_is_local_only():
    g = shard.EmbeddingGather(2, "cpu", mode="auto")
    assert g.mode == "none" and g.describe() == "none"
    outs = g.local_out(5)
    assert [tuple(t.shape) for t in outs] == [(2,) + s for s in shard.LEVEL_SHAPES]
    g.publish(5); g.wait(5); g.drain(); g.close()
