#!/usr/bin/env python
"""Dev tool (torchrun, N >= 2): per-step timeline of the batch-32 encoder with the embedding gather in its different modes."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200 import engine as E, shard  # noqa: E402
from sam2_opt_b200.synth import synth_image_norm  # noqa: E402
from sam2_opt_b200.synth_weights import make_weights  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
dist.init_process_group("nccl", device_id=dev)
eng = E.Engine(make_weights(0), dev)
B, steps = 32, 8
img = synth_image_norm(B, seed=1 + rank).to(dev)
for label, mode, do_publish in (("none", "none", False), ("p2p buffers, no publish", "p2p", False), ("p2p", "p2p", True), ("nccl", "nccl", True), ("none again", "none", False)):
    g = shard.EmbeddingGather(B, dev, mode=mode)
    k = 0
    for _ in range(3):
        eng.image_encode(img, True, False, out=g.local_out(k))
        if do_publish:
            g.publish(k)
        k += 1
    g.drain(); dist.barrier(); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 2)]
    ev[0].record()
    for s in range(steps):
        eng.image_encode(img, True, False, out=g.local_out(k))
        if do_publish:
            g.publish(k)
        k += 1
        ev[s + 1].record()
    g.drain()
    ev[steps + 1].record()
    torch.cuda.synchronize()
    per = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps + 1)]
    if rank == 0:
        print(f"{label:28s} per-step ms {[round(x, 2) for x in per[:-1]]} drain {per[-1]:.2f}  mean {sum(per) / steps:.2f}", flush=True)
    g.close()
eng.release()
dist.destroy_process_group()
