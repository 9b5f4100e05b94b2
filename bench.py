#!/usr/bin/env python
"""Headline benchmark: SAM2.1 Hiera-L image-encoder forward (BASELINE.json configs[1]: batch 32 synthetic
1024x1024 images per GPU) -> images/s.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...     # the reference algorithm's CPU path (oracle port) on the host cores

One "step" = one pass of the hot path (`sam2b_image_encode`: Hiera trunk + FpnNeck + conv_s0/s1, boundary E2) over one
batch of seeded synthetic, already-normalised images.  `value`: inputs resident in HBM, CUDA-event timed, max over ranks.
`e2e`: the same metric through the plugin's host-buffer entry point (`sam2b_image_encode_host`, pinned host tensors,
H2D of the images and D2H of all three outputs inside the timed region).  N > 1: every rank encodes its own batch
(weak scaling, batch-sharded config 5) and the embeddings (fpn2 level) are all-gathered with NCCL inside the timed
region, as north_star prescribes.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GFLOP_PER_IMAGE = 1822.8      # algorithmic work of forward_image per 1024^2 image
METRIC = "Hiera-L 1024^2 image-encoder images/sec"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        # median over samples under load (upper half of the sorted clocks excludes idle samples at the edges)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


CPU_SAMPLE_IMAGES = 4      # bounded sample of the batch-32 workload per CPU step (about 10-30 s of host work per run)


def cpu_reference_images_per_s(n_images, steps, warmup):
    """The reference algorithm's fp32 CPU path (oracle port of the reference torch code) on the host cores."""
    import torch
    from oracle import sam2_oracle as O
    from sam2_opt_b200.synth import synth_image_norm
    torch.set_grad_enabled(False)
    w = O.make_weights(0, only_prefix=("image_encoder.", "sam_mask_decoder.conv_s", "no_mem"))
    img = synth_image_norm(n_images, seed=1)
    for _ in range(warmup):
        O.forward_image(w, img[:1])
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        O.forward_image(w, img)
        times.append(time.perf_counter() - t0)
    return n_images * len(times) / sum(times), torch.get_num_threads(), times


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 5)), min(args.warmup, 1)
    v, cores, times = cpu_reference_images_per_s(CPU_SAMPLE_IMAGES, steps, warmup)
    sample = f"{steps} step(s) x {CPU_SAMPLE_IMAGES} images of the batch-{args.batch} workload (fp32 torch CPU port of the reference path), {warmup} warm-up"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"SAM2.1 Hiera-L image-encoder forward (E2), batch {args.batch} x 1024x1024 per GPU", "sample_images_per_step": CPU_SAMPLE_IMAGES},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="images per GPU per step (configs[1]: 32)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--video-frames", type=int, default=40, help="frames of the propagate_in_video leg (0 = skip it)")
    ap.add_argument("--video-objects", type=int, default=1)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from sam2_opt_b200 import build, engine as E
    from sam2_opt_b200.synth import synth_image_norm

    build.build()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    dev = torch.device(f"cuda:{local}")
    warmup = max(args.warmup, 3)

    # random-init weights of the exact architecture (no checkpoints offline): same generator the parity tests use
    from sam2_opt_b200.synth_weights import make_weights
    eng = E.Engine(make_weights(0), dev)
    B = args.batch
    img_host = synth_image_norm(B, seed=1 + rank).pin_memory()
    img = img_host.to(dev)
    outs_host = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    gathered = None
    def step():
        nonlocal gathered
        f0, f1, f2 = eng.image_encode(img, True, False)
        if world > 1:
            if gathered is None:
                gathered = torch.empty((world * B, 256, 64, 64), device=dev, dtype=torch.float32)
            dist.all_gather_into_tensor(gathered, f2)
        return f2

    for _ in range(warmup):
        step()
    barrier()
    l0 = E.launch_count()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        t0.record()
        for _ in range(args.steps):
            step()
        t1.record()
        barrier()
    launches = E.launch_count() - l0
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = ms.item()
    value = world * B * args.steps / (ms_total / 1e3)

    # ---- e2e: host buffers through the plugin entry point -------------------------------------------------
    for _ in range(2):
        outs_host = eng.image_encode_host(img_host, True, False, outs_host)
    barrier()
    e0 = time.perf_counter()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(args.steps):
        outs_host = eng.image_encode_host(img_host, True, False, outs_host)
    c1.record()
    barrier()
    e_ms = torch.tensor([max(c0.elapsed_time(c1), 0.0), (time.perf_counter() - e0) * 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(e_ms, op=dist.ReduceOp.MAX)
    e2e_value = world * B * args.steps / (max(e_ms.tolist()) / 1e3)
    h2d = img_host.numel() * 4
    d2h = sum(o.numel() * 4 for o in outs_host)

    # ---- the same from raw uint8 frames (what set_image_batch receives): bytes over PCIe, ToTensor/Normalize on the device
    raw_host = torch.randint(0, 256, (B, 1024, 1024, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(7 + rank)).pin_memory()
    for _ in range(2):
        outs_host = eng.image_encode_host_u8(raw_host, False, outs_host)
    barrier()
    u0 = time.perf_counter()
    for _ in range(args.steps):
        outs_host = eng.image_encode_host_u8(raw_host, False, outs_host)
    barrier()
    u_ms = torch.tensor([(time.perf_counter() - u0) * 1e3], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(u_ms, op=dist.ReduceOp.MAX)
    e2e_u8_value = world * B * args.steps / (u_ms.item() / 1e3)

    # ---- roofline of the dominant kernel (tcgen05 GEMM): live CUDA-event pairs around every launch of one more step
    roof = None
    if rank == 0:
        E.profile_begin()
        eng.image_encode(img, True, False)        # rank-local: no collective here (only rank 0 runs this leg)
        rows = E.profile_end()
        g = [(ms_, fl) for name, ms_, fl, by in rows if name.startswith("gemm")]
        all_ms = sum(r[1] for r in rows)
        g_ms, g_fl = sum(x[0] for x in g), sum(x[1] for x in g)
        peak, hbm, how = _peaks()
        ach = g_fl / (g_ms * 1e-3) / 1e12
        traffic, traffic_note = None, None
        tp = os.path.join(ROOT, "profiles", "r1e_traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp))
            traffic = tj["dram_bytes_per_launch_avg"]
            traffic_note = ("bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum), mean of the four stage-3 GEMM forms "
                            "at batch 16 from one ncu --set full capture: " + tj["source"])
        roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "traffic_note": traffic_note,
                "kernel": "gemm_bf16_tcgen05_kernel", "launches_per_step": len(g), "avg_launch_ms": g_ms / max(1, len(g)),
                "share_of_step_kernel_time": g_ms / all_ms, "peak_source": how,
                "flops_per_launch_avg": g_fl / max(1, len(g)),
                "note": "algorithmic 2MNK flops of every GEMM launch of one step / sum of their CUDA-event durations on the launching stream"}

    # ---- second headline metric (BASELINE.json: "video frames/sec propagated"): propagate_in_video through the public
    # predictor API, one synthetic clip and one prompted object per rank (videos shard across GPUs, config 3/4 shapes)
    video = None
    if args.video_frames > 0:
        from sam2_opt_b200.predictor import SAM2VideoPredictor
        from sam2_opt_b200.synth import synth_video
        eng.release()
        vp = SAM2VideoPredictor(make_weights(0), device=dev)
        clip = synth_video(args.video_frames, seed=1 + rank).pin_memory()
        t_steady, n_steady, t_all = 0.0, 0, 0.0
        for rep in range(2):                      # first pass warms allocator / workspaces
            st = vp.init_state(clip)
            for o in range(args.video_objects):
                g = max(1, int(args.video_objects ** 0.5 + 0.999))
                xy = [512.0, 512.0] if args.video_objects == 1 else [128 + 768.0 * (o % g) / max(1, g - 1), 128 + 768.0 * (o // g) / max(1, g - 1)]
                vp.add_new_points_or_box(st, 0, o, points=[xy], labels=[1])
            barrier()
            stamps = [time.perf_counter()]
            for t, ids, masks in vp.propagate_in_video(st):
                (masks > 0).any().item()          # the caller consumes every frame's masks (D2H read per frame)
                stamps.append(time.perf_counter())
            if rep == 1:
                t_all = stamps[-1] - stamps[0]
                s0 = min(16, len(stamps) - 2)      # steady state: full 7-frame bank + 16 object pointers from frame 16 on
                t_steady, n_steady = stamps[-1] - stamps[s0], len(stamps) - 1 - s0
        vt = torch.tensor([t_all, t_steady], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(vt, op=dist.ReduceOp.MAX)
        t_all, t_steady = vt.tolist()
        video = {"metric": "video frames/sec propagated (SAM2VideoPredictor.propagate_in_video)", "unit": "frames/s",
                 "value_steady": world * n_steady / t_steady, "value_overall": world * args.video_frames / t_all,
                 "frames_per_video": args.video_frames, "objects_per_video": args.video_objects, "videos": world,
                 "object_frames_per_s_steady": world * args.video_objects * n_steady / t_steady,
                 "timing": "host wall clock around the generator incl. H2D of frames and a D2H read of every frame's masks, max over ranks; steady = frames >= 16 (7-frame memory bank, 16 pointers)"}
        vp.release()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, times = cpu_reference_images_per_s(CPU_SAMPLE_IMAGES, 2, 1)
        cpu = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
               "sample": f"2 timed forward_image calls on {CPU_SAMPLE_IMAGES} images of the workload (fp32 torch CPU port of the reference path, all host threads), 1 single-image warm-up"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"SAM2.1 Hiera-L image-encoder forward (boundary E2: trunk+FpnNeck+conv_s0/s1), batch {B} x 3x1024x1024 fp32 per GPU, random-init weights",
                       "batch_per_gpu": B, "global_batch": B * world, "sharding": "batch" if world > 1 else "none",
                       "collective": "nccl all_gather_into_tensor(fpn2 embeddings) in the timed region" if world > 1 else "none",
                       "l2": "inputs (403 MB/step) and activations exceed the 126 MB L2; no explicit flush"},
            "tflops_algorithmic": value * GFLOP_PER_IMAGE / 1e3,
            "clocks": clk.summary(), "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "Engine.image_encode_host -> sam2b_image_encode_host (pinned host in/out)"},
            "e2e_u8_frames": {"value": e2e_u8_value, "unit": "images/s", "h2d_bytes_per_step": raw_host.numel(), "d2h_bytes_per_step": d2h,
                              "api": "Engine.image_encode_host_u8 -> sam2b_image_encode_host_u8 (raw uint8 HWC frames, pinned; host wall clock)"},
            "roofline": roof, "cpu_baseline": cpu, "video": video,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
