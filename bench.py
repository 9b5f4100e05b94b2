#!/usr/bin/env python
"""Headline benchmark: SAM2.1 Hiera-L image-encoder forward (BASELINE.json configs[1]: batch 32 synthetic
1024x1024 images per GPU) -> images/s, plus the video legs (configs[2] and configs[3]).

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...     # the UNMODIFIED reference's CPU torch path (oracle/_ref) on the host cores
    python bench.py --sweep                  # config 5 table: batch 1..256 per run, compute-only and with the gather

One "step" = one pass of the hot path (`sam2b_image_encode`: Hiera trunk + FpnNeck + conv_s0/s1, boundary E2) over one
batch of seeded synthetic, already-normalised images.  `value`: inputs resident in HBM, CUDA-event timed, max over ranks.
`e2e`: the same metric through the plugin's host-buffer entry point (`sam2b_image_encode_host`, pinned host tensors,
H2D of the images and D2H of all three outputs inside the timed region).  N > 1: every rank encodes its own batch
(weak scaling, batch-sharded config 5); all three feature levels (feat0, feat1, feat2) are gathered onto every rank inside
the timed region — the neck writes them straight into this rank's slice of the gather buffer and the transfer of step k
overlaps the trunk of step k+1 (sam2_opt_b200/shard.py: `EmbeddingGather`).
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
CPU_SAMPLE_IMAGES = 4         # bounded sample of the batch-32 workload per CPU step (about 10-30 s of host work per run)
CPU_VIDEO_FRAMES = 3          # reference propagate_in_video on the host cores: 1 prompted + 2 tracked frames


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", 1400.0), d.get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json, sustained)"
    return 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


def host_threads():
    """All the host threads this process may use.  torchrun exports OMP_NUM_THREADS=1 to its workers; a CPU baseline must
    not inherit that, so the CPU legs call torch.set_num_threads(host_threads()) explicitly."""
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


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
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# The reference on the host cores / on the same GPU (oracle/_ref = byte-identical copy of the reference sources; the oracle
# port only when that copy is missing)
# ---------------------------------------------------------------------------------------------------------------------
def _reference_model(kind, device="cpu"):
    """-> (callable forward_image(img) or None, model or None, kind string)"""
    import torch
    from oracle import reference_loader as rl
    from oracle import sam2_oracle as O
    w = O.make_weights(0)
    if rl.available():
        m = rl.build_reference(kind)
        m.load_state_dict(w, strict=True)
        return m.to(device).eval(), "reference"
    return w, "port"


def cpu_reference_images_per_s(n_images, steps, warmup):
    """The reference's own fp32 CPU path (`SAM2Base.forward_image`, sam2_base_official.py:548-564) on all host threads."""
    import torch
    from oracle import sam2_oracle as O
    from sam2_opt_b200.synth import synth_image_norm
    torch.set_grad_enabled(False)
    cores = host_threads()
    torch.set_num_threads(cores)
    m, kind = _reference_model("video")
    fwd = (lambda x: m.forward_image(x)) if kind == "reference" else (lambda x: O.forward_image(m, x))
    img = synth_image_norm(n_images, seed=1)
    for _ in range(warmup):
        fwd(img[:1])
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        fwd(img)
        times.append(time.perf_counter() - t0)
    return n_images * len(times) / sum(times), torch.get_num_threads(), times, kind


def cpu_reference_video_fps(n_frames):
    """The reference's SAM2VideoPredictor.propagate_in_video on the host cores, first `n_frames` frames of a 64-frame
    clip (clip length enters the pointer encodings only), one object, one click."""
    import numpy as np
    import torch
    from oracle import reference_loader as rl
    from sam2_opt_b200.synth import synth_video
    torch.set_grad_enabled(False)
    cores = host_threads()
    torch.set_num_threads(cores)
    frames = synth_video(n_frames, seed=1)
    if rl.available():
        m, _ = _reference_model("video")
        import sam2.sam2_video_predictor_official as vp
        vp.load_video_frames = lambda **k: (frames, 1024, 1024)
        st = m.init_state("synthetic")
        m.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
        t0 = time.perf_counter()
        n = sum(1 for _ in m.propagate_in_video(st))
        return (n - 1) / (time.perf_counter() - t0), cores, "reference"      # frame 0 is the prompted frame (served from its stored output)
    from oracle import sam2_oracle as O
    w = O.make_weights(0)
    t0 = time.perf_counter()
    n = sum(1 for _ in O.video_propagate(w, frames))
    return (n - 1) / (time.perf_counter() - t0), cores, "port"


def torch_gpu_baseline(dev, batch):
    """The "same GPU" comparison: the UNMODIFIED reference `forward_image` (cuBLAS / cuDNN /
    SDPA eager torch) on this B200, strict fp32 (TF32 off for matmul and cuDNN) and bf16 autocast (the precision precedent of
    sam2/sam2/benchmark.py:20-92), CUDA-event timed."""
    import torch
    from sam2_opt_b200.synth import synth_image_norm
    torch.set_grad_enabled(False)
    m, kind = _reference_model("video", dev)
    if kind != "reference":
        return {"unavailable": "reference sources (oracle/_ref) not present"}
    out = {"kind": "reference", "api": "SAM2Base.forward_image (unmodified reference, eager torch)", "unit": "images/s"}
    for name in ("bf16_autocast", "fp32"):
        b = batch
        while b >= 1:
            try:
                img = synth_image_norm(b, seed=1).to(dev)
                torch.backends.cuda.matmul.allow_tf32 = False
                torch.backends.cudnn.allow_tf32 = False
                ctx = torch.autocast("cuda", dtype=torch.bfloat16) if name == "bf16_autocast" else torch.autocast("cuda", enabled=False)
                reps = 3 if name == "bf16_autocast" else 2
                with ctx:
                    m.forward_image(img)
                    torch.cuda.synchronize(dev)
                    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    for _ in range(reps):
                        m.forward_image(img)
                    z.record()
                torch.cuda.synchronize(dev)
                ms = a.elapsed_time(z) / reps
                out[name] = {"value": b / (ms / 1e3), "batch": b, "ms_per_step": ms, "timed_steps": reps, "warmup": 1,
                             "tf32": False if name == "fp32" else "n/a"}
                break
            except torch.cuda.OutOfMemoryError:
                torch.cuda.empty_cache()
                b //= 2
        del img
        torch.cuda.empty_cache()
    del m
    torch.cuda.empty_cache()
    return out


def torch_gpu_video_baseline(dev, T):
    """Config 3 on the same GPU through the UNMODIFIED reference `SAM2VideoPredictor` (its own per-object loop and memory-bank
    assembly): (1) eager torch under bf16 autocast, the precision its own scripts use (sam2/sam2/benchmark.py:20-92); (2) the
    same object after `predictor.speedup("b200")` — the drop-in a user of the reference gets before switching to this package's
    batched predictor.  Host wall clock around the generator, every frame's masks read by the consumer."""
    import numpy as np
    import torch
    from oracle import reference_loader as rl
    from sam2_opt_b200.synth import synth_video
    torch.set_grad_enabled(False)
    if not rl.available():
        return {"unavailable": "reference sources (oracle/_ref) not present"}
    m, _ = _reference_model("video", dev)
    import sam2.sam2_video_predictor_official as vpm
    frames = synth_video(T, seed=1)
    old_loader = vpm.load_video_frames
    vpm.load_video_frames = lambda **k: (frames, 1024, 1024)

    def run(n_frames):
        st = m.init_state("synthetic")
        m.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
        torch.cuda.synchronize(dev)
        t0, n = time.perf_counter(), 0
        for _, _, masks in m.propagate_in_video(st, max_frame_num_to_track=n_frames - 1):
            (masks > 0).any().item()
            n += 1
        torch.cuda.synchronize(dev)
        return n / (time.perf_counter() - t0)

    out = {"kind": "reference", "api": "SAM2VideoPredictor.propagate_in_video (unmodified reference), 1 object, one click", "unit": "frames/s",
           "frames_per_video": T}
    try:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            run(10)
            out["bf16_autocast"] = run(T)
        from sam2_opt_b200 import speedup as sp
        sp.install(type(m))
        m.speedup("b200")
        run(20)
        out["speedup_b200"] = run(T)
        m.speedup("torch")
    except Exception as ex:
        out["unavailable"] = f"{type(ex).__name__}: {ex}"[:300]
    finally:
        vpm.load_video_frames = old_loader
    del m
    torch.cuda.empty_cache()
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 5)), min(args.warmup, 1)
    v, cores, times, kind = cpu_reference_images_per_s(CPU_SAMPLE_IMAGES, steps, warmup)
    what = ("the unmodified reference SAM2Base.forward_image (oracle/_ref on the GPU box), fp32 torch" if kind == "reference"
            else "fp32 torch CPU port of the reference path (oracle/sam2_oracle.py)")
    sample = f"{steps} step(s) x {CPU_SAMPLE_IMAGES} images of the batch-{args.batch} workload: {what}, {cores} host threads, {warmup} single-image warm-up"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"SAM2.1 Hiera-L image-encoder forward (E2), batch {args.batch} x 1024x1024 per GPU; CPU arm: a {CPU_SAMPLE_IMAGES}-image sample of that batch per step (CPU throughput is flat in the batch size)",
                   "sample_images_per_step": CPU_SAMPLE_IMAGES},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ---------------------------------------------------------------------------------------------------------------------
# video legs
# ---------------------------------------------------------------------------------------------------------------------
def grid_clicks(n_objects):
    """config 4: clicks on a g x g grid of points (16 objects -> 4 x 4); one object -> the image centre (config 3)."""
    if n_objects == 1:
        return [[512.0, 512.0]]
    g = max(1, int(n_objects ** 0.5 + 0.999))
    return [[128 + 768.0 * (o % g) / max(1, g - 1), 128 + 768.0 * (o // g) / max(1, g - 1)] for o in range(n_objects)]


def attention_roofline(rows, nobj=1):
    """Roofline of the video path's dominant kernel, the cross-attention over the memory bank (flash_attn_hd256_v64_tcgen05_kernel),
    from the library's per-launch CUDA-event profile: full-bank launches only (7 frames x 4096 + 64 pointer tokens = 28 736 keys)."""
    peak, _, how = _peaks()
    cross = [(ms_, fl) for name, ms_, fl, by in rows if name.startswith("attn256v64 ") and " k28736 " in name + " "]
    self_ = [(ms_, fl) for name, ms_, fl, by in rows if name.startswith("attn256 ")]
    c_ms, c_fl = sum(x[0] for x in cross), sum(x[1] for x in cross)
    s_ms, s_fl = sum(x[0] for x in self_), sum(x[1] for x in self_)
    tot = sum(r[1] for r in rows)
    issued = c_fl * 320.0 / 512.0      # the kernel issues 2 T Lk (256 + 64) flops; the reference's formulation needs 4 T Lk 256
    traffic, traffic_note = None, None
    cap = os.path.join(ROOT, "profiles", "r2n_top_kernels_summary.json")
    if nobj == 16 and os.path.exists(cap):      # the committed ncu --set full capture ran this kernel at 16 objects, 28 736 keys
        with open(cap) as f:
            v64 = [r for r in json.load(f) if "hd256_v64" in r["kernel"]]
        if v64:
            traffic = v64[-1]["dram_read_bytes"] + v64[-1]["dram_write_bytes"]
            traffic_note = ("dram__bytes_read.sum + dram__bytes_write.sum of the warm launch at 16 objects x 28 736 keys (profiles/r2n_top_kernels_summary.json, "
                            "tools/ncu_top_kernels.py); algorithmic bytes per launch: q 33.6 + k 235.4 + v 58.9 + out 8.4 = 336 MB")
    return {"bound": "tensor", "kernel": "flash_attn_hd256_v64_tcgen05_kernel", "traffic_note": traffic_note,
            "achieved": issued / (c_ms * 1e-3) / 1e12 if c_ms else None, "peak": peak, "unit": "TFLOP/s",
            "frac": (issued / (c_ms * 1e-3) / 1e12 / peak) if c_ms else None,
            "achieved_algorithmic": c_fl / (c_ms * 1e-3) / 1e12 if c_ms else None,
            "launches": len(cross), "avg_launch_ms": c_ms / max(1, len(cross)),
            "share_of_library_kernel_time": c_ms / tot if tot else None, "peak_source": how, "traffic": traffic,
            "self_attention": {"kernel": "flash_attn_hd256_tcgen05_kernel", "achieved": s_fl / (s_ms * 1e-3) / 1e12 if s_ms else None,
                               "launches": len(self_), "share_of_library_kernel_time": s_ms / tot if tot else None},
            "note": "achieved = flops the kernel issues, 2*4096*Lk*(256+64) per object and layer (v_proj is applied after P V on 64-wide "
                    "values), over the full-bank launches (Lk = 28 736) of a short enqueued clip / sum of their CUDA-event durations; "
                    "achieved_algorithmic counts the reference's formulation (4*4096*Lk*256)"}


def run_clip(vp, clip, clicks, obj_lo=0, obj_hi=None, consume=True):
    """One propagate_in_video pass; returns the per-frame host timestamps (the caller consumes every frame's masks)."""
    st = vp.init_state(clip)
    obj_hi = len(clicks) if obj_hi is None else obj_hi
    for o in range(obj_lo, obj_hi):
        vp.add_new_points_or_box(st, 0, o, points=[clicks[o]], labels=[1])
    stamps = [time.perf_counter()]
    for t, ids, masks in vp.propagate_in_video(st):
        if consume:
            (masks > 0).any().item()          # D2H read per frame
        stamps.append(time.perf_counter())
    return stamps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="images per GPU per step (configs[1]: 32)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-torch-gpu-baseline", action="store_true")
    ap.add_argument("--video-frames", type=int, default=64, help="frames per clip of the video legs (configs[2]/[3]: 64; 0 = skip them)")
    ap.add_argument("--video-objects", type=int, default=16, help="objects per video of the config-4 leg (0 = skip it)")
    ap.add_argument("--videos", type=int, default=8, help="videos of the config-4 leg, sharded over the ranks")
    ap.add_argument("--gather", default="auto", choices=["auto", "p2p", "nccl", "none"], help="N > 1: how the three feature levels are gathered")
    ap.add_argument("--sweep", action="store_true", help="config 5: batch sweep 1..256 (prints one JSON table, no headline line)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from sam2_opt_b200 import build, engine as E, shard
    from sam2_opt_b200.synth import synth_image_norm, synth_video

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
    weights = make_weights(0)
    eng = E.Engine(weights, dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(vals):
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    def timed_encode(B, steps, warm, gather_mode, sampler=None):
        """`steps` passes over a batch of B images per rank with the gather of all three levels; -> (ms total, max over ranks).
        sampler: context manager entered around the timed region only (clock / throttle sampling)."""
        import contextlib
        img = synth_image_norm(B, seed=1 + rank).to(dev)
        g = shard.EmbeddingGather(B, dev, mode=gather_mode if world > 1 else "none")
        k = 0
        for _ in range(warm):
            eng.image_encode(img, True, False, out=g.local_out(k)); g.publish(k); k += 1
        g.drain()
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with (sampler if sampler is not None else contextlib.nullcontext()):
            barrier()
            t0.record()
            for _ in range(steps):
                eng.image_encode(img, True, False, out=g.local_out(k)); g.publish(k); k += 1
            g.drain()                 # the compute stream waits for the last transfers: they are inside the timed region
            t1.record()
            barrier()
        ms = allmax([t0.elapsed_time(t1)])[0]
        desc = g.describe()
        g.close()
        del img
        return ms, desc

    if args.sweep:
        rows = []
        for B in (1, 2, 4, 8, 16, 32, 64, 128, 256):
            steps = 3 if B >= 64 else 5
            ms_c, _ = timed_encode(B, steps, 2, "none")
            row = {"batch_per_gpu": B, "global_batch": B * world, "images_per_s_compute_only": world * B * steps / (ms_c / 1e3)}
            if world > 1:
                ms_g, desc = timed_encode(B, steps, 2, args.gather)
                row.update({"images_per_s_with_gather": world * B * steps / (ms_g / 1e3), "gather": desc})
            rows.append(row)
            torch.cuda.empty_cache()
        if rank == 0:
            print(json.dumps({"config": "BASELINE.json configs[4]: Hiera-L image-encoder throughput sweep, batch-sharded", "n_gpus": world, "rows": rows}))
        if world > 1:
            dist.destroy_process_group()
        return

    B = args.batch
    # ---- headline: device-resident inputs ---------------------------------------------------------------------------
    clk = ClockSampler(local)
    l0 = E.launch_count()
    ms_total, gather_desc = timed_encode(B, args.steps, warmup, args.gather, sampler=clk)
    launches = (E.launch_count() - l0) * args.steps // (args.steps + warmup)
    value = world * B * args.steps / (ms_total / 1e3)

    # ---- e2e: host buffers through the plugin entry point -------------------------------------------------
    img_host = synth_image_norm(B, seed=1 + rank).pin_memory()
    outs_host = None
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
    e_ms = allmax([max(c0.elapsed_time(c1), 0.0), (time.perf_counter() - e0) * 1e3])
    e2e_sync_value = world * B * args.steps / (max(e_ms) / 1e3)
    h2d = img_host.numel() * 4
    d2h = sum(o.numel() * 4 for o in outs_host)
    # the same as a serving loop: two sets of pinned buffers, the call returns once enqueued (sam2b_image_encode_host_async) so
    # that step k + 1's uploads overlap step k's compute; every step's H2D and D2H are inside the timed region (host wall clock
    # from the first call until sam2b_host_sync has returned)
    img_host2 = synth_image_norm(B, seed=101 + rank).pin_memory()
    sets = [(img_host, outs_host), (img_host2, tuple(torch.empty_like(o).pin_memory() for o in outs_host))]
    for k in range(2):
        eng.image_encode_host(sets[k][0], True, False, sets[k][1], wait=False)
    eng.host_sync()
    barrier()
    p0 = time.perf_counter()
    for k in range(args.steps):
        eng.image_encode_host(sets[k & 1][0], True, False, sets[k & 1][1], wait=False)
    eng.host_sync()
    p_ms = allmax([(time.perf_counter() - p0) * 1e3])[0]
    barrier()
    e2e_value = world * B * args.steps / (p_ms / 1e3)
    del img_host2, sets

    # ---- the same from raw uint8 frames (what set_image_batch receives): bytes over PCIe, ToTensor/Normalize on the device
    raw_host = torch.randint(0, 256, (B, 1024, 1024, 3), dtype=torch.uint8, generator=torch.Generator().manual_seed(7 + rank)).pin_memory()
    for _ in range(2):
        outs_host = eng.image_encode_host_u8(raw_host, False, outs_host)
    barrier()
    u0 = time.perf_counter()
    for _ in range(args.steps):
        outs_host = eng.image_encode_host_u8(raw_host, False, outs_host)
    barrier()
    e2e_u8_value = world * B * args.steps / (allmax([(time.perf_counter() - u0) * 1e3])[0] / 1e3)
    raw_bytes = raw_host.numel()
    del raw_host, outs_host

    # ---- roofline of the dominant kernel (tcgen05 GEMM): live CUDA-event pairs around every launch of one more step
    roof = None
    if rank == 0:
        img = img_host.to(dev)
        PROF_STEPS = 3
        for _ in range(2):                        # back to the steady (power-capped) clocks of the timed region first
            eng.image_encode(img, True, False)
        E.profile_begin()
        for _ in range(PROF_STEPS):
            eng.image_encode(img, True, False)    # rank-local: no collective here (only rank 0 runs this leg)
        rows = E.profile_end()
        del img
        g = [(ms_, fl) for name, ms_, fl, by in rows if name.startswith("gemm")]
        all_ms = sum(r[1] for r in rows)
        g_ms, g_fl = sum(x[0] for x in g), sum(x[1] for x in g)
        peak, hbm, how = _peaks()
        ach = g_fl / (g_ms * 1e-3) / 1e12
        traffic, traffic_note = None, None
        tp = next((q for q in (os.path.join(ROOT, "profiles", n) for n in ("r2n_traffic.json", "r2_traffic.json", "r2b_traffic.json")) if os.path.exists(q)), None)
        if tp:
            tj = json.load(open(tp))
            traffic = tj["dram_bytes_per_launch_avg"]
            traffic_note = tj["note"]
        by_kernel = {}
        for name, ms_, fl, by in rows:
            k = name.split(" ")[0]
            by_kernel[k] = by_kernel.get(k, 0.0) + ms_
        forms = {}
        for name, ms_, fl, by in rows:
            if name.startswith(("gemm", "mlp_fused")):
                f = forms.setdefault(name, [0, 0.0, 0.0])
                f[0] += 1; f[1] += ms_; f[2] += fl
        gemm_forms = [{"form": k, "launches": v[0] // PROF_STEPS, "ms": round(v[1] / PROF_STEPS, 3), "tflops": round(v[2] / (v[1] * 1e-3) / 1e12, 1)}
                      for k, v in sorted(forms.items(), key=lambda kv: -kv[1][1])[:14]]
        roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": traffic,
                "traffic_note": traffic_note,
                "kernel": "gemm_bf16_tcgen05_kernel", "launches_per_step": len(g) // PROF_STEPS, "avg_launch_ms": g_ms / max(1, len(g)),
                "share_of_step_kernel_time": g_ms / all_ms, "peak_source": how,
                "flops_per_launch_avg": g_fl / max(1, len(g)),
                "kernel_time_ms_by_family": {k: round(v / PROF_STEPS, 3) for k, v in sorted(by_kernel.items(), key=lambda kv: -kv[1])},
                "profiled_steps": PROF_STEPS,
                "gemm_forms_by_time": gemm_forms,
                "note": "algorithmic 2MNK flops of every GEMM launch of three consecutive steps (after two un-profiled ones) / sum of their CUDA-event durations on the launching stream; per-step figures are the mean"}
    del img_host
    eng.release()
    torch.cuda.empty_cache()

    # ---- video legs (BASELINE.json: "video frames/sec propagated") through the public predictor API ----------------------
    video, video16 = None, None
    if args.video_frames > 0:
        from sam2_opt_b200.predictor import SAM2VideoPredictor
        vp = SAM2VideoPredictor(weights, device=dev)
        T = args.video_frames
        # config 3: one clip, one object per rank (N ranks = N independent videos)
        clip = synth_video(T, seed=1 + rank).pin_memory()
        run_clip(vp, clip, grid_clicks(1))                     # first pass warms allocator / workspaces / the heads graph
        barrier()
        stamps = run_clip(vp, clip, grid_clicks(1))
        s0 = min(16, len(stamps) - 2)      # steady state: full 7-frame bank + 16 object pointers from frame 16 on
        t_all, t_steady = allmax([stamps[-1] - stamps[0], stamps[-1] - stamps[s0]])
        n_steady = len(stamps) - 1 - s0
        video = {"metric": "video frames/sec propagated (SAM2VideoPredictor.propagate_in_video)", "unit": "frames/s",
                 "config": "BASELINE.json configs[2]: 1 object, 64 synthetic 1024x1024 frames, 7-frame memory bank" if T == 64 else f"{T}-frame clip, 1 object",
                 "value_steady": world * n_steady / t_steady, "value_overall": world * T / t_all,
                 "frames_per_video": T, "objects_per_video": 1, "videos": world,
                 "sam_heads": "libsam2b (sam2b_sam_heads)" if vp.native_heads else "torch fp32",
                 "tracked_frame": "one CUDA graph per steady-state frame (E3 + SAM heads + E4 + bank update)" if vp.graph_track else "enqueued per call",
                 "rank0_fps_by_16_frame_block": [round(min(16, len(stamps) - 1 - i) / (stamps[min(i + 16, len(stamps) - 1)] - stamps[i]), 1) for i in range(0, len(stamps) - 1, 16)],
                 "timing": "host wall clock around the generator incl. H2D of frames and a D2H read of every frame's masks, max over ranks; steady = frames >= 16 (7-frame memory bank, 16 pointers)"}
        if rank == 0:
            # roofline of the video path's dominant kernel (hd-256 flash attention over the memory bank), live CUDA events
            vp.graph_track = False            # the per-launch event profile needs the calls enqueued, not replayed from the graph
            E.profile_begin()
            run_clip(vp, clip[:24], grid_clicks(1))
            rows = E.profile_end()
            vp.graph_track = True
            video["roofline"] = attention_roofline(rows)
        del clip
        # config 4: `videos` videos x `video_objects` objects, videos (then objects) sharded over the ranks
        if args.video_objects > 0:
            nobj = args.video_objects
            mine = shard.shard_videos_objects(args.videos, nobj, world, rank)
            clicks = grid_clicks(nobj)
            clips = {v: synth_video(T, seed=100 + v).pin_memory() for v, _, _ in mine}
            if mine:
                v0, lo0, hi0 = mine[0]
                run_clip(vp, clips[v0][:20], clicks, lo0, hi0)      # warm-up: workspaces and the heads graph of this object count
            barrier()
            w0 = time.perf_counter()
            per_video = []
            for v, lo, hi in mine:
                st_ = run_clip(vp, clips[v], clicks, lo, hi)
                per_video.append(st_[-1] - st_[0])
            torch.cuda.synchronize()
            t_mine = time.perf_counter() - w0
            t_job = allmax([t_mine])[0]
            video16 = {"metric": "video propagation, aggregate object-frames/s", "unit": "object-frames/s",
                       "config": f"BASELINE.json configs[3]: {args.videos} videos x {nobj} objects x {T} frames sharded over {world} GPU(s) (shard.shard_videos_objects: videos first, then objects)",
                       "value": args.videos * nobj * T / t_job, "video_frames_per_s": args.videos * T / t_job,
                       "videos": args.videos, "objects_per_video": nobj, "frames_per_video": T, "seconds": t_job,
                       "rank0_seconds_per_video": [round(x, 3) for x in per_video],
                       "timing": "host wall clock from a barrier until the slowest rank has consumed its last frame (H2D of frames and a D2H mask read per frame inside)"}
            if rank == 0 and mine:
                vp.graph_track = False
                E.profile_begin()
                v0, lo0, hi0 = mine[0]
                run_clip(vp, clips[v0][:22], clicks, lo0, hi0)
                video16["roofline"] = attention_roofline(E.profile_end(), nobj)
                vp.graph_track = True
            del clips
        vp.release()
        torch.cuda.empty_cache()

    # ---- baselines: the reference on the host cores and on this GPU (rank 0, N = 1 only) ---------------------------------
    cpu, tgpu = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, cores, times, kind = cpu_reference_images_per_s(CPU_SAMPLE_IMAGES, 2, 1)
        cpu = {"value": v, "unit": "images/s", "cores": cores, "kind": kind,
               "sample": f"2 timed forward_image calls on {CPU_SAMPLE_IMAGES} images of the workload ({'the unmodified reference SAM2Base.forward_image from oracle/_ref' if kind == 'reference' else 'oracle port'}, fp32 torch, {cores} host threads), 1 single-image warm-up"}
        if video is not None:
            fps, cores, kind = cpu_reference_video_fps(CPU_VIDEO_FRAMES)
            video["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": cores, "kind": kind,
                                     "sample": f"the reference SAM2VideoPredictor.propagate_in_video on the first {CPU_VIDEO_FRAMES} frames of the clip: {CPU_VIDEO_FRAMES - 1} tracked frames / wall time of the propagate loop (the prompted frame is served from its stored output; the bank holds 1-2 frames, not 7, so this flatters the CPU)"}
    if rank == 0 and world == 1 and not args.no_torch_gpu_baseline:
        try:
            tgpu = torch_gpu_baseline(dev, B)
        except Exception as ex:      # a baseline must never take the headline line down with it
            tgpu = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}
        if video is not None:
            try:
                video["torch_gpu_baseline"] = torch_gpu_video_baseline(dev, args.video_frames)
            except Exception as ex:
                video["torch_gpu_baseline"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"SAM2.1 Hiera-L image-encoder forward (boundary E2: trunk+FpnNeck+conv_s0/s1), batch {B} x 3x1024x1024 fp32 per GPU, random-init weights",
                       "batch_per_gpu": B, "global_batch": B * world, "sharding": "batch" if world > 1 else "none",
                       "collective": gather_desc,
                       "l2": "inputs (403 MB/step) and activations exceed the 126 MB L2; no explicit flush"},
            "tflops_algorithmic": value * GFLOP_PER_IMAGE / 1e3,
            "clocks": clk.summary(), "gpu_launches": int(launches),
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "Engine.image_encode_host(wait=False) -> sam2b_image_encode_host_async per step + sam2b_host_sync at the end (serving loop: "
                           "two sets of pinned host buffers, step k+1's uploads overlap step k's compute; every step's H2D and D2H inside the "
                           "timed region, host wall clock)",
                    "per_call_sync_value": e2e_sync_value,
                    "per_call_sync_api": "Engine.image_encode_host -> sam2b_image_encode_host (returns after the step's outputs are in host memory; no overlap between steps)"},
            "e2e_u8_frames": {"value": e2e_u8_value, "unit": "images/s", "h2d_bytes_per_step": raw_bytes, "d2h_bytes_per_step": d2h,
                              "api": "Engine.image_encode_host_u8 -> sam2b_image_encode_host_u8 (raw uint8 HWC frames, pinned; host wall clock)"},
            "roofline": roof, "cpu_baseline": cpu, "torch_gpu_baseline": tgpu, "video": video, "video_config4": video16,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
