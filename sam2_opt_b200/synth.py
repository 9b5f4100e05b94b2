"""Seeded synthetic inputs for the workloads BASELINE.json names (no datasets offline).

Plain numpy ``default_rng`` streams so the same arrays are produced in the build container
(where the golden fixtures are made with the reference) and on the GPU box."""
import numpy as np
import torch


def checksum(t) -> float:
    """Order-independent fingerprint stored next to golden outputs so a changed RNG stream is
    reported as such, not as a parity failure."""
    t = torch.as_tensor(t).double().flatten()
    return float((t * torch.cos(torch.arange(t.numel(), dtype=torch.float64) * 0.001)).sum())


def synth_image_u8(seed: int = 0, size: int = 1024) -> np.ndarray:
    """Config 1: uint8 HWC image, rng(seed).integers(0,256)."""
    return np.random.default_rng(seed).integers(0, 256, (size, size, 3), dtype=np.uint8)


def synth_image_norm(n: int, seed: int = 1, size: int = 1024) -> torch.Tensor:
    """Config 2: float32[n,3,size,size] ~ N(0,1) — an already-normalised batch (what the
    reference feeds its export, export_image_onnx.py:49)."""
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.standard_normal((n, 3, size, size), dtype=np.float32))


def synth_video(t: int, seed: int = 1, size: int = 1024) -> torch.Tensor:
    """Config 3: temporally coherent clip — base ~ N(0,1) + 0.1 * per-frame noise."""
    rng = np.random.default_rng(seed)
    base = rng.standard_normal((3, size, size), dtype=np.float32)
    return torch.from_numpy(np.stack([base + 0.1 * rng.standard_normal((3, size, size), dtype=np.float32)
                                      for _ in range(t)]))


def synth_memattn_inputs(L: int, P: int, N: int, seed: int = 3):
    """Boundary E3 inputs: curr, memory, curr_pos, memory_pos, memory_exclude, memory_pos_exclude."""
    rng = np.random.default_rng(seed)
    f = lambda *s: torch.from_numpy(rng.standard_normal(s, dtype=np.float32))
    curr, curr_pos = f(4096, N, 256), f(4096, N, 256)
    mem = f(L, 4096, N, 64).to(torch.bfloat16).float()      # the bank is stored in bf16
    return curr, mem, curr_pos, f(L, 4096, N, 64), f(P, N, 64), f(P, N, 64)


def synth_memenc_inputs(N: int, seed: int = 4):
    """Boundary E4 inputs: pix_feat [N,256,64,64], sigmoid-scaled mask [N,1,1024,1024] in [-10,10]."""
    rng = np.random.default_rng(seed)
    pix = torch.from_numpy(rng.standard_normal((N, 256, 64, 64), dtype=np.float32))
    # smooth blobby logits so the sigmoid mask has real edges
    lo = torch.from_numpy(rng.standard_normal((N, 1, 32, 32), dtype=np.float32))
    logits = torch.nn.functional.interpolate(lo, size=(1024, 1024), mode="bilinear", align_corners=False) * 4
    return pix, torch.sigmoid(logits) * 20.0 - 10.0
