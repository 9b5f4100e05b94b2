#!/usr/bin/env python
"""Dev tool (not a test): two launches of each dominant kernel on its real shape at BATCH 32 (the bench configuration), for
an `ncu --set full` capture of the second (warm) one.  Order: attn.proj and fc2 (LayerNorm producers: fp32 + residual + bf16
copy + row statistics), qkv (LayerNorm-fold consumer), fc1 (consumer + GELU); hd72 attention 16x16 windows and global; 8x8
small-window kernel; hd256 attention (28 736 keys); hd256 / 64-wide-value cross-attention (16 objects).

    ncu --set full --clock-control none --import-source on -k regex:'gemm_bf16|flash_attn|window_attn' -o gpurun_out/r2_top python tools/ncu_top_kernels.py
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sam2_opt_b200 import ops  # noqa: E402

B = int(os.environ.get("NCU_BATCH", "32"))
M, C = 4096 * B, 576
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
att = rnd(M, C).bfloat16()
wp = (rnd(C, C) / C ** 0.5).bfloat16()
hid = rnd(M, 4 * C).bfloat16()
w2 = (rnd(C, 4 * C) / (4 * C) ** 0.5).bfloat16()
bias = rnd(C)
res = rnd(M, C)
for _ in range(2):
    x, xb, stats = ops.gemm_ln_producer(att, wp, bias, res)
torch.cuda.synchronize()
for _ in range(2):
    ops.gemm_ln_producer(hid, w2, bias, res)
torch.cuda.synchronize()
del hid
stats = stats.contiguous()
for N, act in ((3 * C, None), (4 * C, "gelu")):
    w = (rnd(N, C) / C ** 0.5)
    w = (w - w.mean(1, keepdim=True)).bfloat16()
    b = rnd(N)
    for _ in range(2):
        ops.gemm_ln_consumer(xb, w, b, stats, 1e-6, act=act)
    torch.cuda.synchronize()
qkv = rnd(B, 64, 64, 3 * 576).bfloat16()
for ws in (16, 64):
    for _ in range(2):
        ops.window_attention(qkv, 8, ws, False)
    torch.cuda.synchronize()
del qkv
qkv1 = rnd(B, 256, 256, 3 * 144).bfloat16()
for _ in range(2):
    ops.window_attention(qkv1, 2, 8, False)
torch.cuda.synchronize()
del qkv1
q = rnd(1, 4096, 256).bfloat16()
k = rnd(1, 28736, 256).bfloat16()
v = rnd(1, 28736, 256).bfloat16()
for _ in range(2):
    ops.attention_hd256(q, k, v)
torch.cuda.synchronize()
# cross-attention of the memory attention, 16 objects: 64-wide values (v_proj folded behind P V)
q = rnd(16, 4096, 256).bfloat16()
k = rnd(16, 28800, 256).bfloat16()
v = rnd(16, 28800, 64).bfloat16()
for _ in range(2):
    ops.attention_hd256_v64(q, k, v)
torch.cuda.synchronize()
print("ok")
