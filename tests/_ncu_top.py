"""Dev tool (not a test): one launch of each dominant kernel on its real stage-3 shape (batch 16), for an `ncu --set full`
capture.  Order: qkv (LayerNorm-fold consumer), fc1 (consumer + GELU), fc2 and attn.proj (LayerNorm producers: fp32 +
residual + bf16 copy + row statistics); hd72 attention 16x16 windows and global; 8x8 / 4x4 small-window streaming kernel;
hd256 attention (28 736 keys)."""
import sys

import torch

sys.path.insert(0, "/root/repo")
from sam2_opt_b200 import ops

M, C = 65536, 576
g = torch.Generator(device="cuda").manual_seed(0)
rnd = lambda *s: torch.randn(*s, device="cuda", generator=g)
# producer: attn.proj form (K = 576) and fc2 form (K = 2304)
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
# consumers: qkv and fc1 (+GELU) on the raw rows
stats = stats.contiguous()
for N, act in ((3 * C, None), (4 * C, "gelu")):
    w = (rnd(N, C) / C ** 0.5)
    w = (w - w.mean(1, keepdim=True)).bfloat16()
    b = rnd(N)
    for _ in range(2):
        ops.gemm_ln_consumer(xb, w, b, stats, 1e-6, act=act)
    torch.cuda.synchronize()
qkv = rnd(16, 64, 64, 3 * 576).bfloat16()
for ws in (16, 64):
    for _ in range(2):
        ops.window_attention(qkv, 8, ws, False)
    torch.cuda.synchronize()
qkv1 = rnd(16, 256, 256, 3 * 144).bfloat16()
for _ in range(2):
    ops.window_attention(qkv1, 2, 8, False)
torch.cuda.synchronize()
q = rnd(1, 4096, 256).bfloat16()
k = rnd(1, 28736, 256).bfloat16()
v = rnd(1, 28736, 256).bfloat16()
for _ in range(2):
    ops.attention_hd256(q, k, v)
torch.cuda.synchronize()
print("ok")
