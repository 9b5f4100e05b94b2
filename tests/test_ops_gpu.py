"""Parity of the non-GEMM kernels (through the C ABI) against plain torch fp32 on the same inputs."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("C", [144, 256, 288, 576, 1152])
def test_layernorm(C):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(C)
    x = torch.randn(1000, C, device="cuda", generator=g) * 3 + 1
    w = torch.randn(C, device="cuda", generator=g)
    b = torch.randn(C, device="cuda", generator=g)
    ref = F.layer_norm(x, (C,), w, b, 1e-6)
    y32 = ops.layernorm(x, w, b, 1e-6, out_dtype=torch.float32)
    assert (y32 - ref).abs().max().item() < 2e-5 * ref.abs().max().item()
    y16 = ops.layernorm(x, w, b, 1e-6)
    assert (y16.float() - ref).abs().max().item() <= 2 ** -8 * ref.abs().max().item()


def test_cast_and_maxpool():
    from sam2_opt_b200 import ops
    x = torch.randn(2, 64, 64, 288, device="cuda")
    assert torch.equal(ops.cast_bf16(x), x.bfloat16())
    ref = F.max_pool2d(x.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
    assert torch.equal(ops.maxpool2x2_nhwc(x), ref)


def _attn_ref(qkv, heads, ws, q_pool):
    """Reference semantics (hieradet.py:56-81 + window partition) in fp32 on bf16-rounded q/k/v."""
    B, H, W, _ = qkv.shape
    C = heads * 72
    x = qkv.float().view(B, H // ws, ws, W // ws, ws, 3, heads, 72).permute(0, 1, 3, 2, 4, 5, 6, 7)
    x = x.reshape(-1, ws, ws, 3, heads, 72)
    q, k, v = x[:, :, :, 0], x[:, :, :, 1], x[:, :, :, 2]
    Bw = q.shape[0]
    wq = ws
    if q_pool:
        q = F.max_pool2d(q.reshape(Bw, ws, ws, C).permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
        wq = ws // 2
        q = q.reshape(Bw, wq, wq, heads, 72)
    q = q.reshape(Bw, wq * wq, heads, 72).transpose(1, 2)
    k = k.reshape(Bw, ws * ws, heads, 72).transpose(1, 2)
    v = v.reshape(Bw, ws * ws, heads, 72).transpose(1, 2)
    o = F.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(Bw, wq, wq, C)
    Ho, Wo = (H // 2, W // 2) if q_pool else (H, W)
    o = o.reshape(B, H // ws, W // ws, wq, wq, C).permute(0, 1, 3, 2, 4, 5).reshape(B, Ho, Wo, C)
    return o


# (H, heads, ws, q_pool): every window configuration of Hiera-L (scaled-down grids keep the test fast)
ATTN = [(64, 2, 8, False), (64, 4, 8, True), (32, 4, 4, False), (32, 8, 4, True), (32, 8, 16, False),
        (64, 8, 64, False), (32, 16, 16, True), (32, 16, 8, False)]


@pytest.mark.parametrize("H,heads,ws,q_pool", ATTN)
def test_window_attention(H, heads, ws, q_pool):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(H * 31 + heads * 7 + ws)
    B = 2
    qkv = (torch.randn(B, H, H, 3 * heads * 72, device="cuda", generator=g) * 1.5).bfloat16()
    out = ops.window_attention(qkv, heads, ws, q_pool)
    torch.cuda.synchronize()
    ref = _attn_ref(qkv, heads, ws, q_pool)
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * ref.abs().max().item(), err     # P is rounded to bf16 before PV (2^-8) + bf16 output


def test_softmax_rows():
    from sam2_opt_b200 import ops
    s = torch.randn(64, 28736 + 4, device="cuda") * 4
    p = ops.softmax_rows(s)
    ref = torch.softmax(s, -1)
    assert p.shape[1] % 64 == 0 and (p[:, s.shape[1]:] == 0).all()
    assert (p[:, :s.shape[1]].float() - ref).abs().max().item() <= 2 ** -8 * ref.max().item()


# (n, Tq, Lk): self attention, the first propagated frame (L=1, P=4), the full bank (7*4096+64), ragged Lk, multi-object
ATTN256 = [(1, 4096, 4096), (1, 4096, 4100), (1, 4096, 28736), (3, 256, 200), (4, 4096, 8256), (2, 128, 64)]


@pytest.mark.parametrize("n,Tq,Lk", ATTN256)
def test_attention_hd256(n, Tq, Lk):
    """tcgen05 flash attention (K4) against torch SDPA in fp32 on the same bf16 operands."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(n * 7 + Lk)
    q = (torch.randn(n, Tq, 256, device="cuda", generator=g) * 1.2).bfloat16()
    k = (torch.randn(n, Lk, 256, device="cuda", generator=g) * 1.2).bfloat16()
    v = torch.randn(n, Lk, 256, device="cuda", generator=g).bfloat16()
    out = ops.attention_hd256(q, k, v)
    torch.cuda.synchronize()
    ref = F.scaled_dot_product_attention(q.float()[:, None], k.float()[:, None], v.float()[:, None])[:, 0]
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * ref.abs().max().item(), err     # P rounded to bf16 before PV (2^-8) + bf16 output


def test_position_encoding_export_matches_the_reference_formula():
    """sam2b_position_encoding (the constant vision_pos_enc / mask_memory_pos_enc outputs of E2 / E4) against the oracle's
    restatement of PositionEmbeddingSine (position_encoding.py:90-124), which oracle/make_golden.py pinned on the
    reference's own pos outputs.  fp32 sin/cos of arguments up to 2 pi: a few ulp of the argument, 5e-6 absolute."""
    from oracle import sam2_oracle as O
    from sam2_opt_b200.engine import Engine
    from sam2_opt_b200.synth_weights import make_weights
    eng = Engine(make_weights(0), torch.device("cuda"))
    for c, size in ((256, 256), (256, 128), (256, 64), (64, 64)):
        got = eng.position_encoding(c, size, 2)
        assert got.shape == (2, c, size, size) and got.stride(0) == 0          # a view of one constant, not a copy
        want = O.position_sine(c, size, size)
        assert (got[0].cpu() - want).abs().max().item() <= 5e-6, (c, size)
    eng.release()
