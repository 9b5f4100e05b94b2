"""Parity of the tcgen05 GEMM (sam2b_op_gemm, through the C ABI) against a torch fp32 matmul of the same
bf16-rounded operands.  Tolerance: fp32 accumulation-order noise only (1e-3 relative to the output scale);
bf16 outputs add one rounding (2^-8 relative)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, bias=None, act=None, alpha=1.0, gamma=None, residual=None, bias_along_m=False):
    o = alpha * (a.float() @ b.float().transpose(-1, -2))
    if bias is not None:
        o = o + (bias[:, None] if bias_along_m else bias)
    if act == "gelu":
        o = torch.nn.functional.gelu(o)
    elif act == "relu":
        o = torch.relu(o)
    if gamma is not None:
        o = o * gamma
    if residual is not None:
        o = o + residual
    return o


def _close(x, y, tol):
    scale = y.abs().max().item() + 1e-6
    err = (x.float() - y).abs().max().item()
    assert err <= tol * scale, f"max err {err:.4g} vs scale {scale:.4g}"


# (M, N, K): trunk linears of every stage incl. ragged K (144 = 2.25 k-blocks) and ragged N / M tails
SHAPES = [(256, 128, 64), (128, 256, 128), (4096, 1728, 576), (4096, 576, 2304), (8192, 432, 144), (1000, 144, 144),
          (1024, 3456, 1152), (4096, 2304, 576), (300, 72, 160), (65536, 144, 160), (16384, 864, 288)]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_plain(M, N, K):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    out = ops.gemm(a, b)
    torch.cuda.synchronize()
    _close(out, _ref(a, b), 2e-3)


@pytest.mark.parametrize("bn", [32, 64, 96, 128, 160, 192, 224, 256])
def test_gemm_every_block_n(bn):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(bn)
    M, N, K = 640, 1008, 328
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    out = ops.gemm(a, b, bias=bias, block_n=bn)
    torch.cuda.synchronize()
    _close(out, _ref(a, b, bias), 2e-3)


def test_gemm_epilogues():
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(7)
    M, N, K = 2048, 576, 576
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    gamma = torch.rand(N, device="cuda", generator=g) + 0.5
    res = torch.randn(M, N, device="cuda", generator=g)
    _close(ops.gemm(a, b, bias=bias, act="gelu", out_dtype=torch.bfloat16), _ref(a, b, bias, "gelu"), 1e-2)
    _close(ops.gemm(a, b, bias=bias, act="relu"), _ref(a, b, bias, "relu"), 2e-3)
    _close(ops.gemm(a, b, bias=bias, residual=res), _ref(a, b, bias, residual=res), 2e-3)
    _close(ops.gemm(a, b, bias=bias, gamma=gamma, residual=res, alpha=0.5), _ref(a, b, bias, None, 0.5, gamma, res), 2e-3)
    # residual broadcast over row blocks (pos-embed add of the patch embedding)
    res_small = torch.randn(256, N, device="cuda", generator=g)
    _close(ops.gemm(a, b, residual=res_small, res_row_mod=256), _ref(a, b, residual=res_small.repeat(M // 256, 1)), 2e-3)
    # in-place residual (what the trunk does): out aliases residual
    bias_m = torch.randn(M, device="cuda", generator=g)
    _close(ops.gemm(a, b, bias=bias_m, bias_along_m=True), _ref(a, b, bias_m, bias_along_m=True), 2e-3)
    torch.cuda.synchronize()


def test_gemm_batched_shared_a_nchw_out():
    """FpnNeck form: weights [Cout,Cin] as A (shared), per-image tokens [hw,Cin] as B -> NCHW output."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(11)
    n, hw, cin, cout = 3, 4096, 576, 256
    wgt = (torch.randn(cout, cin, device="cuda", generator=g) / cin ** 0.5).bfloat16()
    x = torch.randn(n, hw, cin, device="cuda", generator=g).bfloat16()
    bias = torch.randn(cout, device="cuda", generator=g)
    out = ops.gemm(wgt[None].expand(n, -1, -1), x, bias=bias, bias_along_m=True)
    ref = torch.einsum("ok,nhk->noh", wgt.float(), x.float()) + bias[None, :, None]
    _close(out, ref, 2e-3)
    # small-M (32 rows, one quarter of the MMA tile)
    w32 = (torch.randn(32, cin, device="cuda", generator=g) / cin ** 0.5).bfloat16()
    out = ops.gemm(w32[None].expand(n, -1, -1), x)
    _close(out, torch.einsum("ok,nhk->noh", w32.float(), x.float()), 2e-3)


def test_gemm_many_tiles_persistent_phases():
    """> 148*2 tiles per CTA so every barrier ring wraps several times (both accumulator stages, all phases)."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 128 * 148 * 5, 256, 1152
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    out = ops.gemm(a, b, block_n=64)
    _close(out, _ref(a, b), 2e-3)


# ---- CTA pairs (cluster of 2, tcgen05 cta_group::2, 256 x BLOCK_N tiles): block_n = 1000 + BLOCK_N forces them ----
@pytest.mark.parametrize("bn", [128, 160, 192, 224, 256])
def test_gemm_cta_pairs_every_block_n(bn):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(100 + bn)
    M, N, K = 1000, 1008, 328            # ragged M (not a multiple of 256 or 128), ragged N and K
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    out = ops.gemm(a, b, bias=bias, block_n=1000 + bn)
    torch.cuda.synchronize()
    _close(out, _ref(a, b, bias), 2e-3)


@pytest.mark.parametrize("M,N,K,bn", [(65536, 1728, 576, 256), (65536, 576, 2304, 192), (16384, 2304, 576, 256), (128 * 148 * 3, 256, 1152, 128)])
def test_gemm_cta_pairs_epilogues_persistent(M, N, K, bn):
    """Trunk forms on pair tiles, with enough tiles per pair that every barrier ring wraps several times."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(M % 1000 + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    _close(ops.gemm(a, b, bias=bias, out_dtype=torch.bfloat16, block_n=1000 + bn), _ref(a, b, bias), 1e-2)
    _close(ops.gemm(a, b, bias=bias, act="gelu", out_dtype=torch.bfloat16, block_n=1000 + bn), _ref(a, b, bias, "gelu"), 1e-2)
    _close(ops.gemm(a, b, bias=bias, residual=res, block_n=1000 + bn), _ref(a, b, bias, residual=res), 2e-3)
    res_small = torch.randn(512, N, device="cuda", generator=g)
    _close(ops.gemm(a, b, residual=res_small, res_row_mod=512, block_n=1000 + bn), _ref(a, b, residual=res_small.repeat(M // 512, 1)), 2e-3)
    torch.cuda.synchronize()


def test_gemm_cta_pairs_batched():
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(13)
    n, M, N, K = 3, 4096, 4100, 256
    a = torch.randn(n, M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(n, N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    out = ops.gemm(a, b, alpha=0.0625, block_n=1256)
    _close(out, 0.0625 * torch.einsum("bmk,bnk->bmn", a.float(), b.float()), 2e-3)


# ---- LayerNorm fold: producer GEMM emits bf16 rows + row statistics, consumer GEMM finishes LN in its epilogue ----
@pytest.mark.parametrize("C,N2,M,bn", [(576, 1728, 4096, 0), (144, 432, 8192, 0), (288, 1152, 3000, 0), (1152, 3456, 1024, 0),
                                       (576, 2304, 4096, 1192)])
def test_gemm_layernorm_fold(C, N2, M, bn):
    """x = a Wp^T + bp + res;  y = act(LayerNorm(x; gamma, beta, eps 1e-6) W^T + b)  vs torch fp32 (reference
    MultiScaleBlock: norm -> Linear, hieradet.py:134-166).  Tolerance: bf16 operand rounding (1e-2 of the output scale)."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(C + N2)
    Kp = 256
    a = torch.randn(M, Kp, device="cuda", generator=g).bfloat16()
    wp = (torch.randn(C, Kp, device="cuda", generator=g) / Kp ** 0.5).bfloat16()
    bp = torch.randn(C, device="cuda", generator=g)
    res = torch.randn(M, C, device="cuda", generator=g) * 2 + 0.7          # non-zero row mean
    x, xb, stats = ops.gemm_ln_producer(a, wp, bp, res, block_n=bn)
    x_ref = a.float() @ wp.float().t() + bp + res
    _close(x, x_ref, 2e-3)
    _close(xb, x_ref, 1e-2)
    s = stats.sum(1)
    _close(s[:, 0], x_ref.sum(1), 1e-4)
    _close(s[:, 1], (x_ref * x_ref).sum(1), 1e-4)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.3
    W = torch.randn(N2, C, device="cuda", generator=g) / C ** 0.5
    b = torch.randn(N2, device="cuda", generator=g)
    w_scaled = W * gamma[None, :]
    w_scaled = (w_scaled - w_scaled.mean(1, keepdim=True)).bfloat16()      # zero-mean rows absorb LN's mean subtraction
    bias2 = b + W @ beta
    for act in (None, "gelu"):
        y = ops.gemm_ln_consumer(xb, w_scaled, bias2, stats.contiguous(), 1e-6, act=act)
        y_ref = torch.nn.functional.layer_norm(x_ref, (C,), gamma, beta, 1e-6) @ W.t() + b
        if act == "gelu":
            y_ref = torch.nn.functional.gelu(y_ref)
        _close(y, y_ref, 1.5e-2)
    y32 = ops.gemm_ln_consumer(xb, w_scaled, bias2, stats.contiguous(), 1e-6, out_dtype=torch.float32)
    _close(y32, torch.nn.functional.layer_norm(x_ref, (C,), gamma, beta, 1e-6) @ W.t() + b, 1.5e-2)


# ---- 16 epilogue warps (block_n + 2000), alone and with CTA pairs (+ 1000) ----
@pytest.mark.parametrize("code", [2128, 2160, 2192, 2224, 2256, 3192, 3256])
def test_gemm_sixteen_epilogue_warps(code):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(code)
    M, N, K = 128 * 148 * 2 + 72, 1008, 328
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g)
    _close(ops.gemm(a, b, bias=bias, block_n=code), _ref(a, b, bias), 2e-3)
    _close(ops.gemm(a, b, bias=bias, out_dtype=torch.bfloat16, block_n=code), _ref(a, b, bias), 1e-2)
    _close(ops.gemm(a, b, bias=bias, act="gelu", out_dtype=torch.bfloat16, block_n=code), _ref(a, b, bias, "gelu"), 1e-2)
    _close(ops.gemm(a, b, bias=bias, residual=res, block_n=code), _ref(a, b, bias, residual=res), 2e-3)
    torch.cuda.synchronize()


@pytest.mark.parametrize("code", [2160, 2192, 3192])
def test_gemm_layernorm_fold_sixteen_warps(code):
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(code)
    M, C, Kp, N2 = 5000, 576, 320, 1152
    a = torch.randn(M, Kp, device="cuda", generator=g).bfloat16()
    wp = (torch.randn(C, Kp, device="cuda", generator=g) / Kp ** 0.5).bfloat16()
    bp = torch.randn(C, device="cuda", generator=g)
    res = torch.randn(M, C, device="cuda", generator=g) + 0.5
    x, xb, stats = ops.gemm_ln_producer(a, wp, bp, res, block_n=code)
    x_ref = a.float() @ wp.float().t() + bp + res
    _close(x, x_ref, 2e-3)
    _close(xb, x_ref, 1e-2)
    s = stats.sum(1)
    _close(s[:, 0], x_ref.sum(1), 1e-4)
    _close(s[:, 1], (x_ref * x_ref).sum(1), 1e-4)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.3
    W = torch.randn(N2, C, device="cuda", generator=g) / C ** 0.5
    b = torch.randn(N2, device="cuda", generator=g)
    ws = W * gamma[None, :]
    ws = (ws - ws.mean(1, keepdim=True)).bfloat16()
    y = ops.gemm_ln_consumer(xb, ws, b + W @ beta, stats.contiguous(), 1e-6, act="gelu", block_n=2256)
    _close(y, torch.nn.functional.gelu(torch.nn.functional.layer_norm(x_ref, (C,), gamma, beta, 1e-6) @ W.t() + b), 1.5e-2)


@pytest.mark.parametrize("kind,tol", [("outliers", 1.5e-2), ("mean5", 3e-2), ("mean50", 0.25)])
def test_gemm_layernorm_fold_stress(kind, tol):
    """Where the fold's rounding differs from LayerNorm-then-bf16: the rows reach the consumer GEMM un-centred, so a bf16
    rounding step is relative to |x|, not to |x - mean|.  Outlier channels (a few channels at 100x the others, the 'massive
    activations' of trained ViTs) cost nothing -- the error stays relative per element, as for LayerNorm output rounded to
    bf16 -- while an offset common to all channels of a row grows the error as ~1.7e-3 (1 + |mean| / sigma) of the output
    scale (same figure from a CPU model of the arithmetic; DESIGN.md section 2 'unpinned').  The statistics themselves
    (sum, sum of squares in fp32) stay exact to 1e-4 in all three cases."""
    from sam2_opt_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(len(kind))
    M, C, Kp, N2 = 4096, 576, 256, 1152
    a = torch.randn(M, Kp, device="cuda", generator=g).bfloat16()
    wp = (torch.randn(C, Kp, device="cuda", generator=g) / Kp ** 0.5).bfloat16()
    bp = torch.randn(C, device="cuda", generator=g)
    res = torch.randn(M, C, device="cuda", generator=g) * 2
    if kind == "outliers":
        res[:, [5, 100, 400]] += torch.tensor([300.0, -250.0, 180.0], device="cuda")
    else:
        res += 2.2 * float(kind[4:])
    x, xb, stats = ops.gemm_ln_producer(a, wp, bp, res)
    x_ref = a.float() @ wp.float().t() + bp + res
    s = stats.sum(1)
    _close(s[:, 0], x_ref.sum(1), 1e-4)
    _close(s[:, 1], (x_ref * x_ref).sum(1), 1e-4)
    gamma = torch.rand(C, device="cuda", generator=g) + 0.5
    beta = torch.randn(C, device="cuda", generator=g) * 0.3
    W = torch.randn(N2, C, device="cuda", generator=g) / C ** 0.5
    b = torch.randn(N2, device="cuda", generator=g)
    ws = W * gamma[None, :]
    ws = (ws - ws.mean(1, keepdim=True)).bfloat16()
    y = ops.gemm_ln_consumer(xb, ws, b + W @ beta, stats.contiguous(), 1e-6, out_dtype=torch.float32)
    y_ref = torch.nn.functional.layer_norm(x_ref, (C,), gamma, beta, 1e-6) @ W.t() + b
    err = ((y - y_ref).abs().max() / y_ref.abs().max()).item()
    ratio = (x_ref.mean(1).abs() / x_ref.std(1)).mean().item()
    print(f"LayerNorm fold, {kind}: |mean|/sigma {ratio:.1f}, max err / scale {err:.3e}")
    _close(y, y_ref, tol)
