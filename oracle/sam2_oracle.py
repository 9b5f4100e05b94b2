"""CPU oracle for the SAM2.1 Hiera-L inference hot path  --  TEST INFRASTRUCTURE ONLY.

A plain fp32 torch restatement (functional, over a flat ``{name: tensor}`` weight dict that
uses the reference's own ``state_dict`` names) of what FasterProcess/sam2_opt computes on
the path this repo accelerates.  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu_baseline / ``--impl reference`` legs may import it; the product package
``sam2_opt_b200`` never does.

Parity pinning: the reference ships NO tests or golden vectors for this path.
The oracle is pinned instead against the reference implementation itself, run in the build
container: ``oracle/make_golden.py`` loads these synthetic weights into the real reference
modules (strict ``load_state_dict`` -> the names/shapes below are the reference's), runs the
reference torch path and commits output slices under ``tests/golden/``;
``tests/test_oracle_golden.py`` re-checks this file against them everywhere, and
``tests/test_oracle_vs_reference.py`` re-runs the live comparison where /root/reference exists.

All ``ref:`` citations are relative to /root/reference/sam2/sam2/.
"""
import math
import zlib
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
W = Dict[str, Tensor]

# ---------------------------------------------------------------------------------------
# Hiera-L constants (ref: configs/sam2.1/sam2.1_hiera_l.yaml:10-16, hieradet.py:208-266)
# ---------------------------------------------------------------------------------------
EMBED_DIM = 144
STAGES = (2, 6, 36, 4)
GLOBAL_ATT_BLOCKS = (23, 33, 43)
WINDOW_SPEC = (8, 4, 16, 8)
STAGE_ENDS = (1, 7, 43, 47)
Q_POOL_BLOCKS = (2, 8, 44)
HEAD_DIM = 72
D_MODEL = 256
MEM_DIM = 64
NUM_MASKMEM = 7
IMAGE_SIZE = 1024
NO_OBJ_SCORE = -1024.0
IMG_MEAN = (0.485, 0.456, 0.406)   # ref: utils/transforms.py:27-28
IMG_STD = (0.229, 0.224, 0.225)


def block_cfgs() -> List[dict]:
    """Per-block (dim, dim_out, heads, window, q_pool) — ref: hieradet.py:233-260.

    The window size lags one block behind the stage change (hieradet.py:237-248)."""
    cfgs, dim, heads, cur_stage = [], EMBED_DIM, 2, 1
    for i in range(sum(STAGES)):
        dim_out = dim
        window = WINDOW_SPEC[cur_stage - 1]
        if i in GLOBAL_ATT_BLOCKS:
            window = 0
        if i - 1 in STAGE_ENDS:
            dim_out, heads, cur_stage = dim * 2, heads * 2, cur_stage + 1
        cfgs.append(dict(dim=dim, dim_out=dim_out, heads=heads, window=window,
                         q_pool=i in Q_POOL_BLOCKS))
        dim = dim_out
    return cfgs


# ---------------------------------------------------------------------------------------
# Weight spec + deterministic synthetic ("random-init") weights: shared with the product's bench / smoke
# (sam2_opt_b200/synth_weights.py) so every arm sees bit-identical values.
# ---------------------------------------------------------------------------------------
from sam2_opt_b200.synth_weights import make_weights, weight_spec  # noqa: E402,F401


# ---------------------------------------------------------------------------------------
# Optional operand-rounding emulation (numerics studies only; default = exact fp32)
# ---------------------------------------------------------------------------------------
_EMULATE = {"dtype": None}


def set_emulation(dtype: Optional[torch.dtype]):
    """Round every GEMM/SDPA *input* to ``dtype`` (fp32 accumulate) — predicts what the
    bf16-operand kernels do, on CPU.  ``None`` restores the exact fp32 oracle."""
    _EMULATE["dtype"] = dtype


def _r(x: Tensor) -> Tensor:
    d = _EMULATE["dtype"]
    return x if d is None else x.to(d).float()


def linear(x, w, b=None):
    return F.linear(_r(x), _r(w), b)


def sdpa(q, k, v):
    q, k, v = _r(q), _r(k), _r(v)
    if _EMULATE["dtype"] is None:
        return F.scaled_dot_product_attention(q, k, v)
    s = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(q.shape[-1]), dim=-1)
    return _r(s) @ v


def conv1x1(x, w, b):
    return F.conv2d(_r(x), _r(w), b)


def layer_norm(x, w, b, eps):
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def layer_norm_2d(x, w, b, eps=1e-6):
    """ref: modeling/sam2_utils.py:141-153 (LayerNorm2d over the channel dim of NCHW)."""
    u = x.mean(1, keepdim=True)
    s = (x - u).pow(2).mean(1, keepdim=True)
    x = (x - u) / torch.sqrt(s + eps)
    return w[:, None, None] * x + b[:, None, None]


# ---------------------------------------------------------------------------------------
# Image encoder: Hiera trunk + FpnNeck  
# ---------------------------------------------------------------------------------------
def normalize_image(img01: Tensor) -> Tensor:
    """ref: utils/transforms.py:31 (torchvision Normalize) as used by
    sam2_image_predictor.py:253."""
    mean = torch.tensor(IMG_MEAN, dtype=img01.dtype, device=img01.device)[None, :, None, None]
    std = torch.tensor(IMG_STD, dtype=img01.dtype, device=img01.device)[None, :, None, None]
    return (img01 - mean) / std


def hiera_pos_embed(w: W, h: int, wd: int) -> Tensor:
    """ref: modeling/backbones/hieradet.py:273-281. Returns [1,h,w,C]."""
    t = "image_encoder.trunk."
    pe = F.interpolate(w[t + "pos_embed"], size=(h, wd), mode="bicubic")
    win = w[t + "pos_embed_window"]
    pe = pe + win.tile([x // y for x, y in zip(pe.shape, win.shape)])
    return pe.permute(0, 2, 3, 1)


def window_partition(x: Tensor, ws: int) -> Tensor:
    """ref: modeling/backbones/utils.py:16-36 (no padding at 1024^2)."""
    B, H, Wd, C = x.shape
    x = x.view(B, H // ws, ws, Wd // ws, ws, C)
    return x.permute(0, 1, 3, 2, 4, 5).reshape(-1, ws, ws, C)


def window_unpartition(win: Tensor, ws: int, H: int, Wd: int) -> Tensor:
    """ref: modeling/backbones/utils.py:39-60."""
    B = win.shape[0] // ((H // ws) * (Wd // ws))
    x = win.reshape(B, H // ws, Wd // ws, ws, ws, -1)
    return x.permute(0, 1, 3, 2, 4, 5).reshape(B, H, Wd, -1)


def _maxpool2(x_nhwc: Tensor) -> Tensor:
    """ref: hieradet.py:25-36 (do_pool with MaxPool2d(2,2))."""
    return F.max_pool2d(x_nhwc.permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)


def multiscale_attention(w: W, p: str, x: Tensor, heads: int, q_pool: bool) -> Tensor:
    """ref: hieradet.py:56-81. x: [B', H, W, C] (already window-partitioned)."""
    B, H, Wd, _ = x.shape
    qkv = linear(x, w[p + "qkv.weight"], w[p + "qkv.bias"]).reshape(B, H * Wd, 3, heads, -1)
    q, k, v = torch.unbind(qkv, 2)
    if q_pool:
        q = _maxpool2(q.reshape(B, H, Wd, -1))
        H, Wd = q.shape[1:3]
        q = q.reshape(B, H * Wd, heads, -1)
    o = sdpa(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)).transpose(1, 2)
    o = o.reshape(B, H, Wd, -1)
    return linear(o, w[p + "proj.weight"], w[p + "proj.bias"])


def multiscale_block(w: W, i: int, c: dict, x: Tensor) -> Tensor:
    """ref: hieradet.py:134-166."""
    p = f"image_encoder.trunk.blocks.{i}."
    shortcut = x
    x = layer_norm(x, w[p + "norm1.weight"], w[p + "norm1.bias"], 1e-6)
    if c["dim"] != c["dim_out"]:
        shortcut = _maxpool2(linear(x, w[p + "proj.weight"], w[p + "proj.bias"]))
    ws = c["window"]
    H, Wd = x.shape[1], x.shape[2]
    if ws > 0:
        x = window_partition(x, ws)
    x = multiscale_attention(w, p + "attn.", x, c["heads"], c["q_pool"])
    if c["q_pool"]:
        ws //= 2
        H, Wd = shortcut.shape[1:3]
    if c["window"] > 0:
        x = window_unpartition(x, ws, H, Wd)
    x = shortcut + x
    h = layer_norm(x, w[p + "norm2.weight"], w[p + "norm2.bias"], 1e-6)
    h = linear(h, w[p + "mlp.layers.0.weight"], w[p + "mlp.layers.0.bias"])
    h = F.gelu(h)  # exact erf GELU (nn.GELU default), ref: hieradet.py:93, sam2_utils.py:131-136
    h = linear(h, w[p + "mlp.layers.1.weight"], w[p + "mlp.layers.1.bias"])
    return x + h


def patch_embed(w: W, img: Tensor) -> Tensor:
    """ref: modeling/backbones/utils.py:89-93 (Conv2d 3->144 k7 s4 p3, NCHW->NHWC)."""
    t = "image_encoder.trunk.patch_embed.proj."
    return F.conv2d(_r(img), _r(w[t + "weight"]), w[t + "bias"], stride=4, padding=3).permute(0, 2, 3, 1)


def hiera_trunk(w: W, img: Tensor, num_blocks: Optional[int] = None, return_all=False) -> List[Tensor]:
    """ref: hieradet.py:283-299. img [N,3,H,W] normalised -> 4 NCHW stage outputs."""
    x = patch_embed(w, img)
    x = x + hiera_pos_embed(w, x.shape[1], x.shape[2])
    outs, allx = [], [x]
    for i, c in enumerate(block_cfgs()):
        if num_blocks is not None and i >= num_blocks:
            break
        x = multiscale_block(w, i, c, x)
        if return_all:
            allx.append(x)
        if i in STAGE_ENDS:
            outs.append(x.permute(0, 3, 1, 2))
    return allx if return_all else outs


_SINE_CACHE = {}


def position_sine(num_pos_feats: int, H: int, Wd: int, temperature=10000.0) -> Tensor:
    """ref: modeling/position_encoding.py:90-124 (normalize=True, scale=2*pi). -> [C,H,W]"""
    key = (num_pos_feats, H, Wd)
    if key in _SINE_CACHE:
        return _SINE_CACHE[key]
    npf = num_pos_feats // 2
    y = torch.arange(1, H + 1, dtype=torch.float32).view(-1, 1).repeat(1, Wd)
    x = torch.arange(1, Wd + 1, dtype=torch.float32).view(1, -1).repeat(H, 1)
    eps, scale = 1e-6, 2 * math.pi
    y = y / (y[-1:, :] + eps) * scale
    x = x / (x[:, -1:] + eps) * scale
    dim_t = torch.arange(npf, dtype=torch.float32)
    dim_t = temperature ** (2 * (dim_t // 2) / npf)
    px, py = x[:, :, None] / dim_t, y[:, :, None] / dim_t
    px = torch.stack((px[:, :, 0::2].sin(), px[:, :, 1::2].cos()), dim=3).flatten(2)
    py = torch.stack((py[:, :, 0::2].sin(), py[:, :, 1::2].cos()), dim=3).flatten(2)
    pos = torch.cat((py, px), dim=2).permute(2, 0, 1).contiguous()
    _SINE_CACHE[key] = pos
    return pos


def fpn_neck(w: W, xs: List[Tensor]) -> Tuple[List[Tensor], List[Tensor]]:
    """ref: modeling/backbones/image_encoder.py:102-134 (nearest top-down into level 2 only)."""
    n = len(xs) - 1
    out, pos, prev = [None] * 4, [None] * 4, None
    for i in range(n, -1, -1):
        p = f"image_encoder.neck.convs.{n - i}.conv."
        lat = conv1x1(xs[i], w[p + "weight"], w[p + "bias"])
        if i in (2, 3) and prev is not None:
            prev = lat + F.interpolate(prev, scale_factor=2.0, mode="nearest")
        else:
            prev = lat
        out[i] = prev
        pos[i] = position_sine(D_MODEL, prev.shape[-2], prev.shape[-1])[None].expand(prev.shape[0], -1, -1, -1)
    return out, pos


def forward_image(w: W, img_norm: Tensor):
    """Boundary E2 — ref: modeling/sam2_base_official.py:548-582 (+conv_s0/conv_s1), scalp=1
    (image_encoder.py:32-34). Returns (vision_features, pos0, pos1, pos2, fpn0, fpn1, fpn2)."""
    feats, pos = fpn_neck(w, hiera_trunk(w, img_norm))
    feats, pos = feats[:-1], pos[:-1]
    md = "sam_mask_decoder."
    f0 = conv1x1(feats[0], w[md + "conv_s0.weight"], w[md + "conv_s0.bias"])
    f1 = conv1x1(feats[1], w[md + "conv_s1.weight"], w[md + "conv_s1.bias"])
    return feats[2], pos[0], pos[1], pos[2], f0, f1, feats[2]


def set_image_e2e(w: W, img01: Tensor):
    """Boundary E1 — ref: sam2_image_predictor.py:252-266. img in [0,1], un-normalised.
    Returns (feat0[N,32,256,256], feat1[N,64,128,128], feat2[N,256,64,64] + no_mem_embed)."""
    o = forward_image(w, normalize_image(img01))
    return o[4], o[5], o[6] + w["no_mem_embed"].view(1, -1, 1, 1)


# ---------------------------------------------------------------------------------------
# Memory attention (rows a13-a17)
# ---------------------------------------------------------------------------------------
_ROPE_CACHE = {}


def rope_cos_sin(dim=256, end_x=64, end_y=64, theta=10000.0):
    """ref: modeling/position_encoding_fix.py:166-183. -> cos, sin [end_x*end_y, dim]"""
    key = (dim, end_x, end_y, theta)
    if key not in _ROPE_CACHE:
        freqs = 1.0 / (theta ** (torch.arange(0, dim, 2).float() / dim))
        t = torch.arange(end_x * end_y, dtype=torch.float32)
        tx, ty = (t % end_x).float(), torch.div(t, end_x, rounding_mode="floor").float()
        fr = torch.cat([torch.outer(tx, freqs), torch.outer(ty, freqs)], dim=-1)
        _ROPE_CACHE[key] = (fr.cos(), fr.sin())
    return _ROPE_CACHE[key]


def apply_rope(x: Tensor, cos: Tensor, sin: Tensor) -> Tensor:
    """ref: position_encoding_fix.py:192-205 — interleaved pairs (2i,2i+1) rotate by the
    angle stored at column 2i of the [seq, dim] table."""
    x1, x2 = x[..., ::2], x[..., 1::2]
    c, s = cos[..., ::2], sin[..., ::2]
    return torch.stack([x1 * c - x2 * s, x1 * s + x2 * c], dim=-1).flatten(-2)


def rope_attention(w: W, p: str, q: Tensor, k: Tensor, v: Tensor, num_k_exclude_rope: int) -> Tensor:
    """ref: modeling/sam/transformer.py:345-424 (num_heads=1). q,k,v: [B, n, c]."""
    q = linear(q, w[p + "q_proj.weight"], w[p + "q_proj.bias"])
    k = linear(k, w[p + "k_proj.weight"], w[p + "k_proj.bias"])
    v = linear(v, w[p + "v_proj.weight"], w[p + "v_proj.bias"])
    cos, sin = rope_cos_sin(q.shape[-1], int(math.sqrt(q.shape[-2])), int(math.sqrt(q.shape[-2])))
    q = apply_rope(q, cos, sin)
    rep = k.shape[-2] // q.shape[-2]
    cos_k, sin_k = cos.repeat(rep, 1), sin.repeat(rep, 1)
    n_rope = k.shape[-2] - num_k_exclude_rope
    k = torch.cat([apply_rope(k[:, :n_rope], cos_k[:n_rope], sin_k[:n_rope]), k[:, n_rope:]], dim=1)
    o = sdpa(q[:, None], k[:, None], v[:, None])[:, 0]
    return linear(o, w[p + "out_proj.weight"], w[p + "out_proj.bias"])


def memory_attention(w: W, curr: Tensor, memory: Tensor, curr_pos: Tensor, memory_pos: Tensor,
                     memory_exclude: Tensor, memory_pos_exclude: Tensor, num_layers=4) -> Tensor:
    """Boundary E3 — ref: modeling/memory_attention.py:299-349 (+ layer :59-109).
    curr/curr_pos [4096,N,256]; memory/memory_pos [L,4096,N,64]; *_exclude [P,N,64]."""
    mem = memory.flatten(0, 1)
    mpos = memory_pos.flatten(0, 1)
    P = memory_exclude.shape[0]
    if P > 0:
        mem = torch.cat([mem, memory_exclude], 0)
        mpos = torch.cat([mpos, memory_pos_exclude], 0)
    x = (curr + 0.1 * curr_pos).transpose(0, 1)
    mem, mpos = mem.transpose(0, 1), mpos.transpose(0, 1)
    for l in range(num_layers):
        p = f"memory_attention.layers.{l}."
        t = layer_norm(x, w[p + "norm1.weight"], w[p + "norm1.bias"], 1e-5)
        x = x + rope_attention(w, p + "self_attn.", t, t, t, 0)
        t = layer_norm(x, w[p + "norm2.weight"], w[p + "norm2.bias"], 1e-5)
        x = x + rope_attention(w, p + "cross_attn_image.", t, mem + mpos, mem, P)
        t = layer_norm(x, w[p + "norm3.weight"], w[p + "norm3.bias"], 1e-5)
        t = linear(F.relu(linear(t, w[p + "linear1.weight"], w[p + "linear1.bias"])),
                   w[p + "linear2.weight"], w[p + "linear2.bias"])
        x = x + t
    x = layer_norm(x, w["memory_attention.norm.weight"], w["memory_attention.norm.bias"], 1e-5)
    return x.transpose(0, 1)


# ---------------------------------------------------------------------------------------
# Memory encoder (row a18)
# ---------------------------------------------------------------------------------------
def conv2d(x, w_, b, stride=1, padding=0, groups=1):
    return F.conv2d(_r(x), _r(w_), b, stride=stride, padding=padding, groups=groups)


def mask_downsampler(w: W, masks: Tensor) -> Tensor:
    """ref: modeling/memory_encoder.py:19-60 (k3 s2 p1, 1->4->16->64->256, then 1x1)."""
    e = "memory_encoder.mask_downsampler.encoder."
    x = masks
    for j in range(4):
        x = conv2d(x, w[f"{e}{3 * j}.weight"], w[f"{e}{3 * j}.bias"], stride=2, padding=1)
        x = F.gelu(layer_norm_2d(x, w[f"{e}{3 * j + 1}.weight"], w[f"{e}{3 * j + 1}.bias"]))
    return conv2d(x, w[e + "12.weight"], w[e + "12.bias"])


def cx_block(w: W, p: str, x: Tensor) -> Tensor:
    """ref: memory_encoder.py:106-119."""
    inp = x
    x = F.conv2d(x, w[p + "dwconv.weight"], w[p + "dwconv.bias"], padding=3, groups=x.shape[1])
    x = layer_norm_2d(x, w[p + "norm.weight"], w[p + "norm.bias"]).permute(0, 2, 3, 1)
    x = F.gelu(linear(x, w[p + "pwconv1.weight"], w[p + "pwconv1.bias"]))
    x = w[p + "gamma"] * linear(x, w[p + "pwconv2.weight"], w[p + "pwconv2.bias"])
    return inp + x.permute(0, 3, 1, 2)


def memory_encoder(w: W, pix_feat: Tensor, masks: Tensor) -> Tuple[Tensor, Tensor]:
    """Boundary E4 — ref: memory_encoder.py:233-241. masks already sigmoid-scaled.
    -> (x[N,64,64,64], pos[N,64,64,64])"""
    m = "memory_encoder."
    x = conv2d(pix_feat, w[m + "pix_feat_proj.weight"], w[m + "pix_feat_proj.bias"]) + mask_downsampler(w, masks)
    for l in range(2):
        x = cx_block(w, f"{m}fuser.layers.{l}.", x)
    x = conv2d(x, w[m + "out_proj.weight"], w[m + "out_proj.bias"])
    pos = position_sine(MEM_DIM, x.shape[-2], x.shape[-1])[None].expand(x.shape[0], -1, -1, -1)
    return x, pos


# ---------------------------------------------------------------------------------------
# SAM heads (fp32; SURVEY §8f-1 — not on the accelerated path but needed for mask IoU)
# ---------------------------------------------------------------------------------------
def _mlp(w: W, p: str, x: Tensor, n: int, sigmoid=False) -> Tensor:
    """ref: modeling/sam2_utils.py:112-136 (ReLU between layers)."""
    for i in range(n):
        x = F.linear(x, w[f"{p}layers.{i}.weight"], w[f"{p}layers.{i}.bias"])
        if i < n - 1:
            x = F.relu(x)
    return torch.sigmoid(x) if sigmoid else x


def _pe_encoding(w: W, coords01: Tensor) -> Tensor:
    """ref: modeling/position_encoding.py:146-154."""
    c = (2 * coords01 - 1) @ w["sam_prompt_encoder.pe_layer.positional_encoding_gaussian_matrix"]
    c = 2 * math.pi * c
    return torch.cat([torch.sin(c), torch.cos(c)], dim=-1)


def dense_pe(w: W, size=64) -> Tensor:
    """ref: position_encoding.py:156-167, prompt_encoder.py:113-122. -> [1,256,64,64]"""
    g = torch.ones((size, size), dtype=torch.float32)
    y = (g.cumsum(0) - 0.5) / size
    x = (g.cumsum(1) - 0.5) / size
    return _pe_encoding(w, torch.stack([x, y], dim=-1)).permute(2, 0, 1)[None]


def prompt_encoder(w: W, coords: Optional[Tensor], labels: Optional[Tensor], masks: Optional[Tensor], bs=1):
    """ref: modeling/sam/prompt_encoder.py:124-231 (points + optional dense mask; boxes are
    passed as points with labels 2/3 by both predictors)."""
    pe = "sam_prompt_encoder."
    sparse = torch.empty((bs, 0, D_MODEL))
    if coords is not None:
        bs = coords.shape[0]
        pts = torch.cat([coords + 0.5, torch.zeros((bs, 1, 2))], 1)
        lab = torch.cat([labels, -torch.ones((bs, 1), dtype=labels.dtype)], 1)
        emb = _pe_encoding(w, pts / IMAGE_SIZE)
        emb = torch.where((lab == -1)[..., None], torch.zeros_like(emb) + w[pe + "not_a_point_embed.weight"], emb)
        for i in range(4):
            emb = torch.where((lab == i)[..., None], emb + w[pe + f"point_embeddings.{i}.weight"], emb)
        sparse = emb
    if masks is not None:
        d = pe + "mask_downscaling."
        x = F.conv2d(masks, w[d + "0.weight"], w[d + "0.bias"], stride=2)
        x = F.gelu(layer_norm_2d(x, w[d + "1.weight"], w[d + "1.bias"]))
        x = F.conv2d(x, w[d + "3.weight"], w[d + "3.bias"], stride=2)
        x = F.gelu(layer_norm_2d(x, w[d + "4.weight"], w[d + "4.bias"]))
        dense = F.conv2d(x, w[d + "6.weight"], w[d + "6.bias"])
    else:
        dense = w[pe + "no_mask_embed.weight"].reshape(1, -1, 1, 1).expand(bs, -1, 64, 64)
    return sparse, dense


def _attention(w: W, p: str, q, k, v, heads=8):
    """ref: modeling/sam/transformer.py:222-294."""
    q = F.linear(q, w[p + "q_proj.weight"], w[p + "q_proj.bias"])
    k = F.linear(k, w[p + "k_proj.weight"], w[p + "k_proj.bias"])
    v = F.linear(v, w[p + "v_proj.weight"], w[p + "v_proj.bias"])
    sh = lambda x: x.reshape(x.shape[0], x.shape[1], heads, -1).transpose(1, 2)
    o = F.scaled_dot_product_attention(sh(q), sh(k), sh(v)).transpose(1, 2)
    o = o.reshape(o.shape[0], o.shape[1], -1)
    return F.linear(o, w[p + "out_proj.weight"], w[p + "out_proj.bias"])


def two_way_transformer(w: W, src: Tensor, pos: Tensor, tokens: Tensor):
    """ref: modeling/sam/transformer.py:97-219 (depth 2, 8 heads, downsample 2)."""
    t = "sam_mask_decoder.transformer."
    keys = src.flatten(2).permute(0, 2, 1)
    kpe = pos.flatten(2).permute(0, 2, 1)
    queries, qpe = tokens, tokens
    ln = lambda x, n: layer_norm(x, w[n + "weight"], w[n + "bias"], 1e-5)
    for l in range(2):
        p = f"{t}layers.{l}."
        if l == 0:
            queries = _attention(w, p + "self_attn.", queries, queries, queries)
        else:
            q = queries + qpe
            queries = queries + _attention(w, p + "self_attn.", q, q, queries)
        queries = ln(queries, p + "norm1.")
        queries = ln(queries + _attention(w, p + "cross_attn_token_to_image.", queries + qpe, keys + kpe, keys), p + "norm2.")
        queries = ln(queries + _mlp(w, p + "mlp.", queries, 2), p + "norm3.")
        keys = ln(keys + _attention(w, p + "cross_attn_image_to_token.", keys + kpe, queries + qpe, queries), p + "norm4.")
    queries = queries + _attention(w, t + "final_attn_token_to_image.", queries + qpe, keys + kpe, keys)
    return ln(queries, t + "norm_final_attn."), keys


def mask_decoder(w: W, image_emb: Tensor, sparse: Tensor, dense: Tensor, multimask: bool,
                 high_res: List[Tensor], repeat_image=False, dynamic_multimask=True):
    """ref: modeling/sam/mask_decoder.py:116-224, 262-316, 337-382."""
    md = "sam_mask_decoder."
    out_tok = torch.cat([w[md + "obj_score_token.weight"], w[md + "iou_token.weight"], w[md + "mask_tokens.weight"]], 0)
    tokens = torch.cat((out_tok[None].expand(sparse.shape[0], -1, -1), sparse), dim=1)
    src = torch.repeat_interleave(image_emb, tokens.shape[0], dim=0) if repeat_image else image_emb
    src = src + dense
    pos = torch.repeat_interleave(dense_pe(w), tokens.shape[0], dim=0)
    b, c, h, wd = src.shape
    hs, src = two_way_transformer(w, src, pos, tokens)
    iou_tok, mask_toks = hs[:, 1], hs[:, 2:6]
    src = src.transpose(1, 2).view(b, c, h, wd)
    u = md + "output_upscaling."
    x = F.conv_transpose2d(src, w[u + "0.weight"], w[u + "0.bias"], stride=2) + high_res[1]
    x = F.gelu(layer_norm_2d(x, w[u + "1.weight"], w[u + "1.bias"]))
    x = F.gelu(F.conv_transpose2d(x, w[u + "3.weight"], w[u + "3.bias"], stride=2) + high_res[0])
    hyper = torch.stack([_mlp(w, f"{md}output_hypernetworks_mlps.{i}.", mask_toks[:, i], 3) for i in range(4)], 1)
    b, c, h, wd = x.shape
    masks = (hyper @ x.view(b, c, h * wd)).view(b, -1, h, wd)
    iou = _mlp(w, md + "iou_prediction_head.", iou_tok, 3, sigmoid=True)
    obj = _mlp(w, md + "pred_obj_score_head.", hs[:, 0], 3)
    if multimask:
        masks, iou = masks[:, 1:], iou[:, 1:]
        sam_tok = mask_toks[:, 1:]
    else:
        sam_tok = mask_toks[:, 0:1]
        if dynamic_multimask:
            masks, iou = _dynamic_multimask(masks, iou)
        else:
            masks, iou = masks[:, 0:1], iou[:, 0:1]
    return masks, iou, sam_tok, obj


def _dynamic_multimask(all_logits, all_iou, delta=0.05, thresh=0.98):
    """ref: mask_decoder.py:337-382."""
    mm, mi = all_logits[:, 1:], all_iou[:, 1:]
    best = torch.argmax(mi, dim=-1)
    bi = torch.arange(mi.shape[0])
    best_logits, best_iou = mm[bi, best][:, None], mi[bi, best][:, None]
    single, single_iou = all_logits[:, 0:1], all_iou[:, 0:1]
    fl = single.flatten(-2)
    ai, au = (fl > delta).sum(-1).float(), (fl > -delta).sum(-1).float()
    stable = torch.where(au > 0, ai / au, 1.0) >= thresh
    return (torch.where(stable[..., None, None].expand_as(single), single, best_logits),
            torch.where(stable.expand_as(single_iou), single_iou, best_iou))


def forward_sam_heads(w: W, backbone_features: Tensor, point_inputs=None, mask_inputs=None,
                      high_res_features=None, multimask_output=False):
    """ref: modeling/sam2_base_official.py:338-494."""
    B = backbone_features.shape[0]
    if point_inputs is not None:
        coords, labels = point_inputs["point_coords"], point_inputs["point_labels"]
    else:
        coords, labels = torch.zeros(B, 1, 2), -torch.ones(B, 1, dtype=torch.int32)
    mask_prompt = None
    if mask_inputs is not None:
        mask_prompt = mask_inputs
        if mask_inputs.shape[-2:] != (256, 256):
            mask_prompt = F.interpolate(mask_inputs.float(), size=(256, 256), align_corners=False,
                                        mode="bilinear", antialias=True)
    sparse, dense = prompt_encoder(w, coords, labels, mask_prompt)
    low, ious, sam_tokens, obj_logits = mask_decoder(w, backbone_features, sparse, dense, multimask_output,
                                                     high_res_features)
    is_obj = obj_logits > 0
    low = torch.where(is_obj[:, None, None], low, torch.tensor(NO_OBJ_SCORE)).float()
    high = F.interpolate(low, size=(IMAGE_SIZE, IMAGE_SIZE), mode="bilinear", align_corners=False)
    tok = sam_tokens[:, 0]
    if multimask_output:
        best = torch.argmax(ious, dim=-1)
        bi = torch.arange(B)
        low_best, high_best = low[bi, best][:, None], high[bi, best][:, None]
        if sam_tokens.shape[1] > 1:
            tok = sam_tokens[bi, best]
    else:
        low_best, high_best = low, high
    ptr = _mlp(w, "obj_ptr_proj.", tok, 3)
    lam = is_obj.float()
    ptr = lam * ptr + (1 - lam) * w["no_obj_ptr"]
    return low, high, ious, low_best, high_best, ptr, obj_logits


def use_mask_as_output(w: W, backbone_features: Tensor, high_res_features, mask_inputs: Tensor):
    """A mask prompt turned straight into the frame's output (what `add_new_mask` / the VOS tool feed) — ref:
    modeling/sam2_base_official.py:496-545 (_use_mask_as_output): +10 / -10 logits, antialiased x1/4 for the low-res copy, IoU
    of one, the object pointer from the SAM heads run on `mask_downsample(mask)` as their mask prompt, and the MASK (not the
    decoder) deciding whether the object is there.  Same 7 outputs as forward_sam_heads."""
    m = mask_inputs.float()
    high = m * 20.0 - 10.0
    low = F.interpolate(high, size=(high.shape[-2] // 4, high.shape[-1] // 4), align_corners=False, mode="bilinear", antialias=True)
    ious = torch.ones(m.shape[0], 1)
    prompt = conv2d(m, w["mask_downsample.weight"], w["mask_downsample.bias"], stride=4)
    ptr = forward_sam_heads(w, backbone_features, mask_inputs=prompt, high_res_features=high_res_features)[5]
    lam = (m.flatten(1) > 0).any(dim=1, keepdim=True).float()
    obj_logits = 20.0 * lam - 10.0
    ptr = lam * ptr + (1 - lam) * w["no_obj_ptr"]
    return low, high, ious, low, high, ptr, obj_logits


# ---------------------------------------------------------------------------------------
# Predictor-level glue (configs 1 and 3): image predict + single-direction video propagation
# ---------------------------------------------------------------------------------------
def image_predict(w: W, image_u8_hwc, point_xy, point_labels, multimask_output=True):
    """Config 1 — ref: sam2_image_predictor.py:141-202 (set_image) + :388-589 (predict), for a
    1024x1024 input (resize is the identity). Returns (mask_logits[C,H,W], iou[C], low_res[C,256,256])."""
    img = torch.as_tensor(image_u8_hwc).permute(2, 0, 1)[None].float() / 255.0
    f0, f1, f2 = set_image_e2e(w, img)
    H, Wd = image_u8_hwc.shape[:2]
    pts = torch.as_tensor(point_xy, dtype=torch.float32).clone()
    pts[..., 0] = pts[..., 0] / Wd * IMAGE_SIZE
    pts[..., 1] = pts[..., 1] / H * IMAGE_SIZE
    lab = torch.as_tensor(point_labels, dtype=torch.int32)
    sparse, dense = prompt_encoder(w, pts[None], lab[None], None)
    low, iou, _, _ = mask_decoder(w, f2, sparse, dense, multimask_output, [f0, f1])
    masks = F.interpolate(low.float(), (H, Wd), mode="bilinear", align_corners=False)
    return masks[0], iou[0], torch.clamp(low, -32.0, 32.0)[0]


def get_1d_sine_pe(pos: Tensor, dim: int, temperature=10000.0) -> Tensor:
    """ref: modeling/sam2_utils.py:64-74."""
    pe_dim = dim // 2
    dim_t = torch.arange(pe_dim, dtype=torch.float32)
    dim_t = temperature ** (2 * (dim_t // 2) / pe_dim)
    e = pos.unsqueeze(-1) / dim_t
    return torch.cat([e.sin(), e.cos()], dim=-1)


def assemble_memory(w: W, frame_idx: int, cond_outputs: dict, non_cond_outputs: dict, num_frames: int):
    """ref: sam2_base_official.py:825-950 for forward tracking, stride 1, no cond-frame limit.
    Each stored output holds 'maskmem_features' [1,64,64,64] (bf16-rounded, see
    sam2_video_predictor_official.py:887), 'maskmem_pos' [1,64,64,64] and 'obj_ptr' [1,256].
    Returns the six tensors of boundary E3 after `curr`/`curr_pos`: memory [L,4096,1,64],
    memory_pos, memory_exclude [P,1,64], memory_pos_exclude."""
    t_pos_prev = [(0, o) for o in cond_outputs.values()]
    for t_pos in range(1, NUM_MASKMEM):
        t_rel = NUM_MASKMEM - t_pos
        t_pos_prev.append((t_pos, non_cond_outputs.get(frame_idx - t_rel)))
    mem, mpos = [], []
    for t_pos, prev in t_pos_prev:
        if prev is None:
            continue
        mem.append(prev["maskmem_features"].float().flatten(2).permute(2, 0, 1))
        e = prev["maskmem_pos"].flatten(2).permute(2, 0, 1)
        mpos.append(e + w["maskmem_tpos_enc"][NUM_MASKMEM - t_pos - 1])
    max_ptrs = min(num_frames, 16)
    pos_ptrs = [(frame_idx - t, o["obj_ptr"]) for t, o in cond_outputs.items() if t <= frame_idx]
    for t_diff in range(1, max_ptrs):
        t = frame_idx - t_diff
        if t < 0:
            break
        o = non_cond_outputs.get(t)
        if o is not None:
            pos_ptrs.append((t_diff, o["obj_ptr"]))
    pos_list, ptrs = zip(*pos_ptrs)
    ptrs = torch.stack(ptrs, 0)                                             # [n,1,256]
    opos = get_1d_sine_pe(torch.tensor(pos_list, dtype=torch.float32) / (max_ptrs - 1), D_MODEL)
    opos = F.linear(opos, w["obj_ptr_tpos_proj.weight"], w["obj_ptr_tpos_proj.bias"])
    opos = opos.unsqueeze(1).expand(-1, 1, MEM_DIM)
    ptrs = ptrs.reshape(-1, 1, D_MODEL // MEM_DIM, MEM_DIM).permute(0, 2, 1, 3).flatten(0, 1)
    opos = opos.repeat_interleave(D_MODEL // MEM_DIM, dim=0)
    return torch.stack(mem, 0), torch.stack(mpos, 0), ptrs, opos


def encode_new_memory(w: W, pix_feat: Tensor, high_res_masks: Tensor, obj_logits: Tensor, from_points: bool):
    """ref: sam2_base_official.py:978-1026 (binarize on clicked frames, sigmoid*20-10, +no_obj_embed_spatial)."""
    m = (high_res_masks > 0).float() if from_points else torch.sigmoid(high_res_masks)
    x, pos = memory_encoder(w, pix_feat, m * 20.0 - 10.0)
    is_obj = (obj_logits > 0).float()
    x = x + (1 - is_obj[..., None, None]) * w["no_obj_embed_spatial"][..., None, None].expand(*x.shape)
    return x, pos


def video_propagate(w: W, frames_norm: Tensor, click_xy=(512.0, 512.0), fwd_image=None, mem_attn=None, mem_enc=None):
    """Config 3 — one object, one positive click on frame 0, forward propagation.
    ref: sam2_video_predictor_official.py:267-399 (add_new_points_or_box), :586-736
    (preflight + propagate), :843-909; sam2_base_official.py:1028-1179 (track_step).
    ``fwd_image``/``mem_attn``/``mem_enc`` let a test swap in another implementation of the
    E2/E3/E4 boundaries (same signatures as the oracle's) while keeping this glue.
    Yields (frame_idx, low_res_mask_logits[1,1,256,256])."""
    fwd_image = fwd_image or (lambda x: forward_image(w, x))
    mem_attn = mem_attn or (lambda *a: memory_attention(w, *a))
    mem_enc = mem_enc or (lambda p, m: memory_encoder(w, p, m))
    T = frames_norm.shape[0]

    def feats(t):
        o = fwd_image(frames_norm[t:t + 1])
        return o[6], o[3], [o[4], o[5]]      # fpn2 [1,256,64,64], pos2, high-res

    def new_memory(f2, high, obj_logits, from_points):
        m = (high > 0).float() if from_points else torch.sigmoid(high)
        x, pos = mem_enc(f2, m * 20.0 - 10.0)
        is_obj = (obj_logits > 0).float()
        x = x + (1 - is_obj[..., None, None]) * w["no_obj_embed_spatial"][..., None, None].expand(*x.shape)
        return x.to(torch.bfloat16), pos

    # frame 0: click -> SAM heads on (fpn2 + no_mem_embed); memory encoded in the preflight
    f2, pos2, hr = feats(0)
    pts = {"point_coords": torch.tensor([[list(click_xy)]], dtype=torch.float32),
           "point_labels": torch.tensor([[1]], dtype=torch.int32)}
    pix = f2 + w["no_mem_embed"].view(1, -1, 1, 1)
    _, _, _, low, _, ptr, obj = forward_sam_heads(w, pix, point_inputs=pts, high_res_features=hr, multimask_output=True)
    high = F.interpolate(low, size=(IMAGE_SIZE, IMAGE_SIZE), mode="bilinear", align_corners=False)
    mf, mp = new_memory(f2, high, obj, True)
    cond = {0: dict(maskmem_features=mf, maskmem_pos=mp, obj_ptr=ptr)}
    non_cond = {}
    yield 0, low
    for t in range(1, T):
        f2, pos2, hr = feats(t)
        curr = f2.flatten(2).permute(2, 0, 1)
        curr_pos = pos2.flatten(2).permute(2, 0, 1)
        mem = assemble_memory(w, t, cond, non_cond, T)
        pix = mem_attn(curr, mem[0], curr_pos, mem[1], mem[2], mem[3])
        pix = pix.permute(1, 2, 0).view(1, D_MODEL, 64, 64)
        _, _, _, low, high, ptr, obj = forward_sam_heads(w, pix, high_res_features=hr, multimask_output=True)
        mf, mp = new_memory(f2, high, obj, False)
        non_cond[t] = dict(maskmem_features=mf, maskmem_pos=mp, obj_ptr=ptr)
        yield t, low


def mask_iou(a_logits: Tensor, b_logits: Tensor) -> float:
    a, b = a_logits > 0, b_logits > 0
    u = (a | b).sum().item()
    return 1.0 if u == 0 else (a & b).sum().item() / u
