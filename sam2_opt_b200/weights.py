"""One-time weight preparation at ``speedup()`` time: constants that the reference recomputes on every call
(or bakes into its ONNX/TensorRT engine) are derived once here with torch and handed to the engine as
``derived.*`` tensors.  Nothing in this file runs on the per-frame path.

All ``ref:`` citations are relative to /root/reference/sam2/sam2/.
"""
import math
from typing import Dict

import torch
import torch.nn.functional as F

IMAGE_SIZE = 1024
D_MODEL = 256
MEM_DIM = 64


def hiera_pos_embed(w: Dict[str, torch.Tensor], size: int = 256) -> torch.Tensor:
    """ref: modeling/backbones/hieradet.py:273-281 (bicubic background + tiled window embedding) -> [size*size, C]"""
    t = "image_encoder.trunk."
    pe = F.interpolate(w[t + "pos_embed"].float(), size=(size, size), mode="bicubic")
    win = w[t + "pos_embed_window"].float()
    pe = pe + win.tile([x // y for x, y in zip(pe.shape, win.shape)])
    return pe.permute(0, 2, 3, 1).reshape(size * size, -1).contiguous()


def position_sine(num_pos_feats: int, H: int, Wd: int, device, temperature: float = 10000.0) -> torch.Tensor:
    """ref: modeling/position_encoding.py:90-124 (PositionEmbeddingSine, normalize=True, scale=2*pi) -> [C,H,W].
    The reference materialises this with .repeat(B,...) on every call; we keep one constant and expand views."""
    npf = num_pos_feats // 2
    y = torch.arange(1, H + 1, dtype=torch.float32, device=device).view(-1, 1).repeat(1, Wd)
    x = torch.arange(1, Wd + 1, dtype=torch.float32, device=device).view(1, -1).repeat(H, 1)
    eps, scale = 1e-6, 2 * math.pi
    y = y / (y[-1:, :] + eps) * scale
    x = x / (x[:, -1:] + eps) * scale
    dim_t = torch.arange(npf, dtype=torch.float32, device=device)
    dim_t = temperature ** (2 * torch.div(dim_t, 2, rounding_mode="floor") / npf)
    px, py = x[:, :, None] / dim_t, y[:, :, None] / dim_t
    px = torch.stack((px[:, :, 0::2].sin(), px[:, :, 1::2].cos()), dim=3).flatten(2)
    py = torch.stack((py[:, :, 0::2].sin(), py[:, :, 1::2].cos()), dim=3).flatten(2)
    return torch.cat((py, px), dim=2).permute(2, 0, 1).contiguous()


def rope_tables(device, dim: int = 256, end: int = 64, theta: float = 10000.0):
    """ref: modeling/position_encoding_fix.py:166-183 -> (cos, sin) [end*end, dim/2]: angle of pair i at position r
    (pairs i < dim/4 rotate with x, the rest with y)."""
    freqs = 1.0 / (theta ** (torch.arange(0, dim, 4, device=device)[: dim // 4].float() / dim))
    t = torch.arange(end * end, dtype=torch.float32, device=device)
    tx, ty = (t % end).float(), torch.div(t, end, rounding_mode="floor").float()
    ang = torch.cat([torch.outer(tx, freqs), torch.outer(ty, freqs)], dim=-1)     # [4096, dim/2]
    return ang.cos().contiguous(), ang.sin().contiguous()


def split3(wt: torch.Tensor) -> torch.Tensor:
    """[N,K] fp32 -> [N,3K] = (hi | hi | lo) with hi = bf16(w), lo = bf16(w - hi), every entry exactly representable in bf16:
    the B operand of the split-bf16 GEMMs of the SAM heads (csrc/sam_heads.cu; the A operand is laid out (hi | lo | hi), so one
    bf16 GEMM over 3K columns accumulates hi*hi + lo*hi + hi*lo in fp32)."""
    wt = wt.float()
    hi = wt.to(torch.bfloat16).float()
    lo = (wt - hi).to(torch.bfloat16).float()
    return torch.cat([hi, hi, lo], 1).contiguous()


def derive_heads(w: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Constants of the SAM heads (engine.cu sam_heads_mb).  ref: modeling/sam/transformer.py:97-219 (the image-side
    projections of TwoWayAttentionBlock all act on keys + dense_pe: the dense_pe part is a per-position constant, folded here
    into a table that the GEMM adds as a row-modulo residual), modeling/sam/mask_decoder.py:283-300 (ConvTranspose2d k2 s2 as
    a GEMM over (dy,dx,cout))."""
    md, tr = "sam_mask_decoder.", "sam_mask_decoder.transformer."
    if md + "iou_token.weight" not in w or "sam_prompt_encoder.pe_layer.positional_encoding_gaussian_matrix" not in w:
        return {}
    from .sam_heads import dense_pe
    d = {}
    pe = dense_pe(w).flatten(2)[0].t().double()                                  # [4096, 256], position-major
    lin = lambda p: (w[p + "weight"].double(), w[p + "bias"].double())
    for l in range(2):
        p = f"{tr}layers.{l}."
        (wk, bk), (wv, bv) = lin(p + "cross_attn_token_to_image.k_proj."), lin(p + "cross_attn_token_to_image.v_proj.")
        wq, bq = lin(p + "cross_attn_image_to_token.q_proj.")
        d[f"derived.heads.{l}.kvq.weight"] = split3(torch.cat([wk, wv, wq], 0).float())
        d[f"derived.heads.{l}.kvq.table"] = torch.cat([pe @ wk.t() + bk, bv[None].expand(4096, -1), pe @ wq.t() + bq], 1).float().contiguous()
        d[f"derived.heads.{l}.i2t_out.weight"] = split3(w[p + "cross_attn_image_to_token.out_proj.weight"])
    (wk, bk), (wv, bv) = lin(tr + "final_attn_token_to_image.k_proj."), lin(tr + "final_attn_token_to_image.v_proj.")
    d["derived.heads.final.kv.weight"] = split3(torch.cat([wk, wv], 0).float())
    d["derived.heads.final.kv.table"] = torch.cat([pe @ wk.t() + bk, bv[None].expand(4096, -1)], 1).float().contiguous()
    u0, u3 = w[md + "output_upscaling.0.weight"].float(), w[md + "output_upscaling.3.weight"].float()
    d["derived.heads.up1.weight"] = split3(u0.permute(2, 3, 1, 0).reshape(256, 256))         # [(dy,dx,cout), cin]
    d["derived.heads.up2.weight"] = u3.permute(0, 2, 3, 1).reshape(64, 4, 32).contiguous()      # [cin][(ey,ex)][cout]
    return d


def check_architecture(w: Dict[str, torch.Tensor]) -> None:
    """The kernels are built for SAM2.1 Hiera-L (ref: configs/sam2.1/sam2.1_hiera_l.yaml: embed_dim 144, stages (2, 6, 36, 4),
    window_spec (8, 4, 16, 8), 7x7 / 8x8 position-embedding background / window).  Any other trunk (Hiera-T / S / B+ have head
    dims 96 / 56 and 14 x 14 windows that need padding) is refused here, by name, before anything is uploaded."""
    t = "image_encoder.trunk."
    pe = w.get(t + "patch_embed.proj.weight")
    if pe is None:
        raise ValueError("state_dict has no image_encoder.trunk.patch_embed.proj.weight: not a SAM2 model")
    n_blocks = 1 + max((int(k[len(t) + 7:].split(".")[0]) for k in w if k.startswith(t + "blocks.")), default=-1)
    win = tuple(w[t + "pos_embed_window"].shape[-2:]) if t + "pos_embed_window" in w else None
    got = (int(pe.shape[0]), n_blocks, win)
    if got != (144, 48, (8, 8)):
        raise ValueError(f"sam2_opt_b200 implements the SAM2.1 Hiera-L trunk (embed_dim 144, 48 blocks, 8x8 window embedding); this "
                         f"state_dict has embed_dim {got[0]}, {got[1]} blocks, window embedding {got[2]}: keep it on the torch path")


def derive(w: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Derived constants for the engine (all fp32, same device as the weights)."""
    check_architecture(w)
    d = {}
    dev = w["image_encoder.trunk.pos_embed"].device
    pw = w["image_encoder.trunk.patch_embed.proj.weight"].float().reshape(144, 147)
    d["derived.patch_embed.weight"] = F.pad(pw, (0, 13)).contiguous()            # K 147 -> 160
    d["derived.pos_embed"] = hiera_pos_embed(w)
    # FpnNeck laterals: level i uses convs[3 - i]  (ref: modeling/backbones/image_encoder.py:112-114)
    nk = "image_encoder.neck.convs.{}.conv."
    lat = {i: (w[nk.format(3 - i) + "weight"].double().flatten(1), w[nk.format(3 - i) + "bias"].double()) for i in range(4)}
    d["derived.neck.lat3.weight"] = lat[3][0].float().contiguous()
    d["derived.neck.lat3.bias"] = lat[3][1].float().contiguous()
    d["derived.neck.lat2.weight"] = lat[2][0].float().contiguous()
    d["derived.neck.lat2.bias"] = lat[2][1].float().contiguous()
    d["derived.neck.lat2.bias_nomem"] = (lat[2][1] + w["no_mem_embed"].double().view(-1)).float().contiguous()
    # conv_s0 / conv_s1 (ref: sam2_base_official.py:571-577) are 1x1 convs applied right after the level-0/1 lateral
    # 1x1 convs (no top-down term reaches those levels), so they compose exactly: W = Ws @ Wl, b = Ws @ bl + bs.
    for lvl, name in ((1, "s1"), (0, "s0")):
        ws_ = w[f"sam_mask_decoder.conv_{name}.weight"].double().flatten(1)
        bs_ = w[f"sam_mask_decoder.conv_{name}.bias"].double()
        d[f"derived.neck.{name}.weight"] = (ws_ @ lat[lvl][0]).float().contiguous()
        d[f"derived.neck.{name}.bias"] = (ws_ @ lat[lvl][1] + bs_).float().contiguous()
    # memory attention: self-attention q and k projections share their input -> one GEMM
    for l in range(4):
        sa = f"memory_attention.layers.{l}.self_attn."
        if sa + "q_proj.weight" in w:
            d[f"derived.memattn.{l}.sa_qk.weight"] = torch.cat([w[sa + "q_proj.weight"], w[sa + "k_proj.weight"]], 0).float().contiguous()
            d[f"derived.memattn.{l}.sa_qk.bias"] = torch.cat([w[sa + "q_proj.bias"], w[sa + "k_proj.bias"]], 0).float().contiguous()
    # cross-attention: the attention weights of a row sum to one, so softmax(.) (m Wv^T + bv) = (softmax(.) m) Wv^T + bv and
    # v_proj composes with out_proj into ONE 64 -> 256 map applied after the attention over the un-projected memory
    # (ref: modeling/sam/transformer.py:265-296 with kv_in_dim = 64; engine: attn_hd256_v64.cu)
    for l in range(4):
        ca = f"memory_attention.layers.{l}.cross_attn_image."
        if ca + "v_proj.weight" in w:
            Wv, bv = w[ca + "v_proj.weight"].double(), w[ca + "v_proj.bias"].double()
            Wo, bo = w[ca + "out_proj.weight"].double(), w[ca + "out_proj.bias"].double()
            d[f"derived.memattn.{l}.ca_ov.weight"] = (Wo @ Wv).float().contiguous()          # [256, 64]
            d[f"derived.memattn.{l}.ca_ov.bias"] = (Wo @ bv + bo).float().contiguous()
    # memory encoder: last down-sampler conv as an im2col GEMM with k = (ky,kx,cin); depthwise weights tap-major
    k9 = "memory_encoder.mask_downsampler.encoder.9.weight"
    if k9 in w:
        d["derived.memenc.ds4.weight"] = w[k9].float().permute(0, 2, 3, 1).reshape(256, 576).contiguous()
        # the 16 -> 64 channel stage in the same form: [64, (ky,kx,cin) = 144]
        d["derived.memenc.ds3.weight"] = w["memory_encoder.mask_downsampler.encoder.6.weight"].float().permute(0, 2, 3, 1).reshape(64, 144).contiguous()
        for l in range(2):
            dw = w[f"memory_encoder.fuser.layers.{l}.dwconv.weight"].float()
            d[f"derived.memenc.fuser.{l}.dw.weight"] = dw.reshape(256, 49).t().contiguous()
    # LayerNorm fold for the trunk (engine.cu image_encode_mb): every norm -> Linear pair of MultiScaleBlock
    # (ref: modeling/backbones/hieradet.py:134-166; norm1 -> attn.qkv [and -> proj on dim-change blocks], norm2 -> mlp.layers.0)
    # becomes one GEMM on the raw rows with W'' = row-centred W diag(gamma) and b' = b + W beta; rstd is applied per row in the
    # epilogue from statistics the producing GEMM emitted.
    t = "image_encoder.trunk.blocks."
    i = 0
    while f"{t}{i}.norm1.weight" in w:
        b = f"{t}{i}."
        pairs = [("norm1.", "attn.qkv.", "qkv"), ("norm2.", "mlp.layers.0.", "fc1")]
        if b + "proj.weight" in w:
            pairs.append(("norm1.", "proj.", "skip"))
        for norm, lin, short in pairs:
            gamma, beta = w[b + norm + "weight"].double(), w[b + norm + "bias"].double()
            W, bias = w[b + lin + "weight"].double(), w[b + lin + "bias"].double()
            ws = W * gamma[None, :]
            ws = ws - ws.mean(1, keepdim=True)        # zero-mean rows: sum_k x_k ws[n][k] = sum_k (x_k - mean(x)) (W gamma)[n][k]
            d[f"derived.blocks.{i}.{short}.weight"] = ws.float().contiguous()
            d[f"derived.blocks.{i}.{short}.bias"] = (bias + W @ beta).float().contiguous()
        i += 1
    d.update(derive_heads(w))
    cos, sin = rope_tables(dev)
    # (cos, sin) of column pair i at columns (2i, 2i + 1): the layout the GEMM's RoPE epilogue reads as 32 x 32 boxes
    d["derived.rope.cs"] = torch.stack([cos, sin], -1).reshape(cos.shape[0], -1).contiguous()
    return d
