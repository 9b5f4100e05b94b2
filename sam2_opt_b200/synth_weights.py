"""Deterministic synthetic ("random-init") SAM2.1 Hiera-L weights keyed by the reference's own state_dict names.

No checkpoint is available offline, so the benchmark, the smoke test, the oracle and the golden-fixture generator all
draw the weights from this one generator (values depend only on the tensor NAME and the seed, never on construction
order).  `oracle/make_golden.py` loads them into the unmodified reference with a strict `load_state_dict`, which is what
proves the names / shapes below are the reference's (903 tensors, 224.45 M parameters).
ref: sam2/sam2/configs/sam2.1/sam2.1_hiera_l.yaml, modeling/backbones/hieradet.py:208-266.
"""
import math
import zlib
from typing import Dict, List, Optional, Tuple

import torch

W = Dict[str, torch.Tensor]
EMBED_DIM = 144
STAGES = (2, 6, 36, 4)
GLOBAL_ATT_BLOCKS = (23, 33, 43)
WINDOW_SPEC = (8, 4, 16, 8)
STAGE_ENDS = (1, 7, 43, 47)
Q_POOL_BLOCKS = (2, 8, 44)
D_MODEL = 256
MEM_DIM = 64
NUM_MASKMEM = 7


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
        cfgs.append(dict(dim=dim, dim_out=dim_out, heads=heads, window=window, q_pool=i in Q_POOL_BLOCKS))
        dim = dim_out
    return cfgs


# ---------------------------------------------------------------------------------------
# Weight spec + deterministic synthetic ("random-init") weights
# ---------------------------------------------------------------------------------------
def _attn_spec(p, emb, internal, kv_in=None):
    kv_in = kv_in or emb
    return [(p + "q_proj.weight", (internal, emb), "lin"), (p + "q_proj.bias", (internal,), "lin_b", emb),
            (p + "k_proj.weight", (internal, kv_in), "lin"), (p + "k_proj.bias", (internal,), "lin_b", kv_in),
            (p + "v_proj.weight", (internal, kv_in), "lin"), (p + "v_proj.bias", (internal,), "lin_b", kv_in),
            (p + "out_proj.weight", (emb, internal), "lin"), (p + "out_proj.bias", (emb,), "lin_b", internal)]


def _lin(p, o, i):
    return [(p + "weight", (o, i), "lin"), (p + "bias", (o,), "lin_b", i)]


def _ln(p, c):
    return [(p + "weight", (c,), "ln_w"), (p + "bias", (c,), "ln_b")]


def _conv(p, o, i, k, groups=1):
    return [(p + "weight", (o, i // groups, k, k), "lin"), (p + "bias", (o,), "lin_b", (i // groups) * k * k)]


def weight_spec() -> List[tuple]:
    """(name, shape, kind[, fan_in]) for every tensor of the reference SAM2VideoPredictor
    state_dict built from sam2.1_hiera_l.yaml (903 tensors, 224.45 M parameters)."""
    s = [("maskmem_tpos_enc", (NUM_MASKMEM, 1, 1, MEM_DIM), "tn"),
         ("no_mem_embed", (1, 1, D_MODEL), "tn"), ("no_mem_pos_enc", (1, 1, D_MODEL), "tn"),
         ("no_obj_ptr", (1, D_MODEL), "tn"), ("no_obj_embed_spatial", (1, MEM_DIM), "tn")]
    t = "image_encoder.trunk."
    s += [(t + "pos_embed", (1, EMBED_DIM, 7, 7), "pos"), (t + "pos_embed_window", (1, EMBED_DIM, 8, 8), "pos")]
    s += _conv(t + "patch_embed.proj.", EMBED_DIM, 3, 7)
    for i, c in enumerate(block_cfgs()):
        b = f"{t}blocks.{i}."
        s += _ln(b + "norm1.", c["dim"])
        s += _lin(b + "attn.qkv.", 3 * c["dim_out"], c["dim"])
        s += _lin(b + "attn.proj.", c["dim_out"], c["dim_out"])
        s += _ln(b + "norm2.", c["dim_out"])
        s += _lin(b + "mlp.layers.0.", 4 * c["dim_out"], c["dim_out"])
        s += _lin(b + "mlp.layers.1.", c["dim_out"], 4 * c["dim_out"])
        if c["dim"] != c["dim_out"]:
            s += _lin(b + "proj.", c["dim_out"], c["dim"])
    for n, ch in enumerate((1152, 576, 288, 144)):
        s += _conv(f"image_encoder.neck.convs.{n}.conv.", D_MODEL, ch, 1)
    s += _conv("mask_downsample.", 1, 1, 4)
    for l in range(4):
        p = f"memory_attention.layers.{l}."
        s += _attn_spec(p + "self_attn.", 256, 256)
        s += _attn_spec(p + "cross_attn_image.", 256, 256, 64)
        s += _lin(p + "linear1.", 2048, 256) + _lin(p + "linear2.", 256, 2048)
        s += _ln(p + "norm1.", 256) + _ln(p + "norm2.", 256) + _ln(p + "norm3.", 256)
    s += _ln("memory_attention.norm.", 256)
    e = "memory_encoder.mask_downsampler.encoder."
    cin = 1
    for j in range(4):
        s += _conv(f"{e}{3 * j}.", cin * 4, cin, 3) + _ln(f"{e}{3 * j + 1}.", cin * 4)
        cin *= 4
    s += _conv(e + "12.", 256, 256, 1)
    s += _conv("memory_encoder.pix_feat_proj.", 256, 256, 1)
    for l in range(2):
        p = f"memory_encoder.fuser.layers.{l}."
        s += [(p + "gamma", (256,), "gamma")]
        s += _conv(p + "dwconv.", 256, 256, 7, groups=256) + _ln(p + "norm.", 256)
        s += _lin(p + "pwconv1.", 1024, 256) + _lin(p + "pwconv2.", 256, 1024)
    s += _conv("memory_encoder.out_proj.", 64, 256, 1)
    pe = "sam_prompt_encoder."
    s += [(pe + "pe_layer.positional_encoding_gaussian_matrix", (2, 128), "normal")]
    s += [(pe + f"point_embeddings.{i}.weight", (1, 256), "normal") for i in range(4)]
    s += [(pe + "not_a_point_embed.weight", (1, 256), "normal")]
    s += _conv(pe + "mask_downscaling.0.", 4, 1, 2) + _ln(pe + "mask_downscaling.1.", 4)
    s += _conv(pe + "mask_downscaling.3.", 16, 4, 2) + _ln(pe + "mask_downscaling.4.", 16)
    s += _conv(pe + "mask_downscaling.6.", 256, 16, 1)
    s += [(pe + "no_mask_embed.weight", (1, 256), "normal")]
    md = "sam_mask_decoder."
    for l in range(2):
        p = f"{md}transformer.layers.{l}."
        s += _attn_spec(p + "self_attn.", 256, 256) + _ln(p + "norm1.", 256)
        s += _attn_spec(p + "cross_attn_token_to_image.", 256, 128) + _ln(p + "norm2.", 256)
        s += _lin(p + "mlp.layers.0.", 2048, 256) + _lin(p + "mlp.layers.1.", 256, 2048)
        s += _ln(p + "norm3.", 256) + _ln(p + "norm4.", 256)
        s += _attn_spec(p + "cross_attn_image_to_token.", 256, 128)
    s += _attn_spec(md + "transformer.final_attn_token_to_image.", 256, 128)
    s += _ln(md + "transformer.norm_final_attn.", 256)
    s += [(md + "iou_token.weight", (1, 256), "normal"), (md + "mask_tokens.weight", (4, 256), "normal"),
          (md + "obj_score_token.weight", (1, 256), "normal")]
    # ConvTranspose2d weight is (in, out, k, k); torch's fan_in for it is out*k*k
    s += [(md + "output_upscaling.0.weight", (256, 64, 2, 2), "lin_f", 64 * 4),
          (md + "output_upscaling.0.bias", (64,), "lin_b", 64 * 4)]
    s += _ln(md + "output_upscaling.1.", 64)
    s += [(md + "output_upscaling.3.weight", (64, 32, 2, 2), "lin_f", 32 * 4),
          (md + "output_upscaling.3.bias", (32,), "lin_b", 32 * 4)]
    s += _conv(md + "conv_s0.", 32, 256, 1) + _conv(md + "conv_s1.", 64, 256, 1)
    for i in range(4):
        p = f"{md}output_hypernetworks_mlps.{i}.layers."
        s += _lin(p + "0.", 256, 256) + _lin(p + "1.", 256, 256) + _lin(p + "2.", 32, 256)
    p = md + "iou_prediction_head.layers."
    s += _lin(p + "0.", 256, 256) + _lin(p + "1.", 256, 256) + _lin(p + "2.", 4, 256)
    p = md + "pred_obj_score_head.layers."
    s += _lin(p + "0.", 256, 256) + _lin(p + "1.", 256, 256) + _lin(p + "2.", 1, 256)
    p = "obj_ptr_proj.layers."
    s += _lin(p + "0.", 256, 256) + _lin(p + "1.", 256, 256) + _lin(p + "2.", 256, 256)
    s += _lin("obj_ptr_tpos_proj.", 64, 256)
    return s


def make_weights(seed: int = 0, perturb: bool = True, only_prefix: Optional[Tuple[str, ...]] = None) -> W:
    """Deterministic random-init weights keyed by tensor NAME (independent of construction
    order, so the oracle, the product and the real reference get bit-identical values).

    Distributions follow torch's defaults (Linear/Conv: U(+-1/sqrt(fan_in)); Embedding: N(0,1);
    trunc_normal(0.02) params).  With ``perturb`` the tensors torch initialises to constants
    (LayerNorm affine, Hiera pos_embed*, CXBlock gamma) get non-trivial values so that a kernel
    which drops them cannot pass; ``perturb=False`` reproduces the torch-default constants."""
    out = {}
    for ent in weight_spec():
        name, shape, kind = ent[0], ent[1], ent[2]
        if only_prefix is not None and not name.startswith(only_prefix):
            continue
        g = torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
        if kind in ("lin", "lin_b", "lin_f"):
            fan_in = ent[3] if len(ent) > 3 else math.prod(shape[1:])
            bound = 1.0 / math.sqrt(fan_in)
            v = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif kind == "normal":
            v = torch.randn(shape, generator=g)
        elif kind == "tn":
            v = torch.randn(shape, generator=g).clamp_(-2, 2) * 0.02
        elif kind == "ln_w":
            v = 1 + 0.1 * torch.randn(shape, generator=g) if perturb else torch.ones(shape)
        elif kind == "ln_b":
            v = 0.05 * torch.randn(shape, generator=g) if perturb else torch.zeros(shape)
        elif kind == "pos":
            v = 0.02 * torch.randn(shape, generator=g) if perturb else torch.zeros(shape)
        elif kind == "gamma":
            v = 0.1 * (0.5 + torch.rand(shape, generator=g)) if perturb else torch.full(shape, 1e-6)
        else:
            raise KeyError(kind)
        out[name] = v.float().contiguous()
    return out
