"""CPU check of the host-side constants of the native SAM heads (weights.derive_heads): a torch restatement of the data flow of
engine.cu `sam_heads_mb` — split-bf16 operands (hi | lo | hi) x (hi | hi | lo), the dense-PE tables folded into the k / v / q
projections of the image side, ConvTranspose2d as a GEMM over (dy, dx, cout) followed by the per-position tail — must
reproduce the oracle's fp32 SAM heads (oracle/sam2_oracle.py, itself pinned against the unmodified reference by
tests/test_oracle_golden.py).  No GPU and no CUDA library involved: this pins the algebra and the layouts the kernels consume."""
import math

import torch
import torch.nn.functional as F

from oracle import sam2_oracle as O
from sam2_opt_b200 import synth_weights, weights


def _split_a(x):                      # activations: (hi | lo | hi)
    hi = x.to(torch.bfloat16).float()
    lo = (x - hi).to(torch.bfloat16).float()
    return torch.cat([hi, lo, hi], -1)


def _emulate(w, d, pix, s0, s1, coords, labels, multimask):
    n = pix.shape[0]
    md, tr, pe = "sam_mask_decoder.", "sam_mask_decoder.transformer.", "sam_prompt_encoder."
    P = coords.shape[1]
    # tokens
    pts = torch.cat([coords + 0.5, torch.zeros(n, 1, 2)], 1) / 1024
    lab = torch.cat([labels, -torch.ones(n, 1, dtype=labels.dtype)], 1)
    c = 2 * math.pi * ((2 * pts - 1) @ w[pe + "pe_layer.positional_encoding_gaussian_matrix"])
    emb = torch.cat([c.sin(), c.cos()], -1)
    emb = torch.where((lab == -1)[..., None], w[pe + "not_a_point_embed.weight"].expand_as(emb), emb)
    for i in range(4):
        emb = torch.where((lab == i)[..., None], emb + w[pe + f"point_embeddings.{i}.weight"], emb)
    out_tok = torch.cat([w[md + "obj_score_token.weight"], w[md + "iou_token.weight"], w[md + "mask_tokens.weight"]], 0)
    tok = torch.cat([out_tok[None].expand(n, -1, -1), emb], 1)
    qpe = tok.clone()
    T = tok.shape[1]
    assert T == 6 + P + 1
    keys = pix.flatten(2).permute(0, 2, 1) + w[pe + "no_mask_embed.weight"].view(1, 1, -1)       # [n,4096,256]
    lin = lambda x, p: F.linear(x, w[p + "weight"], w[p + "bias"])
    ln = lambda x, p: F.layer_norm(x, (256,), w[p + "weight"], w[p + "bias"], 1e-5)

    def attn(q, k, v, heads):
        sh = lambda x: x.reshape(x.shape[0], x.shape[1], heads, -1).transpose(1, 2)
        qh, kh, vh = sh(q), sh(k), sh(v)
        o = torch.softmax(qh @ kh.transpose(-1, -2) / math.sqrt(qh.shape[-1]), -1) @ vh
        return o.transpose(1, 2).reshape(q.shape[0], q.shape[1], -1)

    def t2i(p, kv, norm, tok):
        q = lin(tok + qpe, p + "q_proj.")
        o = attn(q, kv[..., :128], kv[..., 128:256], 8)
        return ln(tok + lin(o, p + "out_proj."), norm)

    for l in range(2):
        p, dd = f"{tr}layers.{l}.", f"derived.heads.{l}."
        sa = p + "self_attn."
        x = tok + qpe if l else tok
        o = lin(attn(lin(x, sa + "q_proj."), lin(x, sa + "k_proj."), lin(tok, sa + "v_proj."), 8), sa + "out_proj.")
        tok = ln(tok + o if l else o, p + "norm1.")
        kvq = _split_a(keys) @ d[dd + "kvq.weight"].t() + d[dd + "kvq.table"][None]
        tok = t2i(p + "cross_attn_token_to_image.", kvq, p + "norm2.", tok)
        tok = ln(tok + lin(F.relu(lin(tok, p + "mlp.layers.0.")), p + "mlp.layers.1."), p + "norm3.")
        ca = p + "cross_attn_image_to_token."
        o = attn(kvq[..., 256:], lin(tok + qpe, ca + "k_proj."), lin(tok, ca + "v_proj."), 8)
        keys = ln(keys + _split_a(o) @ d[dd + "i2t_out.weight"].t() + w[ca + "out_proj.bias"], p + "norm4.")
    kv = _split_a(keys) @ d["derived.heads.final.kv.weight"].t() + d["derived.heads.final.kv.table"][None]
    hs = t2i(tr + "final_attn_token_to_image.", kv, tr + "norm_final_attn.", tok)

    def mlp3(x, p, sigmoid=False):
        for i in range(3):
            x = F.linear(x, w[f"{p}layers.{i}.weight"], w[f"{p}layers.{i}.bias"])
            if i < 2:
                x = F.relu(x)
        return torch.sigmoid(x) if sigmoid else x

    iou = mlp3(hs[:, 1], md + "iou_prediction_head.", True)
    obj = mlp3(hs[:, 0], md + "pred_obj_score_head.")
    hyper = torch.stack([mlp3(hs[:, 2 + i], f"{md}output_hypernetworks_mlps.{i}.") for i in range(4)], 1)       # [n,4,32]
    ptr4 = torch.stack([mlp3(hs[:, 2 + i], "obj_ptr_proj.") for i in range(4)], 1)
    u1 = (_split_a(keys) @ d["derived.heads.up1.weight"].t()).reshape(n, 64, 64, 2, 2, 64)                       # [n,y,x,dy,dx,c]
    x = u1.permute(0, 1, 3, 2, 4, 5).reshape(n, 128, 128, 64) + w[md + "output_upscaling.0.bias"] + s1.permute(0, 2, 3, 1)
    x = F.gelu(F.layer_norm(x, (64,), w[md + "output_upscaling.1.weight"], w[md + "output_upscaling.1.bias"], 1e-6))
    w2 = d["derived.heads.up2.weight"]                                                                          # [64,4,32]
    y = torch.einsum("nyxc,cqo->nyxqo", x, w2).reshape(n, 128, 128, 2, 2, 32).permute(0, 1, 3, 2, 4, 5).reshape(n, 256, 256, 32)
    y = F.gelu(y + w[md + "output_upscaling.3.bias"] + s0.permute(0, 2, 3, 1))
    masks = torch.einsum("nko,nyxo->nkyx", hyper, y)
    best = 1 + torch.argmax(iou[:, 1:], -1)
    bi = torch.arange(n)
    is_obj = obj[:, 0] > 0
    if multimask:
        low_multi, ious, tok_idx, mask_idx = masks[:, 1:], iou[:, 1:], best, best
    else:
        fl = masks[:, 0].flatten(1)
        ai, au = (fl > 0.05).sum(-1).float(), (fl > -0.05).sum(-1).float()
        stable = torch.where(au > 0, ai / au, torch.ones_like(au)) >= 0.98
        mask_idx = torch.where(stable, torch.zeros_like(best), best)
        low_multi, ious, tok_idx = masks[bi, mask_idx][:, None], iou[bi, mask_idx][:, None], torch.zeros_like(best)
    low_multi = torch.where(is_obj[:, None, None, None], low_multi, torch.full_like(low_multi, -1024.0))
    low_best = torch.where(is_obj[:, None, None], masks[bi, mask_idx], torch.full_like(masks[:, 0], -1024.0))[:, None]
    lam = is_obj.float()[:, None]
    ptr = lam * ptr4[bi, tok_idx] + (1 - lam) * w["no_obj_ptr"]
    return low_multi, ious, low_best, ptr, obj


def _case(seed, n, P, multimask):
    g = torch.Generator().manual_seed(seed)
    w = synth_weights.make_weights(0, only_prefix=("sam_", "obj_ptr", "no_obj"))
    d = weights.derive_heads(w)
    pix = torch.randn(n, 256, 64, 64, generator=g)
    s0 = 0.5 * torch.randn(n, 32, 256, 256, generator=g)
    s1 = 0.5 * torch.randn(n, 64, 128, 128, generator=g)
    coords = torch.rand(n, P, 2, generator=g) * 1024
    labels = torch.randint(-1, 4, (n, P), generator=g, dtype=torch.int32)
    got = _emulate(w, d, pix, s0, s1, coords, labels, multimask)
    low, _, ious, low_best, _, ptr, obj = O.forward_sam_heads(w, pix, point_inputs={"point_coords": coords, "point_labels": labels},
                                                              high_res_features=[s0, s1], multimask_output=multimask)
    return got, (low, ious, low_best, ptr, obj)


def test_heads_data_flow_matches_oracle():
    for seed, n, P, mm in ((0, 2, 1, True), (1, 1, 3, False)):
        got, want = _case(seed, n, P, mm)
        for name, a, b in zip(("low_res_multimasks", "ious", "low_res_masks", "obj_ptr", "object_score_logits"), got, want):
            assert a.shape == b.shape, (name, a.shape, b.shape)
            err = (a - b).norm() / b.norm().clamp_min(1e-12)
            assert err < 2e-4, (name, float(err))
