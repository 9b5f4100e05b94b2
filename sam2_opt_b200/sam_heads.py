"""Host-side SAM heads (prompt encoder, two-way mask decoder, object-pointer head) and memory-bank assembly, in plain
fp32 torch on the engine's device.

These are the callers either side of the accelerated path: in the reference they stay torch
modules when ``speedup()`` is engaged, and they must stay fp32 for the mask-IoU gate.  The functional
form over the reference's own ``state_dict`` names lets the standalone predictors in ``predictor.py`` run without the
reference package.  All ``ref:`` citations are relative to /root/reference/sam2/sam2/.
"""
import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

W = Dict[str, torch.Tensor]
IMAGE_SIZE = 1024
D_MODEL = 256
MEM_DIM = 64
NUM_MASKMEM = 7
NO_OBJ_SCORE = -1024.0


def layer_norm_2d(x, w, b, eps=1e-6):
    """ref: modeling/sam2_utils.py:141-153"""
    u = x.mean(1, keepdim=True)
    s = (x - u).pow(2).mean(1, keepdim=True)
    return w[:, None, None] * ((x - u) / torch.sqrt(s + eps)) + b[:, None, None]


def mlp(w: W, p: str, x, n: int, sigmoid=False):
    """ref: modeling/sam2_utils.py:112-136"""
    for i in range(n):
        x = F.linear(x, w[f"{p}layers.{i}.weight"], w[f"{p}layers.{i}.bias"])
        if i < n - 1:
            x = F.relu(x)
    return torch.sigmoid(x) if sigmoid else x


def pe_encoding(w: W, coords01):
    """ref: modeling/position_encoding.py:146-154 (PositionEmbeddingRandom._pe_encoding)"""
    c = (2 * coords01 - 1) @ w["sam_prompt_encoder.pe_layer.positional_encoding_gaussian_matrix"]
    c = 2 * math.pi * c
    return torch.cat([torch.sin(c), torch.cos(c)], dim=-1)


def dense_pe(w: W, size=64):
    """ref: modeling/position_encoding.py:156-167 via prompt_encoder.py:113-122 -> [1,256,size,size]"""
    dev = w["sam_prompt_encoder.no_mask_embed.weight"].device
    g = torch.ones((size, size), dtype=torch.float32, device=dev)
    y = (g.cumsum(0) - 0.5) / size
    x = (g.cumsum(1) - 0.5) / size
    return pe_encoding(w, torch.stack([x, y], dim=-1)).permute(2, 0, 1)[None]


def prompt_encoder(w: W, coords, labels, masks, bs=1):
    """ref: modeling/sam/prompt_encoder.py:124-231 (points, boxes passed as points with labels 2/3, optional mask)"""
    pe = "sam_prompt_encoder."
    dev = w[pe + "no_mask_embed.weight"].device
    sparse = torch.empty((bs, 0, D_MODEL), device=dev)
    if coords is not None:
        bs = coords.shape[0]
        pts = torch.cat([coords + 0.5, torch.zeros((bs, 1, 2), device=dev)], 1)
        lab = torch.cat([labels, -torch.ones((bs, 1), dtype=labels.dtype, device=dev)], 1)
        emb = pe_encoding(w, pts / IMAGE_SIZE)
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
    """ref: modeling/sam/transformer.py:222-294"""
    q = F.linear(q, w[p + "q_proj.weight"], w[p + "q_proj.bias"])
    k = F.linear(k, w[p + "k_proj.weight"], w[p + "k_proj.bias"])
    v = F.linear(v, w[p + "v_proj.weight"], w[p + "v_proj.bias"])
    sh = lambda x: x.reshape(x.shape[0], x.shape[1], heads, -1).transpose(1, 2)
    # explicit fp32 softmax(q k^T / sqrt(d)) v: the token sets are tiny (<= 9 tokens against 4096 pixels), for which three
    # plain kernels take ~30 us where the fused mem-efficient SDPA kernel takes ~150 us per call (7 calls per frame)
    qh, kh, vh = sh(q), sh(k), sh(v)
    o = (torch.softmax((qh * (qh.shape[-1] ** -0.5)) @ kh.transpose(-1, -2), dim=-1) @ vh).transpose(1, 2)
    return F.linear(o.reshape(o.shape[0], o.shape[1], -1), w[p + "out_proj.weight"], w[p + "out_proj.bias"])


def two_way_transformer(w: W, src, pos, tokens):
    """ref: modeling/sam/transformer.py:97-219 (depth 2, 8 heads, attention downsample 2)"""
    t = "sam_mask_decoder.transformer."
    keys, kpe = src.flatten(2).permute(0, 2, 1), pos.flatten(2).permute(0, 2, 1)
    queries, qpe = tokens, tokens
    ln = lambda x, n: F.layer_norm(x, (x.shape[-1],), w[n + "weight"], w[n + "bias"], 1e-5)
    for l in range(2):
        p = f"{t}layers.{l}."
        if l == 0:
            queries = _attention(w, p + "self_attn.", queries, queries, queries)
        else:
            q = queries + qpe
            queries = queries + _attention(w, p + "self_attn.", q, q, queries)
        queries = ln(queries, p + "norm1.")
        queries = ln(queries + _attention(w, p + "cross_attn_token_to_image.", queries + qpe, keys + kpe, keys), p + "norm2.")
        queries = ln(queries + mlp(w, p + "mlp.", queries, 2), p + "norm3.")
        keys = ln(keys + _attention(w, p + "cross_attn_image_to_token.", keys + kpe, queries + qpe, queries), p + "norm4.")
    queries = queries + _attention(w, t + "final_attn_token_to_image.", queries + qpe, keys + kpe, keys)
    return ln(queries, t + "norm_final_attn."), keys


def _dynamic_multimask(all_logits, all_iou, delta=0.05, thresh=0.98):
    """ref: modeling/sam/mask_decoder.py:337-382"""
    mm, mi = all_logits[:, 1:], all_iou[:, 1:]
    best = torch.argmax(mi, dim=-1)
    bi = torch.arange(mi.shape[0], device=mi.device)
    best_logits, best_iou = mm[bi, best][:, None], mi[bi, best][:, None]
    single, single_iou = all_logits[:, 0:1], all_iou[:, 0:1]
    fl = single.flatten(-2)
    ai, au = (fl > delta).sum(-1).float(), (fl > -delta).sum(-1).float()
    stable = torch.where(au > 0, ai / au, 1.0) >= thresh
    return (torch.where(stable[..., None, None].expand_as(single), single, best_logits),
            torch.where(stable.expand_as(single_iou), single_iou, best_iou))


def mask_decoder(w: W, image_emb, sparse, dense, multimask: bool, high_res: List[torch.Tensor], repeat_image=False,
                 dynamic_multimask=True, pe=None):
    """ref: modeling/sam/mask_decoder.py:116-224, 262-316"""
    md = "sam_mask_decoder."
    out_tok = torch.cat([w[md + "obj_score_token.weight"], w[md + "iou_token.weight"], w[md + "mask_tokens.weight"]], 0)
    tokens = torch.cat((out_tok[None].expand(sparse.shape[0], -1, -1), sparse), dim=1)
    src = torch.repeat_interleave(image_emb, tokens.shape[0], dim=0) if repeat_image else image_emb
    src = src + dense
    pe = dense_pe(w) if pe is None else pe
    pos = pe.expand(tokens.shape[0], -1, -1, -1)
    b, c, h, wd = src.shape
    hs, src = two_way_transformer(w, src, pos, tokens)
    iou_tok, mask_toks = hs[:, 1], hs[:, 2:6]
    src = src.transpose(1, 2).reshape(b, c, h, wd)
    u = md + "output_upscaling."
    x = F.conv_transpose2d(src, w[u + "0.weight"], w[u + "0.bias"], stride=2) + high_res[1]
    x = F.gelu(layer_norm_2d(x, w[u + "1.weight"], w[u + "1.bias"]))
    x = F.gelu(F.conv_transpose2d(x, w[u + "3.weight"], w[u + "3.bias"], stride=2) + high_res[0])
    hyper = torch.stack([mlp(w, f"{md}output_hypernetworks_mlps.{i}.", mask_toks[:, i], 3) for i in range(4)], 1)
    b, c, h, wd = x.shape
    masks = (hyper @ x.reshape(b, c, h * wd)).reshape(b, -1, h, wd)
    iou = mlp(w, md + "iou_prediction_head.", iou_tok, 3, sigmoid=True)
    obj = mlp(w, md + "pred_obj_score_head.", hs[:, 0], 3)
    if multimask:
        masks, iou, sam_tok = masks[:, 1:], iou[:, 1:], mask_toks[:, 1:]
    else:
        sam_tok = mask_toks[:, 0:1]
        if dynamic_multimask:
            masks, iou = _dynamic_multimask(masks, iou)
        else:
            masks, iou = masks[:, 0:1], iou[:, 0:1]
    return masks, iou, sam_tok, obj


def mask_prompt_256(mask_inputs):
    """ref: modeling/sam2_base_official.py:397-411 - a mask prompt of another size is resized to the prompt encoder's 256 x 256"""
    if mask_inputs.shape[-2:] != (256, 256):
        return F.interpolate(mask_inputs.float(), size=(256, 256), align_corners=False, mode="bilinear", antialias=True)
    return mask_inputs


def forward_sam_heads(w: W, backbone_features, point_inputs=None, mask_inputs=None, high_res_features=None,
                      multimask_output=False, pe=None):
    """ref: modeling/sam2_base_official.py:338-494"""
    B, dev = backbone_features.shape[0], backbone_features.device
    if point_inputs is not None:
        coords, labels = point_inputs["point_coords"], point_inputs["point_labels"]
    else:
        coords = torch.zeros(B, 1, 2, device=dev)
        labels = -torch.ones(B, 1, dtype=torch.int32, device=dev)
    mask_prompt = None
    if mask_inputs is not None:
        mask_prompt = mask_inputs
        if mask_inputs.shape[-2:] != (256, 256):
            mask_prompt = F.interpolate(mask_inputs.float(), size=(256, 256), align_corners=False, mode="bilinear", antialias=True)
    sparse, dense = prompt_encoder(w, coords, labels, mask_prompt)
    low, ious, sam_tokens, obj_logits = mask_decoder(w, backbone_features, sparse, dense, multimask_output, high_res_features, pe=pe)
    is_obj = obj_logits > 0
    low = torch.where(is_obj[:, None, None], low, torch.full_like(low, NO_OBJ_SCORE)).float()
    high = F.interpolate(low, size=(IMAGE_SIZE, IMAGE_SIZE), mode="bilinear", align_corners=False)
    tok = sam_tokens[:, 0]
    if multimask_output:
        best = torch.argmax(ious, dim=-1)
        bi = torch.arange(B, device=dev)
        low_best, high_best = low[bi, best][:, None], high[bi, best][:, None]
        if sam_tokens.shape[1] > 1:
            tok = sam_tokens[bi, best]
    else:
        low_best, high_best = low, high
    ptr = mlp(w, "obj_ptr_proj.", tok, 3)
    lam = is_obj.float()
    ptr = lam * ptr + (1 - lam) * w["no_obj_ptr"]
    return low, high, ious, low_best, high_best, ptr, obj_logits


def get_1d_sine_pe(pos, dim, temperature=10000.0):
    """ref: modeling/sam2_utils.py:64-74"""
    pe_dim = dim // 2
    dim_t = torch.arange(pe_dim, dtype=torch.float32, device=pos.device)
    dim_t = temperature ** (2 * torch.div(dim_t, 2, rounding_mode="floor") / pe_dim)
    e = pos.unsqueeze(-1) / dim_t
    return torch.cat([e.sin(), e.cos()], dim=-1)


def _ptr_tpos_table(w: W, num_frames: int, max_ptrs: int, cache: dict):
    """Projected temporal encodings of every integer frame distance 0..num_frames, [num_frames + 1, 64], built once per
    video length: the per-call `torch.tensor(pos_list, device=...)` of the reference (sam2_base_official.py:930-937) is a
    pageable host-to-device copy, i.e. a stream synchronisation per object per frame."""
    key = (num_frames, max_ptrs)
    if key not in cache:
        d = torch.arange(num_frames + 1, dtype=torch.float32, device=w["obj_ptr_tpos_proj.weight"].device)
        pe = get_1d_sine_pe(d / (max_ptrs - 1), D_MODEL)
        cache[key] = F.linear(pe, w["obj_ptr_tpos_proj.weight"], w["obj_ptr_tpos_proj.bias"])
    return cache[key]


def assemble_pointers(w: W, frame_idx: int, cond_outputs: dict, non_cond_outputs: dict, num_frames: int, cache: dict = None):
    """Object-pointer tokens of the memory bank (ref: modeling/sam2_base_official.py:886-950): the pointers of the conditioning
    frames up to `frame_idx` and of up to 15 previous tracked frames, each split into 4 tokens of 64, with the projected sine
    encoding of its temporal distance -> (tokens [P,n,64], positions [P,n,64])."""
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
    ptrs = torch.stack(ptrs, 0)                                                   # [np,n,256]
    dev, n = ptrs.device, ptrs.shape[1]
    if cache is not None and max_ptrs > 1:
        table = _ptr_tpos_table(w, num_frames, max_ptrs, cache)
        opos = torch.stack([table[d] for d in pos_list], 0)
    else:
        opos = get_1d_sine_pe(torch.tensor(pos_list, dtype=torch.float32, device=dev) / (max_ptrs - 1), D_MODEL)
        opos = F.linear(opos, w["obj_ptr_tpos_proj.weight"], w["obj_ptr_tpos_proj.bias"])
    opos = opos.unsqueeze(1).expand(-1, n, MEM_DIM)
    ptrs = ptrs.reshape(-1, n, D_MODEL // MEM_DIM, MEM_DIM).permute(0, 2, 1, 3).flatten(0, 1)
    opos = opos.repeat_interleave(D_MODEL // MEM_DIM, dim=0)
    return ptrs, opos


def assemble_memory(w: W, frame_idx: int, cond_outputs: dict, non_cond_outputs: dict, num_frames: int, cache: dict = None):
    """Memory bank + object-pointer tokens for `frame_idx` (forward tracking, stride 1) of one object or of n objects that
    share the same conditioning / tracked frames.  ref: modeling/sam2_base_official.py:825-950.  Stored outputs hold
    'maskmem_features' [n,64,64,64] (bf16, as the reference stores them, sam2_video_predictor_official.py:887),
    'maskmem_pos' [n,64,64,64], 'obj_ptr' [n,256].
    -> memory [L,4096,n,64], memory_pos [L,4096,n,64], ptr tokens [P,n,64], ptr pos [P,n,64]"""
    t_pos_prev = [(0, o) for o in cond_outputs.values()]
    for t_pos in range(1, NUM_MASKMEM):
        t_pos_prev.append((t_pos, non_cond_outputs.get(frame_idx - (NUM_MASKMEM - t_pos))))
    mem, mpos = [], []
    for t_pos, prev in t_pos_prev:
        if prev is None:
            continue
        mem.append(prev["maskmem_features"].float().flatten(2).permute(2, 0, 1))
        mpos.append(prev["maskmem_pos"].flatten(2).permute(2, 0, 1) + w["maskmem_tpos_enc"][NUM_MASKMEM - t_pos - 1])
    ptrs, opos = assemble_pointers(w, frame_idx, cond_outputs, non_cond_outputs, num_frames, cache)
    return torch.stack(mem, 0), torch.stack(mpos, 0), ptrs, opos
