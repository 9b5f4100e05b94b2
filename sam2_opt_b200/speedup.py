"""``predictor.speedup("b200")`` for the REFERENCE's own predictor objects (drop-in).

The reference switches backends by rebinding method pointers and storing executors in ``backend_contexts``
(ref: sam2/sam2/sam2_image_predictor.py:94-138,204-250; sam2/sam2/sam2_video_predictor_official.py:45-145;
sam2/sam2/modeling/sam2_base_official.py:230-276; modeling/memory_attention.py:134-261; modeling/memory_encoder.py:160-215).
``speedup(predictor, "b200")`` does exactly that with ``B200Executor`` objects over one shared engine, so
``SAM2ImagePredictor`` / ``SAM2VideoPredictor`` keep their API.  ``install()`` additionally teaches the reference classes'
own ``speedup`` method the new backend name, making ``predictor.speedup("b200")`` work verbatim.

The reference tree is read-only, so this is duck-typed: anything with the same attributes works (the CPU tests drive the
real reference predictors with a stub engine; ``tests/test_speedup_reference_gpu.py`` drives them with the real engine
on the GPU box, where the unmodified reference sources travel as the git-ignored ``oracle/_ref`` copy).
"""
import torch

from .executor import B200Executor

B200_NAMES = ("b200", "sm100", "sm_100a", "cuda")


def _release_contexts(obj):
    for ctx in (getattr(obj, "backend_contexts", None) or []):
        ctx.Release()
    obj.backend_contexts = []


def _make_engine(model, device, engine_factory):
    if engine_factory is None:
        from .engine import Engine
        engine_factory = Engine
    return engine_factory(model.state_dict(), device)


def _native_forward_sam_heads(model, engine):
    """SAM2Base._forward_sam_heads (ref: modeling/sam2_base_official.py:338-494) on the engine: same arguments, same 7 outputs.
    The mask prompt's down-scaling convolutions (sam/prompt_encoder.py:152-155) are one library kernel (sam2b_mask_downscale) whose
    result enters sam2b_sam_heads as the dense embedding; everything else is that one call."""
    import torch.nn.functional as F

    def forward(backbone_features, point_inputs=None, mask_inputs=None, high_res_features=None, multimask_output=False):
        if high_res_features is None or (point_inputs is not None and point_inputs["point_coords"].shape[1] > 25):
            return type(model)._forward_sam_heads(model, backbone_features, point_inputs, mask_inputs, high_res_features, multimask_output)
        coords = labels = dense = None
        if point_inputs is not None:
            coords, labels = point_inputs["point_coords"], point_inputs["point_labels"]
        if mask_inputs is not None:
            size = tuple(model.sam_prompt_encoder.mask_input_size)
            m = mask_inputs.float()
            if tuple(m.shape[-2:]) != size:
                m = F.interpolate(m, size=size, align_corners=False, mode="bilinear", antialias=True)
            dense = engine.mask_downscale(m) if tuple(size) == (256, 256) else model.sam_prompt_encoder._embed_masks(m)
        dev = backbone_features.device
        low_multi, high_multi, ious, low, high, ptr, obj = engine.sam_heads(
            backbone_features.float(), [f.float() for f in high_res_features], coords, labels, dense,
            multimask_output=multimask_output, high_res_multimasks=True)
        return tuple(t.to(dev) for t in (low_multi, high_multi, ious, low, high, ptr, obj))

    return forward


def _native_mask_decoder_forward(model, engine):
    """MaskDecoder.forward (ref: modeling/sam/mask_decoder.py:116-175) on the engine, for the callers that drive the decoder
    themselves: SAM2ImagePredictor._predict (sam2_image_predictor.py:548-570) and, through it, SAM2AutomaticMaskGenerator with
    its batches of 64 one-click prompts on one image.  `image_pe` is not read: every caller in the reference passes the prompt
    encoder's own dense positional encoding, which the library carries folded into its projection tables."""
    md = model.sam_mask_decoder
    no_mask = model.sam_prompt_encoder.no_mask_embed.weight

    def forward(image_embeddings, image_pe, sparse_prompt_embeddings, dense_prompt_embeddings, multimask_output, repeat_image,
                high_res_features=None):
        n = sparse_prompt_embeddings.shape[0]
        if high_res_features is None or sparse_prompt_embeddings.shape[1] > 26 or (not repeat_image and image_embeddings.shape[0] != n):
            return type(md).forward(md, image_embeddings, image_pe, sparse_prompt_embeddings, dense_prompt_embeddings,
                                    multimask_output, repeat_image, high_res_features)
        dev = image_embeddings.device
        dense = dense_prompt_embeddings
        if dense.data_ptr() == no_mask.data_ptr() and dense.stride(2) == 0 and dense.stride(3) == 0:
            dense = None                                   # the expanded no_mask_embed: the library adds it itself
        else:
            dense = dense.expand(n, -1, -1, -1)
        emb = image_embeddings.float().expand(n, -1, -1, -1)
        low, _, iou, _, _, _, obj, toks = engine.sam_heads(emb, [f.float() for f in high_res_features], None, None, dense,
                                                           multimask_output=multimask_output, decoder_only=True, high_res=False,
                                                           sparse_embeddings=sparse_prompt_embeddings.float(), mask_tokens=True)
        sam_tokens = toks[:, 1:] if (multimask_output and md.use_multimask_token_for_obj_ptr) else toks[:, 0:1]
        return low.to(dev), iou.to(dev), sam_tokens.to(dev), obj.to(dev)

    return forward


def _encode_ahead_image_feature(model, encode_ahead):
    """SAM2VideoPredictor._get_image_feature (ref: sam2_video_predictor_official.py:810-841) encodes ONE frame per call and
    keeps a one-entry cache: at batch 1 the encoder runs at 224 images/s against 360 at batch 16 (profiles/r2_config5_sweep_n1).
    This wrapper fills that cache from a block of `encode_ahead` frames encoded in ONE engine call, in the direction the
    caller is walking through the video, and then runs the class's own method — which finds its frame cached.  Per-frame
    results do not change: the encoder is batch-invariant bit for bit (tested)."""
    cls_fn = type(model)._get_image_feature

    def get(inference_state, frame_idx, batch_size):
        images = inference_state.get("images")
        if frame_idx not in inference_state["cached_features"] and torch.is_tensor(images):
            ahead = inference_state.setdefault("_b200_ahead", {})
            if frame_idx not in ahead:
                ahead.clear()
                back = inference_state.get("_b200_last") == frame_idx + 1            # tracking in reverse
                lo = max(0, frame_idx - encode_ahead + 1) if back else frame_idx
                hi = frame_idx + 1 if back else min(images.shape[0], frame_idx + encode_ahead)
                batch = images[lo:hi].to(inference_state["device"]).float()
                out = model.forward_image(batch)
                for i in range(hi - lo):
                    ahead[lo + i] = (batch[i:i + 1], {"vision_features": out["vision_features"][i:i + 1],
                                                      "vision_pos_enc": [p[i:i + 1] for p in out["vision_pos_enc"]],
                                                      "backbone_fpn": [f[i:i + 1] for f in out["backbone_fpn"]]})
            inference_state["cached_features"] = {frame_idx: ahead[frame_idx]}
        inference_state["_b200_last"] = frame_idx
        return cls_fn(model, inference_state, frame_idx, batch_size)

    return get


def _heads_config_supported(model) -> bool:
    """The library implements the SAM2.1 head configuration (configs/sam2.1/sam2.1_hiera_*.yaml:98-117): high-resolution
    features, object scores from an MLP, fixed (hard) no-object pointer, the multimask token for the pointer, dynamic
    multimask fall-back with the default thresholds, sigmoid IoU head.  Any other configuration keeps the torch heads."""
    md = getattr(model, "sam_mask_decoder", None)
    if md is None:
        return False
    # the fork's DAM4SAM track (SAM2_VERSION_TRACK=dam4sam, modeling/sam2_base_dam4sam.py:338-503) returns an eighth value from
    # `_forward_sam_heads` (all candidate masks and IoUs as numpy, for its distractor memory) and overwrites the multimask
    # outputs with the chosen mask: it keeps its own torch heads, E2 / E3 / E4 are rebound as usual
    if "dam4sam" in (getattr(getattr(type(model), "_forward_sam_heads", None), "__module__", "") or ""):
        return False
    want_model = dict(use_high_res_features_in_sam=True, pred_obj_scores=True, fixed_no_obj_ptr=True, soft_no_obj_ptr=False,
                      use_obj_ptrs_in_encoder=True)
    want_md = dict(use_multimask_token_for_obj_ptr=True, dynamic_multimask_via_stability=True, pred_obj_scores=True,
                   use_high_res_features=True, dynamic_multimask_stability_delta=0.05, dynamic_multimask_stability_thresh=0.98)
    return all(getattr(model, k, None) == v for k, v in want_model.items()) and \
        all(getattr(md, k, None) == v for k, v in want_md.items()) and not getattr(model, "training", False)


def speedup(predictor, backend="b200", engine=None, engine_factory=None, warmup=True, heads=True, encode_ahead=16):
    """Engage (``backend="b200"``) or disengage (``"torch"``) the sm_100a engine on a reference predictor.

    Image predictor (has ``set_image_e2e``): rebinds E1.  Video predictor / SAM2Base (has ``inference_image``): rebinds
    E2, E3 (both ``none`` and ``exclude`` variants) and E4 and, with ``heads``, the model's
    ``_forward_sam_heads`` (prompt encoder + mask decoder + pointer head); ``encode_ahead`` > 1 makes the video predictor encode
    its frames in blocks of that many (see _encode_ahead_image_feature).  Returns the engine in use (None for "torch")."""
    b = backend.lower()
    is_image = hasattr(predictor, "set_image_e2e") and hasattr(predictor, "model")
    model = predictor.model if is_image else predictor
    if b == "torch":
        predictor.set_runtime_backend(backend="torch")
        if not is_image:
            model.memory_attention.set_runtime_backend(backend="torch")
            model.memory_encoder.set_runtime_backend(backend="torch")
            model.__dict__.pop("_forward_sam_heads", None)          # back to the class's own torch heads
            model.__dict__.pop("_get_image_feature", None)
        if getattr(model, "sam_mask_decoder", None) is not None:
            model.sam_mask_decoder.__dict__.pop("forward", None)
        return None
    if b not in B200_NAMES:
        raise Exception(f"unsupported runtime backend={backend}")
    device = next(model.parameters()).device if hasattr(model, "parameters") else torch.device("cuda")
    if engine is None:
        engine = _make_engine(model, device, engine_factory)

    if is_image:
        _release_contexts(predictor)
        ex = B200Executor(engine, "set_image_e2e")
        predictor.backend_contexts.append(ex)
        predictor.set_image_e2e = lambda img, _ex=ex: tuple(o.to(img.device) for o in _ex.Inference([img], output_type="torch"))
        if heads and hasattr(engine, "sam_heads") and _heads_config_supported(model):
            model.sam_mask_decoder.forward = _native_mask_decoder_forward(model, engine)
        if warmup:
            ex.warmup([torch.randn(1, 3, 1024, 1024)])
        return engine

    # ---- SAM2Base / SAM2VideoPredictor ----
    _release_contexts(model)
    ex_img = B200Executor(engine, "inference_image")
    model.backend_contexts.append(ex_img)
    model.inference_image = lambda img, _ex=ex_img: tuple(o.to(img.device) for o in _ex.Inference([img], output_type="torch"))
    ma = model.memory_attention
    _release_contexts(ma)
    ma.backend_contexts = [B200Executor(engine, "memory_attention_none"), B200Executor(engine, "memory_attention_exclude")]
    ma.inference_memory_attention_none = ma.inference_memory_attention_speedup_none
    ma.inference_memory_attention_exclude = ma.inference_memory_attention_speedup_exclude
    me = model.memory_encoder
    _release_contexts(me)
    ex_me = B200Executor(engine, "memory_encoder")
    me.backend_contexts = [ex_me]

    def inference_memory(pix_feat, masks, _ex=ex_me):
        x, pos = _ex.Inference([pix_feat, masks], output_type="torch")
        return x.to(pix_feat.device), pos.to(pix_feat.device)

    me.inference_memory = inference_memory
    if heads and hasattr(engine, "sam_heads") and _heads_config_supported(model):
        model._forward_sam_heads = _native_forward_sam_heads(model, engine)
        model.sam_mask_decoder.forward = _native_mask_decoder_forward(model, engine)
    if encode_ahead and encode_ahead > 1 and hasattr(type(model), "_get_image_feature"):
        model._get_image_feature = _encode_ahead_image_feature(model, int(encode_ahead))
    if warmup:
        ex_img.warmup([torch.randn(1, 3, 1024, 1024)])
        for L in (1, 7):
            base = [torch.randn(4096, 1, 256), torch.randn(L, 4096, 1, 64), torch.randn(4096, 1, 256), torch.randn(L, 4096, 1, 64)]
            ma.backend_contexts[0].warmup(base)
            ma.backend_contexts[1].warmup(base + [torch.randn(64, 1, 64), torch.randn(64, 1, 64)])
        ex_me.warmup([torch.randn(1, 256, 64, 64), torch.rand(1, 1, 1024, 1024)])
    return engine


def install(*classes):
    """Patch reference predictor classes so that ``obj.speedup("b200")`` routes here and every other backend name keeps
    the reference behaviour.  Call as ``install(SAM2ImagePredictor, SAM2VideoPredictor)``."""
    for cls in classes:
        orig = cls.speedup
        if getattr(orig, "_sam2b_patched", False):
            continue

        def patched(self, backend="b200", *a, _orig=orig, **k):
            if isinstance(backend, str) and backend.lower() in B200_NAMES:
                self._sam2b_engine = speedup(self, backend, **{x: k[x] for x in ("engine", "engine_factory", "warmup", "heads", "encode_ahead") if x in k})
                return None
            if isinstance(backend, str) and backend.lower() == "torch" and getattr(self, "_sam2b_engine", None) is not None:
                # the reference's own speedup("torch") resets the image / memory-attention hooks only
                # (sam2_video_predictor_official.py:58-61); E4 was engaged by us, so it is disengaged here as well
                speedup(self, "torch")
                return None
            return _orig(self, backend, *a, **k)

        patched._sam2b_patched = True
        cls.speedup = patched


def install_postprocess(misc_module):
    """Give the reference's hole filling its labelling kernel back: ``sam2.utils.misc.get_connected_components`` imports
    the native extension ``sam2._C`` (misc.py:47-63) and ``fill_holes_in_mask_scores`` silently skips the step when it is
    missing (misc.py:325-336).  After ``install_postprocess(sam2.utils.misc)`` the reference predictors' ``fill_hole_area``
    works on the B200 box through ``sam2b_op_connected_components`` (same (labels, counts) contract)."""
    from . import postprocess
    misc_module.get_connected_components = postprocess.get_connected_components
    return misc_module
