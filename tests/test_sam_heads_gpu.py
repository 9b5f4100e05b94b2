"""GPU parity of the native SAM heads (sam2b_sam_heads) through the C ABI against the fp32 CPU
oracle (oracle/sam2_oracle.py forward_sam_heads / mask_decoder, pinned against the unmodified reference).

Tolerances (the reference keeps these heads in fp32; the image side here is split-bf16 on the tensor cores, ~2^-16 per
product): mask logits / object pointers rel-L2 <= 1e-3, predicted IoUs and object scores abs <= 1e-3, mask IoU >= 0.999."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-3


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


def mask_iou(a, b):
    a, b = (torch.as_tensor(a).cpu() > 0), (torch.as_tensor(b).cpu() > 0)
    u = (a | b).sum().item()
    return 1.0 if u == 0 else (a & b).sum().item() / u


@pytest.fixture(scope="module")
def weights():
    from oracle import sam2_oracle as O
    return O.make_weights(0)


@pytest.fixture(scope="module")
def engine(weights):
    from sam2_opt_b200.engine import Engine
    e = Engine(weights)
    yield e
    e.release()


def _inputs(n_img, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n_img, 256, 64, 64, generator=g), 0.5 * torch.randn(n_img, 32, 256, 256, generator=g),
            0.5 * torch.randn(n_img, 64, 128, 128, generator=g), g)


def _check(got, want, names):
    for name, a, b in zip(names, got, want):
        if a is None:
            continue
        assert tuple(a.shape) == tuple(b.shape), (name, a.shape, b.shape)
        if name in ("ious", "object_score_logits"):
            err = (a.cpu() - b).abs().max().item()
        else:
            err = rel_l2(a, b)
        print(f"sam_heads {name}: err {err:.3e}")
        assert err <= TOL, (name, err)


NAMES = ("low_res_multimasks", "high_res_multimasks", "ious", "low_res_masks", "high_res_masks", "obj_ptr", "object_score_logits")


def test_tracked_frame_e3_layout_shared_image(engine, weights):
    """A tracked frame: no prompt, multimask output, three objects whose features arrive in the E3 output layout
    [4096, n, 256] (viewed as [n,256,64,64]) and share the frame's high-resolution levels."""
    from oracle import sam2_oracle as O
    n = 3
    pix, s0, s1, g = _inputs(n, 1)
    s0, s1 = s0[:1], s1[:1]
    seq = pix.flatten(2).permute(2, 0, 1).contiguous().cuda()                     # [4096, n, 256]
    view = seq.permute(1, 2, 0).reshape(n, 256, 64, 64)
    got = engine.sam_heads(view, [s0.cuda(), s1.cuda()], multimask_output=True, high_res_multimasks=True)
    torch.cuda.synchronize()
    with torch.no_grad():
        want = O.forward_sam_heads(weights, pix, high_res_features=[s0.expand(n, -1, -1, -1), s1.expand(n, -1, -1, -1)],
                                   multimask_output=True)
    _check(got, want, NAMES)
    for b in range(n):
        assert mask_iou(got[3][b], want[3][b]) >= 0.999
        assert mask_iou(got[4][b], want[4][b]) >= 0.999


def test_prompted_points_dense_dynamic_single_mask(engine, weights):
    """Clicks and box corners with a mask prompt, single-mask output with the dynamic stability fall-back; per-object images."""
    from oracle import sam2_oracle as O
    n, P = 2, 4
    pix, s0, s1, g = _inputs(n, 2)
    coords = torch.rand(n, P, 2, generator=g) * 1024
    labels = torch.tensor([[1, 0, -1, 1], [2, 3, 1, 0]], dtype=torch.int32)
    mask_in = 4.0 * torch.randn(n, 1, 256, 256, generator=g)
    with torch.no_grad():
        _, dense = O.prompt_encoder(weights, coords, labels, mask_in)
        want = O.forward_sam_heads(weights, pix, point_inputs={"point_coords": coords, "point_labels": labels}, mask_inputs=mask_in,
                                   high_res_features=[s0, s1], multimask_output=False)
    got = engine.sam_heads(pix.cuda(), [s0.cuda(), s1.cuda()], point_coords=coords.cuda(), point_labels=labels.cuda(),
                           dense_embedding=dense.contiguous().cuda(), multimask_output=False)
    torch.cuda.synchronize()
    _check(got, want, NAMES)
    for b in range(n):
        assert mask_iou(got[3][b], want[3][b]) >= 0.999


def test_decoder_only_repeated_image_many_prompts(engine, weights):
    """SAM2ImagePredictor.predict with a batch of prompts on one image (mask_decoder.py:252-259 repeat_image): 70 one-click
    prompts (more than one micro-batch of 64) against the oracle's plain mask decoder (no NO_OBJ_SCORE fill)."""
    from oracle import sam2_oracle as O
    n = 70
    pix, s0, s1, g = _inputs(1, 3)
    coords = torch.rand(n, 1, 2, generator=g) * 1024
    labels = torch.ones(n, 1, dtype=torch.int32)
    got = engine.sam_heads(pix.cuda().expand(n, -1, -1, -1), [s0.cuda(), s1.cuda()], point_coords=coords.cuda(),
                           point_labels=labels.cuda(), multimask_output=True, decoder_only=True, high_res=False)
    torch.cuda.synchronize()
    sel = [0, 1, 63, 64, 69]
    with torch.no_grad():
        sparse, dense = O.prompt_encoder(weights, coords[sel], labels[sel], None)
        masks, iou, _, obj = O.mask_decoder(weights, pix, sparse, dense, True, [s0, s1], repeat_image=True)
    assert rel_l2(got[0][sel], masks) <= TOL
    assert (got[2][sel].cpu() - iou).abs().max().item() <= TOL
    assert (got[6][sel].cpu() - obj).abs().max().item() <= TOL


def test_matches_torch_heads_and_is_deterministic(engine, weights):
    """Against the package's own fp32 torch heads on the same GPU (the path this entry point replaces in the predictors),
    and bit-identical across two calls (no atomics on floating-point data)."""
    from sam2_opt_b200 import sam_heads as H
    n = 2
    pix, s0, s1, g = _inputs(1, 4)
    w = {k: v.cuda() for k, v in weights.items() if k.startswith(("sam_", "obj_ptr", "no_obj"))}
    pixn = (pix + 0.1 * torch.randn(n, 256, 64, 64, generator=g)).cuda()
    hr = [s0.cuda(), s1.cuda()]
    a = engine.sam_heads(pixn, hr, multimask_output=True)
    b = engine.sam_heads(pixn, hr, multimask_output=True)
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        if x is not None:
            assert torch.equal(x, y)
    prev = torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = False
    try:
        with torch.no_grad():
            low, high, ious, low_best, high_best, ptr, obj = H.forward_sam_heads(
                w, pixn, high_res_features=[h.expand(n, -1, -1, -1) for h in hr], multimask_output=True)
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = prev
    _check(a, (low.cpu(), None, ious.cpu(), low_best.cpu(), high_best.cpu(), ptr.cpu(), obj.cpu()), NAMES)


def test_sparse_embeddings_given_and_mask_tokens_out(engine, weights):
    """MaskDecoder.forward semantics: the caller's own sparse prompt embeddings (here the oracle prompt encoder's, box + point)
    instead of points, and no sparse token at all with a mask prompt (SAM2ImagePredictor.predict with mask_input only)."""
    from oracle import sam2_oracle as O
    pix, s0, s1, g = _inputs(1, 6)
    n = 3
    coords = torch.rand(n, 3, 2, generator=g) * 1024
    labels = torch.tensor([[2, 3, 1]] * n, dtype=torch.int32)
    with torch.no_grad():
        sparse, dense = O.prompt_encoder(weights, coords, labels, None)
        masks, iou, sam_tok, obj = O.mask_decoder(weights, pix, sparse, dense, True, [s0, s1], repeat_image=True)
    got = engine.sam_heads(pix.cuda().expand(n, -1, -1, -1), [s0.cuda(), s1.cuda()], sparse_embeddings=sparse.cuda(), multimask_output=True,
                           decoder_only=True, high_res=False, mask_tokens=True)
    assert rel_l2(got[0], masks) <= TOL and (got[2].cpu() - iou).abs().max().item() <= TOL
    assert rel_l2(got[7][:, 1:], sam_tok) <= TOL and (got[6].cpu() - obj).abs().max().item() <= TOL
    mask_in = 4.0 * torch.randn(1, 1, 256, 256, generator=g)
    with torch.no_grad():
        sparse0, dense0 = O.prompt_encoder(weights, None, None, mask_in)
        masks0, iou0, tok0, obj0 = O.mask_decoder(weights, pix, sparse0, dense0, False, [s0, s1])
    assert sparse0.shape[1] == 0
    got0 = engine.sam_heads(pix.cuda(), [s0.cuda(), s1.cuda()], dense_embedding=dense0.contiguous().cuda(), sparse_embeddings=sparse0.cuda(),
                            multimask_output=False, decoder_only=True, high_res=False, mask_tokens=True)
    assert rel_l2(got0[0], masks0) <= TOL and (got0[2].cpu() - iou0).abs().max().item() <= TOL
    assert rel_l2(got0[7][:, 0:1], tok0) <= TOL


def test_rejects_bad_arguments(engine):
    """Errors are loud and name the problem (no silent fallback): too many points, mismatched shapes, host tensors."""
    pix = torch.randn(1, 256, 64, 64, device="cuda")
    hr = [torch.randn(1, 32, 256, 256, device="cuda"), torch.randn(1, 64, 128, 128, device="cuda")]
    with pytest.raises(RuntimeError, match="points per prompt"):
        engine.sam_heads(pix, hr, torch.rand(1, 26, 2, device="cuda") * 1024, torch.ones(1, 26, dtype=torch.int32, device="cuda"))
    with pytest.raises(RuntimeError, match="sparse embeddings per prompt"):
        engine.sam_heads(pix, hr, sparse_embeddings=torch.randn(1, 27, 256, device="cuda"))
    with pytest.raises(ValueError, match="backbone_features"):
        engine.sam_heads(pix[:, :128], hr)
    with pytest.raises(ValueError, match="high-res feature"):
        engine.sam_heads(pix, [hr[1], hr[0]])
    with pytest.raises(RuntimeError, match="CUDA fp32"):
        engine.sam_heads(pix.cpu(), hr)
    with pytest.raises(ValueError, match="memory_encode_low_res"):
        engine.memory_encode_low_res(pix, torch.randn(1, 1, 128, 128, device="cuda"))


def test_mask_downscale_matches_the_prompt_encoder(engine, weights):
    """sam2b_mask_downscale against the oracle's PromptEncoder mask branch (prompt_encoder.py:97-109,152-155), fp32."""
    from oracle import sam2_oracle as O
    g = torch.Generator().manual_seed(8)
    m = 5.0 * torch.randn(3, 1, 256, 256, generator=g)
    with torch.no_grad():
        _, want = O.prompt_encoder(weights, None, None, m)
    got = engine.mask_downscale(m.cuda())
    r = rel_l2(got, want)
    print(f"mask_downscale rel-L2 vs oracle = {r:.2e}")
    assert got.shape == (3, 256, 64, 64) and r <= 1e-5
