"""uint8 -> fp32 image preprocessing: the oracle is pinned on CPU against the operators the reference
calls (torch `div`/`sub`, sam2/utils/transforms.py:27-46; numpy float64 route of utils/misc.py:96), the CUDA op is
compared with the oracle bit for bit through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import preprocess_oracle as O


def _frames(n, h, w, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, size=(n, h, w, 3), dtype=np.uint8)
    x[0, 0, :min(w, 256), 0] = np.arange(min(w, 256), dtype=np.uint8)     # every byte value at least once
    return x


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


# ---- CPU: pin the oracle --------------------------------------------------------------------------------------------
def test_oracle_equals_torch_totensor_and_normalize():
    x = _frames(2, 8, 256)
    t = torch.from_numpy(x).permute(0, 3, 1, 2).to(torch.float32).div(255)            # ToTensor
    assert np.array_equal(_bits(O.to_tensor(x)), _bits(t.numpy()))
    mean = torch.tensor(O.IMG_MEAN, dtype=torch.float32)[None, :, None, None]
    std = torch.tensor(O.IMG_STD, dtype=torch.float32)[None, :, None, None]
    t = t.clone().sub_(mean).div_(std)                                                 # Normalize
    assert np.array_equal(_bits(O.preprocess_u8(x, True)), _bits(t.numpy()))


def test_oracle_float64_loader_route_rounds_to_the_same_fp32():
    x = np.zeros((1, 1, 256, 3), np.uint8)
    x[0, 0, :, :] = np.arange(256, dtype=np.uint8)[:, None]
    assert np.array_equal(_bits(O.video_loader_route(x)), _bits(O.preprocess_u8(x, True)))


def test_oracle_known_answers():
    x = np.array([[[[0, 255, 51]]]], np.uint8)
    y = O.preprocess_u8(x, False)
    assert y.shape == (1, 3, 1, 1)
    assert y[0, 0, 0, 0] == 0.0 and y[0, 1, 0, 0] == 1.0 and y[0, 2, 0, 0] == np.float32(0.2)
    z = O.preprocess_u8(x, True)
    assert z[0, 0, 0, 0] == np.float32(np.float32(0.0 - np.float32(0.485)) / np.float32(0.229))
    assert z[0, 1, 0, 0] == np.float32(np.float32(np.float32(1.0) - np.float32(0.456)) / np.float32(0.224))


def test_wrapper_rejects_bad_inputs_without_a_gpu():
    from sam2_opt_b200 import ops
    with pytest.raises(ValueError):
        ops.preprocess_u8(torch.zeros(4, 4, 3, dtype=torch.uint8))          # host tensor
    with pytest.raises(ValueError):
        ops.preprocess_u8(torch.zeros(1, 4, 4, 3, dtype=torch.float32))


# ---- GPU: parity through the C ABI --------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n,h,w", [(1, 1024, 1024), (3, 64, 260), (2, 5, 4), (1, 1, 4)])
@pytest.mark.parametrize("normalise", [False, True])
def test_preprocess_u8_bit_exact(n, h, w, normalise):
    from sam2_opt_b200 import ops
    x = _frames(n, h, w, seed=n * 7 + h)
    y = ops.preprocess_u8(torch.from_numpy(x).cuda(), normalise)
    assert y.shape == (n, 3, h, w) and y.dtype == torch.float32
    assert np.array_equal(_bits(y.cpu().numpy()), _bits(O.preprocess_u8(x, normalise)))


@pytest.mark.gpu
def test_preprocess_u8_single_frame_and_width_error():
    from sam2_opt_b200 import ops
    x = _frames(1, 16, 16, seed=3)
    y = ops.preprocess_u8(torch.from_numpy(x[0]).cuda())
    assert np.array_equal(_bits(y.cpu().numpy()), _bits(O.preprocess_u8(x, False)))
    with pytest.raises(RuntimeError, match="multiple of 4"):
        ops.preprocess_u8(torch.zeros(1, 8, 6, 3, dtype=torch.uint8, device="cuda"))


@pytest.mark.gpu
def test_predictor_prep_paths_use_the_op_and_match_the_oracle():
    """_prep_image / init_state on 1024x1024 uint8 frames == oracle (un-normalised for the image predictor, which
    normalises inside the engine; normalised for the video loader route)."""
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    from oracle import sam2_oracle
    vp = SAM2VideoPredictor(sam2_oracle.make_weights(0))
    x = _frames(2, 1024, 1024, seed=11)
    a = vp._prep_image(x[0])
    assert np.array_equal(_bits(a.cpu().numpy()), _bits(O.preprocess_u8(x[:1], False)))
    st = vp.init_state([x[0], x[1]])
    assert np.array_equal(_bits(st["images"].cpu().numpy()), _bits(O.video_loader_route(x)))


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 18])
def test_image_encode_host_u8_equals_the_fp32_host_entry_bit_for_bit(n):
    """sam2b_image_encode_host_u8 (raw frames over PCIe, ToTensor -> Normalize on the device) == sam2b_image_encode_host on
    the frames the oracle preprocessed (n = 18: two micro-batch slots, the second ragged)."""
    from oracle import sam2_oracle
    from sam2_opt_b200.engine import Engine
    e = Engine(sam2_oracle.make_weights(0))
    try:
        rng = np.random.default_rng(21)
        base = rng.integers(0, 256, size=(2, 64, 64, 3), dtype=np.uint8)
        x = np.stack([np.kron(base[i % 2], np.ones((16, 16, 1), np.uint8)) + np.uint8(i) for i in range(n)])   # blocky 1024^2 frames
        ref = e.image_encode_host(torch.from_numpy(O.preprocess_u8(x, True)).pin_memory(), True, False)
        ref = [t.clone() for t in ref]
        got = e.image_encode_host_u8(torch.from_numpy(x).pin_memory(), False)
        for a, b in zip(got, ref):
            assert torch.equal(a, b)
        with pytest.raises(ValueError):
            e.image_encode_host_u8(torch.zeros(1, 512, 512, 3, dtype=torch.uint8), False)
    finally:
        e.release()


# ---- antialiased bilinear resize (frames that are not at the model resolution) --------------------------------------------
AA_SIZES = [(540, 960), (1200, 1800), (333, 517), (1024, 768), (2160, 3840)]


@pytest.mark.parametrize("hw", AA_SIZES[:4])
def test_resize_oracle_is_pinned_against_the_reference_frameworks_cpu_operator(hw):
    """oracle/preprocess_oracle.py:resize_bilinear_aa against torch's CPU F.interpolate(mode="bilinear", antialias=True,
    align_corners=False) — the operator behind the reference's Resize((1024, 1024)) (sam2/utils/transforms.py:30).  Stated
    bound: 1 unit in the last place of 1.0 (1.2e-7) on pixel values in [0, 1]; on this CPU the two agree bit for bit."""
    h, w = hw
    rng = np.random.default_rng(h * 7 + w)
    u8 = rng.integers(0, 256, (1, h, w, 3), dtype=np.uint8)
    got = O.resize_frames_u8(u8, 1024, 1024, False)
    x = torch.from_numpy(u8).permute(0, 3, 1, 2).float() / 255.0
    want = torch.nn.functional.interpolate(x, size=(1024, 1024), mode="bilinear", align_corners=False, antialias=True).numpy()
    d = np.abs(got - want)
    print(f"{hw}: max |oracle - torch| = {d.max():.3g}, bit-identical fraction {(d == 0).mean():.6f}")
    assert d.max() <= 1.2e-7
    assert (d == 0).mean() >= 0.99


def test_resize_oracle_known_answers():
    # identity when the size does not change; a constant image stays constant (weights are normalised to sum 1 up to rounding)
    x = np.random.default_rng(0).random((1, 3, 8, 8), dtype=np.float32)
    assert np.array_equal(O.resize_bilinear_aa(x, 8, 8), x)
    c = np.full((1, 3, 30, 50), 0.5, np.float32)
    assert np.abs(O.resize_bilinear_aa(c, 64, 64) - 0.5).max() <= 6e-8
    # 2x down-scaling of one axis: support 2, taps (1, 3, 3, 1) / 8 around each output centre in the interior
    r = np.arange(16, dtype=np.float32).reshape(1, 1, 1, 16).repeat(3, 1)
    y = O.resize_bilinear_aa(r, 1, 8)
    assert np.allclose(y[0, 0, 0, 1:-1], (r[0, 0, 0, 1:-3:2] + 3 * r[0, 0, 0, 2:-2:2] + 3 * r[0, 0, 0, 3:-1:2] + r[0, 0, 0, 4::2]) / 8, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("hw", AA_SIZES)
@pytest.mark.parametrize("normalise", [False, True])
def test_resize_frames_kernel_bit_exact_vs_oracle(hw, normalise):
    from sam2_opt_b200 import ops
    h, w = hw
    n = 1 if h * w > 4e6 else 2
    u8 = np.random.default_rng(h + w).integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    got = ops.resize_frames(torch.from_numpy(u8).cuda(), (1024, 1024), normalise=normalise).cpu().numpy()
    want = O.resize_frames_u8(u8, 1024, 1024, normalise)
    assert got.shape == (n, 3, 1024, 1024)
    assert np.array_equal(got, want), f"{hw}: {np.abs(got - want).max()} max abs difference"
    # fp32 planar input takes the same path after the division
    xf = (torch.from_numpy(u8).permute(0, 3, 1, 2).float() / 255.0).cuda()
    got_f = ops.resize_frames(xf, (1024, 1024), normalise=normalise).cpu().numpy()
    assert np.array_equal(got_f, want)


@pytest.mark.gpu
def test_predictors_resize_non_model_resolution_frames_with_the_kernel():
    """set_image on a 1800 x 1200 image and init_state on 960 x 540 frames: the tensors handed to the engine are the oracle's,
    bit for bit (the reference's own sample media have these sizes: notebooks/images/truck.jpg, notebooks/videos/bedroom)."""
    from sam2_opt_b200.predictor import SAM2ImagePredictor, SAM2VideoPredictor
    from sam2_opt_b200.synth_weights import make_weights
    w = make_weights(0)
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (1200, 1800, 3), dtype=np.uint8)
    p = SAM2ImagePredictor(w)
    assert np.array_equal(p._prep_image(img).cpu().numpy(), O.resize_frames_u8(img[None], 1024, 1024, False))
    p.set_image(img)
    masks, iou, low = p.predict(point_coords=np.array([[900, 600]]), point_labels=np.array([1]))
    assert masks.shape == (3, 1200, 1800)
    p.release()
    frames = [rng.integers(0, 256, (540, 960, 3), dtype=np.uint8) for _ in range(3)]
    vp = SAM2VideoPredictor(w)
    st = vp.init_state(frames)
    assert (st["video_height"], st["video_width"]) == (540, 960)
    assert np.array_equal(st["images"].cpu().numpy(), O.resize_frames_u8(np.stack(frames), 1024, 1024, True))
    vp.add_new_points_or_box(st, 0, 1, points=[[480, 270]], labels=[1])
    outs = list(vp.propagate_in_video(st))
    assert len(outs) == 3 and outs[0][2].shape == (1, 1, 540, 960)
    vp.release()
