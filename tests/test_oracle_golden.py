"""CPU: the oracle (oracle/sam2_oracle.py) against the committed golden fixtures, which were produced by running the
UNMODIFIED reference (oracle/make_golden.py, in the build container).  fp32 vs fp32 -> tight tolerances."""
import numpy as np
import pytest
import torch

from oracle import sam2_oracle as O
from oracle.synth import (checksum, synth_image_norm, synth_image_u8, synth_memattn_inputs, synth_memenc_inputs,
                          synth_video)

torch.set_grad_enabled(False)


@pytest.fixture(scope="module")
def w():
    return O.make_weights(0)


def _sl16(a):
    s = a.shape[-1] // 16
    return a[0, :16, ::s, ::s].numpy()


def test_weights_and_inputs_are_the_fixture_ones(w, golden):
    np.testing.assert_allclose(
        [checksum(w["image_encoder.trunk.blocks.9.attn.qkv.weight"]), checksum(w["memory_attention.layers.2.linear1.weight"])],
        golden["weights_checksum"], rtol=1e-9)
    np.testing.assert_allclose(checksum(synth_image_norm(1, 1)), golden["e2_input_checksum"][0], rtol=1e-9)


def test_weight_spec_matches_reference_param_count(w):
    assert len(w) == 903
    assert abs(sum(v.numel() for v in w.values()) / 1e6 - 224.45) < 0.01


def test_forward_image_e2(w, golden):
    img = synth_image_norm(1, 1)
    xs = O.hiera_trunk(w, img, return_all=True)
    for i, x in enumerate(xs):
        s = max(1, x.shape[1] // 4)
        np.testing.assert_allclose(x[0, ::s, ::s, :8].numpy(), golden[f"e2_block_probe_{i}"], rtol=2e-4, atol=2e-4)
    outs = O.forward_image(w, img)
    for i, o in enumerate(outs):
        np.testing.assert_allclose(_sl16(o), golden[f"e2_out{i}"], rtol=2e-4, atol=2e-4)
        assert abs(o.double().norm().item() - golden[f"e2_out{i}_stats"][3]) < 1e-4 * golden[f"e2_out{i}_stats"][3]


def test_set_image_and_predict_c1(w, golden):
    img_u8 = synth_image_u8(0)
    outs = O.set_image_e2e(w, torch.as_tensor(img_u8).permute(2, 0, 1)[None].float() / 255.0)
    for i, o in enumerate(outs):
        np.testing.assert_allclose(_sl16(o), golden[f"e1_out{i}"], rtol=2e-4, atol=2e-4)
    masks, iou, low = O.image_predict(w, img_u8, [[512.0, 512.0]], [1], True)
    np.testing.assert_allclose(iou.numpy(), golden["c1_iou"], atol=1e-5)
    np.testing.assert_allclose(low[:, ::16, ::16].numpy(), golden["c1_low_res_slice"], atol=1e-4)
    bits = np.unpackbits(golden["c1_low_res_bits"]).reshape(3, 256, 256).astype(bool)
    assert ((low > 0).numpy() != bits).mean() < 1e-4
    np.testing.assert_allclose((masks > 0).sum(dim=(1, 2)).numpy(), golden["c1_mask_area"][0], rtol=1e-3)


@pytest.mark.parametrize("tag,L,P,N", [("a", 2, 8, 1), ("b", 1, 0, 1), ("c", 3, 64, 2)])
def test_memory_attention_e3(w, golden, tag, L, P, N):
    ins = synth_memattn_inputs(L, P, N, seed=3)
    np.testing.assert_allclose([checksum(t) for t in ins], golden[f"e3{tag}_input_checksum"], rtol=1e-9)
    out = O.memory_attention(w, *ins)
    np.testing.assert_allclose(out[::64, :, ::8].numpy(), golden[f"e3{tag}_out"], rtol=1e-3, atol=2e-4)


def test_memory_encoder_e4(w, golden):
    pix, msk = synth_memenc_inputs(2, seed=4)
    x, pos = O.memory_encoder(w, pix, msk)
    np.testing.assert_allclose(x[:, ::4, ::8, ::8].numpy(), golden["e4_out"], rtol=1e-3, atol=1e-4)
    np.testing.assert_allclose(pos[0, ::4, ::8, ::8].numpy(), golden["e4_pos"], atol=1e-6)


def test_video_propagate_c3(w, golden):
    T = 4
    frames = synth_video(T, seed=1)
    np.testing.assert_allclose(checksum(frames), golden["c3_input_checksum"][0], rtol=1e-9)
    for t, low in O.video_propagate(w, frames):
        np.testing.assert_allclose(low[0, 0, ::16, ::16].numpy(), golden["c3_low_res_slice"][t], atol=2e-4)
        bits = np.unpackbits(golden["c3_low_res_bits"][t]).reshape(256, 256).astype(bool)
        assert ((low[0, 0] > 0).numpy() != bits).mean() < 1e-3
