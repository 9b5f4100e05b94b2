"""GPU parity of the MemoryEncoder sub-graph (boundary E4) through the C ABI against the reference golden slices and the
fp32 CPU oracle.  Tolerance: relative L2 <= 2e-2."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 2e-2


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return ((a - b).norm() / b.norm()).item()


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


def test_e4_vs_reference_golden_and_oracle(engine, weights, golden):
    from oracle import sam2_oracle as O
    from oracle.synth import synth_memenc_inputs
    pix, msk = synth_memenc_inputs(2, seed=4)
    x, pos = engine.memory_encode(pix.cuda(), msk.cuda())
    torch.cuda.synchronize()
    assert x.shape == (2, 64, 64, 64) and pos.shape == (2, 64, 64, 64)
    r = rel_l2(x[:, ::4, ::8, ::8], golden["e4_out"])
    print(f"E4: rel-L2 vs reference golden slice = {r:.4%}")
    assert r <= TOL
    assert (pos[0, ::4, ::8, ::8].cpu() - torch.as_tensor(golden["e4_pos"])).abs().max().item() < 1e-5
    with torch.no_grad():
        ref, _ = O.memory_encoder(weights, pix, msk)
    r = rel_l2(x, ref)
    print(f"E4: rel-L2 vs oracle (full tensor) = {r:.4%}")
    assert r <= TOL


def test_e4_binary_mask_and_single_object(engine, weights):
    """Clicked frames feed a binarised mask (+10 / -10), sam2_base_official.py:1000-1013."""
    from oracle import sam2_oracle as O
    from oracle.synth import synth_memenc_inputs
    pix, msk = synth_memenc_inputs(1, seed=9)
    msk = (msk > 0).float() * 20.0 - 10.0
    x, _ = engine.memory_encode(pix.cuda(), msk.cuda())
    with torch.no_grad():
        ref, _ = O.memory_encoder(weights, pix, msk)
    assert rel_l2(x, ref) <= TOL


def test_e4_from_low_res_logits_matches_the_materialised_mask(engine, weights):
    """sam2b_memory_encode_low_res: mask_for_mem = sigmoid(bilinear_x4(low)) * 20 - 10 (or its binarised form) evaluated inside
    the first down-sampler convolution, the no-object embedding added in the library — against the E4 entry point fed with
    the materialised 1024^2 mask (what SAM2Base._encode_new_memory builds, sam2_base_official.py:978-1026)."""
    import torch.nn.functional as F
    from oracle.synth import synth_memenc_inputs
    n = 3
    pix, _ = synth_memenc_inputs(n, seed=11)
    g = torch.Generator().manual_seed(12)
    low = (6.0 * torch.randn(n, 1, 256, 256, generator=g)).cuda()
    low[1] = -1024.0                                            # an object the heads declared absent (NO_OBJ_SCORE fill)
    obj = torch.tensor([[3.0], [-2.0], [0.5]]).cuda()
    nos = weights["no_obj_embed_spatial"].cuda()
    for binarize in (False, True):
        for shared in (False, True):
            p = pix[:1].cuda() if shared else pix.cuda()
            high = F.interpolate(low, size=(1024, 1024), mode="bilinear", align_corners=False)
            m = (high > 0).float() if binarize else torch.sigmoid(high)
            want, _ = engine.memory_encode(p.expand(n, -1, -1, -1), m * 20.0 - 10.0)
            want = want + (1 - (obj > 0).float()[..., None, None]) * nos[..., None, None]
            got, pos = engine.memory_encode_low_res(p, low, obj, binarize=binarize)
            r = rel_l2(got, want)
            print(f"E4 from low-res logits (binarize={binarize}, shared frame={shared}): rel-L2 vs the materialised-mask path = {r:.2e}")
            assert got.shape == (n, 64, 64, 64) and r <= 1e-3
