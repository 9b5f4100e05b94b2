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
