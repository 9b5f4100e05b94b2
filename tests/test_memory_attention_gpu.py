"""GPU parity of the MemoryAttention sub-graph (boundary E3) through the C ABI against the reference golden slices and
the fp32 CPU oracle.  Tolerance: relative L2 <= 2e-2 (bf16 operands, fp32 accumulation / statistics)."""
import numpy as np
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


@pytest.mark.parametrize("tag,L,P,N", [("a", 2, 8, 1), ("b", 1, 0, 1), ("c", 3, 64, 2)])
def test_e3_vs_reference_golden(engine, golden, tag, L, P, N):
    from oracle.synth import synth_memattn_inputs
    ins = [t.cuda() for t in synth_memattn_inputs(L, P, N, seed=3)]
    out = engine.memory_attention(ins[0], ins[1], ins[2], ins[3], ins[4] if P else None, ins[5] if P else None)
    torch.cuda.synchronize()
    assert out.shape == (4096, N, 256)
    r = rel_l2(out[::64, :, ::8], golden[f"e3{tag}_out"])
    print(f"E3{tag} (L={L},P={P},N={N}): rel-L2 vs reference golden slice = {r:.4%}")
    assert r <= TOL


def test_e3_full_vs_oracle_steady_state_shape(engine, weights):
    """Full 7-frame bank + 64 pointer tokens (28 736 keys), whole tensor against the CPU oracle."""
    from oracle import sam2_oracle as O
    from oracle.synth import synth_memattn_inputs
    ins = synth_memattn_inputs(7, 64, 1, seed=11)
    with torch.no_grad():
        ref = O.memory_attention(weights, *ins)
    out = engine.memory_attention(*[t.cuda() for t in ins])
    torch.cuda.synchronize()
    r = rel_l2(out, ref)
    print(f"E3 L=7 P=64: rel-L2 vs oracle = {r:.4%}, max-abs {(out.cpu() - ref).abs().max().item():.4g}")
    assert r <= TOL


def test_e3_more_than_seven_memory_frames(engine, weights):
    """Several conditioning frames make the bank longer than the 7 slots of the reference's exported engines (its torch
    path has no such limit, sam2_base_official.py:839-872): L = 9, P = 16 against the CPU oracle."""
    from oracle import sam2_oracle as O
    from oracle.synth import synth_memattn_inputs
    ins = synth_memattn_inputs(9, 16, 1, seed=12)
    with torch.no_grad():
        ref = O.memory_attention(weights, *ins)
    out = engine.memory_attention(*[t.cuda() for t in ins])
    torch.cuda.synchronize()
    r = rel_l2(out, ref)
    print(f"E3 L=9 P=16: rel-L2 vs oracle = {r:.4%}")
    assert r <= TOL


def test_e3_argument_errors(engine):
    from oracle.synth import synth_memattn_inputs
    ins = [t.cuda() for t in synth_memattn_inputs(1, 4, 1, seed=1)]
    with pytest.raises(RuntimeError, match="L="):
        engine.memory_attention(ins[0], torch.zeros(17, 4096, 1, 64, device="cuda"), ins[2],
                                torch.zeros(17, 4096, 1, 64, device="cuda"), ins[4], ins[5])
