"""GPU parity of the image-encoder sub-graph (boundaries E1 / E2) through the C ABI against
(a) the committed golden fixtures produced by the unmodified reference and (b) the fp32 oracle run on the host CPU.
Tolerance (BASELINE.json north_star): relative error <= 2e-2 for embeddings and high-res features."""
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


def _sl16(a):
    s = a.shape[-1] // 16
    return a[0, :16, ::s, ::s]


def test_e2_forward_image_vs_reference_golden(engine, golden):
    from oracle.synth import synth_image_norm
    img = synth_image_norm(1, 1).cuda()
    f0, f1, f2 = engine.image_encode(img, normalised=True, add_no_mem_embed=False)
    torch.cuda.synchronize()
    for name, got, key in (("fpn0", f0, "e2_out4"), ("fpn1", f1, "e2_out5"), ("fpn2", f2, "e2_out6")):
        r = rel_l2(_sl16(got), golden[key])
        print(f"E2 {name}: rel-L2 vs reference golden slice = {r:.4%}")
        assert r <= TOL, (name, r)
        ref_norm = golden[key + "_stats"][3]
        assert abs(got.double().norm().item() - ref_norm) <= TOL * ref_norm


def test_e1_set_image_vs_reference_golden(engine, golden):
    from oracle.synth import synth_image_u8
    img = (torch.as_tensor(synth_image_u8(0)).permute(2, 0, 1)[None].float() / 255.0).cuda()
    outs = engine.image_encode(img, normalised=False, add_no_mem_embed=True)
    torch.cuda.synchronize()
    for i, got in enumerate(outs):
        r = rel_l2(_sl16(got), golden[f"e1_out{i}"])
        print(f"E1 feat{i}: rel-L2 vs reference golden slice = {r:.4%}")
        assert r <= TOL, (i, r)


def test_e2_full_tensors_vs_oracle_and_batch_invariance(engine, weights):
    """Whole-tensor comparison against the CPU oracle on a fresh seed, inside a batch of 3 (different images), plus:
    each image's result must not depend on its batch neighbours."""
    from oracle import sam2_oracle as O
    from oracle.synth import synth_image_norm
    imgs = synth_image_norm(3, seed=5)
    with torch.no_grad():
        ref = O.forward_image(weights, imgs[1:2])
    f0, f1, f2 = engine.image_encode(imgs.cuda(), normalised=True, add_no_mem_embed=False)
    s0, s1, s2 = engine.image_encode(imgs[1:2].cuda(), normalised=True, add_no_mem_embed=False)
    torch.cuda.synchronize()
    for name, got, single, want in (("fpn0", f0, s0, ref[4]), ("fpn1", f1, s1, ref[5]), ("fpn2", f2, s2, ref[6])):
        r = rel_l2(got[1:2], want)
        mx = (got[1:2].cpu() - want).abs().max().item()
        print(f"E2 {name}: rel-L2 vs oracle = {r:.4%}, max-abs = {mx:.4g}")
        assert r <= TOL, (name, r)
        assert rel_l2(got[1:2], single) < 1e-3


def test_rejects_bad_inputs(engine):
    with pytest.raises(ValueError):
        engine.image_encode(torch.zeros(1, 3, 512, 512, device="cuda"), True, False)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        engine.image_encode(torch.zeros(1, 3, 1024, 1024), True, False)


def test_host_buffer_entry_point_matches_device_path(engine):
    """sam2b_image_encode_host (pinned host in / out, uploads, compute and downloads overlapped over two micro-batch slots,
    level-0/1 outputs leaving the device early) must return exactly what sam2b_image_encode returns: 18 images = one full
    micro-batch of 16 plus a ragged one, called twice so that both staging slots are reused."""
    from oracle.synth import synth_image_norm
    imgs = synth_image_norm(18, seed=9)
    dev = engine.image_encode(imgs.cuda(), normalised=True, add_no_mem_embed=False)
    torch.cuda.synchronize()
    host = None
    for _ in range(2):
        host = engine.image_encode_host(imgs.pin_memory(), True, False, host)
    for d, h in zip(dev, host):
        assert torch.equal(d.cpu(), h)


def test_async_host_calls_overlap_without_changing_results(engine):
    """sam2b_image_encode_host_async: three calls in flight over two sets of pinned buffers (an odd number of micro-batches
    per call, so that consecutive calls start on different staging slots), one sam2b_host_sync at the end; every call must
    return exactly what the device path returns, and a synchronous call afterwards still works."""
    from oracle.synth import synth_image_norm
    batches = [synth_image_norm(20, seed=21 + k) for k in range(3)]
    want = []
    for b in batches:
        want.append([t.cpu() for t in engine.image_encode(b.cuda(), normalised=True, add_no_mem_embed=False)])
    torch.cuda.synchronize()
    pinned = [b.pin_memory() for b in batches]
    outs = [engine.image_encode_host(pinned[k], True, False, None, wait=False) for k in range(3)]
    engine.host_sync()
    for k in range(3):
        for w, h in zip(want[k], outs[k]):
            assert torch.equal(w, h), f"async call {k} differs from the device path"
    again = engine.image_encode_host(pinned[0], True, False)
    for w, h in zip(want[0], again):
        assert torch.equal(w, h)


def test_fused_mlp_path_matches_the_two_gemm_path(monkeypatch):
    """mlp_fused.cu (fc1 -> GELU -> fc2 in one kernel for the stage-1 blocks, hidden layer in tensor memory; the default)
    against the fc1 / fc2 GEMM pair (SAM2B_FUSED_MLP=0) on the same engine: same arithmetic up to the order of the fp32 sums
    (and one bf16 rounding of the hidden layer each way)."""
    from oracle import sam2_oracle as O
    from sam2_opt_b200.engine import Engine
    e = Engine(O.make_weights(0))
    g = torch.Generator(device="cuda").manual_seed(5)
    img = torch.randn(2, 3, 1024, 1024, device="cuda", generator=g)
    from sam2_opt_b200.engine import profile_begin, profile_end
    monkeypatch.setenv("SAM2B_FUSED_MLP", "0")
    ref = [o.clone() for o in e.image_encode(img, True, False)]
    monkeypatch.delenv("SAM2B_FUSED_MLP", raising=False)
    profile_begin()
    got = e.image_encode(img, True, False)
    assert sum(1 for r in profile_end() if r[0].startswith("mlp_fused")) == 2, "the fused MLP is not the default for blocks 0 and 1"
    for a, b in zip(got, ref):
        r = float((a - b).norm() / b.norm())
        print(f"fused MLP (default) vs GEMM pair: rel-L2 {r:.3e}")
        assert r < 5e-3
    e.release()
