"""GPU (needs two devices; skipped otherwise): one process driving engines on two GPUs.  Launch attributes (opt-in shared
memory) are per device, the library must not leave the caller on another device, and calls enqueue on the current stream of
the ENGINE's device."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_engines_on_two_devices_in_one_process():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from oracle import sam2_oracle as O
    from oracle.synth import synth_image_norm
    from sam2_opt_b200 import ops
    from sam2_opt_b200.engine import Engine
    w = O.make_weights(0)
    img = synth_image_norm(2, seed=3)
    torch.cuda.set_device(0)
    e0 = Engine(w, device="cuda:0")
    ref = [t.cpu() for t in e0.image_encode(img.to("cuda:0"), True, False)]
    e1 = Engine(w, device="cuda:1")                      # created and called while cuda:0 is the current device
    assert torch.cuda.current_device() == 0
    got = e1.image_encode(img.to("cuda:1"), True, False)
    assert torch.cuda.current_device() == 0, "the library changed the caller's current device"
    for a, b in zip(got, ref):
        assert a.device.index == 1
        assert torch.equal(a.cpu(), b)
    # the video-path entries and a raw op on the second device
    g = torch.Generator(device="cuda:1").manual_seed(0)
    curr = torch.randn(4096, 1, 256, device="cuda:1", generator=g)
    mem = torch.randn(2, 4096, 1, 64, device="cuda:1", generator=g)
    pos = torch.randn(4096, 1, 256, device="cuda:1", generator=g)
    mpos = torch.randn(2, 4096, 1, 64, device="cuda:1", generator=g)
    o1 = e1.memory_attention(curr, mem, pos, mpos)
    o0 = e0.memory_attention(curr.to("cuda:0"), mem.to("cuda:0"), pos.to("cuda:0"), mpos.to("cuda:0"))
    assert torch.equal(o1.cpu(), o0.cpu())
    with torch.cuda.device(1):
        q = torch.randn(1, 4096, 256, device="cuda:1", generator=g).bfloat16()
        k = torch.randn(1, 8192, 256, device="cuda:1", generator=g).bfloat16()
        y1 = ops.attention_hd256(q, k, k)
    y0 = ops.attention_hd256(q.to("cuda:0"), k.to("cuda:0"), k.to("cuda:0"))
    assert torch.equal(y1.cpu(), y0.cpu())
    e0.release(); e1.release()
