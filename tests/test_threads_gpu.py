"""GPU: one engine shared by two host threads with private CUDA streams — the reference's own threading model
(video_multi_thread.py:36-88: two threads, `torch.cuda.stream(private)` each, one predictor; its executor serialises on the
host with a lock and on the device through a private work stream, tensorrt_executor.py:248-252,288-291).  All engine calls
share the named workspaces, so they must be serialised on the device as well: every result must be bit-identical to the
sequential single-stream run."""
import threading

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(dev, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    r = lambda *s: torch.randn(*s, device=dev, generator=g)
    return {
        "img": r(2, 3, 1024, 1024),
        "ma": (r(4096, 2, 256), r(3, 4096, 2, 64), r(4096, 2, 256), r(3, 4096, 2, 64), r(8, 2, 64), r(8, 2, 64)),
        "me": (r(2, 256, 64, 64), torch.rand(2, 1, 1024, 1024, device=dev, generator=g) * 20 - 10),
    }


def _work(eng, x, rounds):
    outs = []
    for _ in range(rounds):
        outs.append([t.clone() for t in eng.image_encode(x["img"], True, False)])
        outs.append([eng.memory_attention(*x["ma"]).clone()])
        outs.append([eng.memory_encode(*x["me"])[0].clone()])
    return outs


def test_two_threads_two_streams_one_engine():
    from sam2_opt_b200.engine import Engine
    from sam2_opt_b200.synth_weights import make_weights
    dev = torch.device("cuda", torch.cuda.current_device())
    eng = Engine(make_weights(0), dev)
    xs = [_inputs(dev, 11), _inputs(dev, 12)]
    rounds = 3
    want = [_work(eng, x, rounds) for x in xs]           # sequential, default stream
    torch.cuda.synchronize()

    got, errs = [None, None], []
    start = threading.Barrier(2)

    def run(i):
        try:
            s = torch.cuda.Stream(device=dev)
            s.wait_stream(torch.cuda.default_stream(dev))
            with torch.cuda.stream(s):
                start.wait()
                got[i] = _work(eng, xs[i], rounds)
            s.synchronize()
        except Exception as e:      # surfaced in the main thread below
            errs.append(e)

    th = [threading.Thread(target=run, args=(i,)) for i in range(2)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    torch.cuda.synchronize()
    for i in range(2):
        for step, (a, b) in enumerate(zip(got[i], want[i])):
            for u, v in zip(a, b):
                assert torch.equal(u, v), f"thread {i}, call {step}: result differs from the sequential run"
    eng.release()


def test_predictor_on_default_cuda_device_and_non_current_device():
    """device="cuda" binds to the CURRENT device (under torchrun that is cuda:LOCAL_RANK, not cuda:0); single-kernel front
    ends launch on the device of their tensors whatever the caller's current device is."""
    from sam2_opt_b200 import ops, postprocess
    from sam2_opt_b200.engine import Engine, normalise_device
    from sam2_opt_b200.synth_weights import make_weights
    assert normalise_device("cuda").index == torch.cuda.current_device()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs for the non-current-device half")
    w = make_weights(0)
    with torch.cuda.device(1):
        e = Engine(w)                                    # no device given -> cuda:1
        assert e.device.index == 1
        e.release()
    assert torch.cuda.current_device() == 0
    u8 = torch.randint(0, 256, (1, 1024, 1024, 3), dtype=torch.uint8)
    a = ops.preprocess_u8(u8.to("cuda:1"))               # cuda:0 is current
    b = ops.preprocess_u8(u8.to("cuda:0"))
    assert a.device.index == 1 and torch.equal(a.cpu(), b.cpu())
    low = torch.randn(2, 1, 256, 256)
    ma = postprocess.upsample_masks(low.to("cuda:1"), (540, 960))
    mb = postprocess.upsample_masks(low.to("cuda:0"), (540, 960))
    assert torch.equal(ma.cpu(), mb.cpu())
    la, ca = postprocess.get_connected_components((low > 0).to("cuda:1"))
    lb, cb = postprocess.get_connected_components((low > 0).to("cuda:0"))
    assert torch.equal(la.cpu(), lb.cpu()) and torch.equal(ca.cpu(), cb.cpu())
