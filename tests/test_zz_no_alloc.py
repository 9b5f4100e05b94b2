"""GPU: the allocation-free entry point `sam2b_op_resize_frames_static` (include/sam2b.h; no counterpart in the reference)
against the existing `sam2b_op_resize_frames`, both through the C ABI.

The resize kernels are bit-exact against the oracle whatever memory holds their workspace, so the tolerance here is the one
the resize tests use: bit-identical.  Two calls in a row share the static scratch array (the second one finds the first
one's contents in it); a problem beyond the array's size must be rejected with a clear error and no launch.
This file is named to be collected last."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_BYTES = 16 << 20          # SAM2B_RESIZE_STATIC_MAX_BYTES


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(lib, src, u8, n, h, w, oh, ow, normalise):
    out = torch.empty((n, 3, oh, ow), device=src.device, dtype=torch.float32)
    ws = torch.empty(int(lib.sam2b_op_resize_frames_workspace(n, h, w, oh, ow)), device=src.device, dtype=torch.uint8)
    assert lib.sam2b_op_resize_frames(_ptr(src), int(u8), n, h, w, _ptr(out), oh, ow, int(normalise), _ptr(ws), _stream()) == 0
    return out


def test_static_scratch_resize_matches_the_workspace_entry_point_twice_in_a_row():
    from sam2_opt_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(7)
    # call 1: one 1080p uint8 frame (the case the 16 MiB limit is sized for); call 2: two smaller fp32 frames, up-scaled in
    # one axis and down-scaled in the other, normalised - a different layout of the same scratch array
    a = torch.randint(0, 256, (1, 1080, 1920, 3), generator=g, dtype=torch.uint8).to(dev)
    b = torch.rand((2, 3, 540, 1280), generator=g).to(dev)
    cases = [(a, True, 1, 1080, 1920, False), (b, False, 2, 540, 1280, True)]
    want = [_existing(lib, src, u8, n, h, w, 1024, 1024, norm) for src, u8, n, h, w, norm in cases]
    outs = [torch.empty((n, 3, 1024, 1024), device=dev, dtype=torch.float32) for _, _, n, _, _, _ in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev)       # the calls below must not allocate through torch either
    for (src, u8, n, h, w, norm), out in zip(cases, outs):
        assert int(lib.sam2b_op_resize_frames_workspace(n, h, w, 1024, 1024)) <= MAX_BYTES
        rc = lib.sam2b_op_resize_frames_static(_ptr(src), int(u8), n, h, w, _ptr(out), 1024, 1024, int(norm), _stream())
        assert rc == 0, lib.sam2b_last_error().decode()
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated(dev) == before
    for got, ref in zip(outs, want):
        assert torch.isfinite(got).all()
        assert torch.equal(got, ref)
    # and once more in the other order: results do not depend on what the previous call left in the array
    for i in (1, 0):
        src, u8, n, h, w, norm = cases[i]
        outs[i].zero_()
        assert lib.sam2b_op_resize_frames_static(_ptr(src), int(u8), n, h, w, _ptr(outs[i]), 1024, 1024, int(norm), _stream()) == 0
    torch.cuda.synchronize()
    for got, ref in zip(outs, want):
        assert torch.equal(got, ref)


def test_static_scratch_resize_rejects_problems_beyond_its_array():
    from sam2_opt_b200 import _lib
    lib = _lib.load()
    dev = torch.device("cuda:0")
    n, h, w = 2, 1080, 1920
    assert int(lib.sam2b_op_resize_frames_workspace(n, h, w, 1024, 1024)) > MAX_BYTES
    src = torch.zeros((n, h, w, 3), device=dev, dtype=torch.uint8)
    out = torch.full((n, 3, 1024, 1024), -7.0, device=dev)
    rc = lib.sam2b_op_resize_frames_static(_ptr(src), 1, n, h, w, _ptr(out), 1024, 1024, 0, _stream())
    assert rc != 0
    msg = lib.sam2b_last_error().decode()
    assert "static array holds" in msg and str(MAX_BYTES) in msg, msg
    torch.cuda.synchronize()
    assert (out == -7.0).all()          # nothing was launched
