"""CPU: the host logic of predictor._GraphedTrack (device rings, control vector, counts) against sam_heads.assemble_memory — for
every frame after the prompt, the (memory row, position row) pairs and the (pointer token, pointer position) pairs the graph body
hands to E3, restricted to the counts it announces, must be exactly the set the reference's selection rule produces
(ref: sam2_base_official.py:825-950).  The engine is replaced by recorders; no GPU, no CUDA library."""
import types

import pytest
import torch


@pytest.mark.parametrize("t_cond", [0, 3])
def test_rings_and_control_vector_reproduce_the_memory_bank(monkeypatch, t_cond):
    monkeypatch.setattr(torch.Tensor, "pin_memory", lambda self, *a, **k: self)
    monkeypatch.setattr(torch.cuda, "Event", lambda *a, **k: types.SimpleNamespace(record=lambda *a: None, synchronize=lambda: None))
    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: None)
    from sam2_opt_b200 import predictor as P, sam_heads as H
    from sam2_opt_b200.synth_weights import make_weights
    w = make_weights(0, only_prefix=("maskmem", "obj_ptr", "no_obj"))
    n, T = 2, t_cond + 26
    g = torch.Generator().manual_seed(0)
    cap, frame_out = {}, {}

    def memory_attention(curr, mem, curr_pos, pos, ptr_tok, ppos, counts_dev=None):
        cap.update(mem=mem.clone(), pos=pos.clone(), ptr=ptr_tok.clone(), ppos=ppos.clone(), counts=counts_dev.clone())
        return torch.zeros(4096, n, 256)

    engine = types.SimpleNamespace(position_encoding=lambda c, s, k: torch.randn(k, c, s, s, generator=g), workspace_generation=lambda: 0,
                                   memory_attention=memory_attention,
                                   sam_heads=lambda pix, hr, **kw: (None, None, None, torch.zeros(n, 1, 256, 256), None, frame_out[cap["t"]]["obj_ptr"], torch.ones(n, 1)))
    vp = types.SimpleNamespace(device=torch.device("cpu"), w=w, engine=engine, _tpos_cache={},
                               _encode_memory_low=lambda f2, low, obj, fp: (frame_out[cap["t"]]["maskmem_features"], frame_out[cap["t"]]["maskmem_pos"]))
    pe = torch.randn(1, 64, 64, 64, generator=g)
    rec = lambda: {"maskmem_features": torch.randn(n, 64, 64, 64, generator=g).to(torch.bfloat16), "maskmem_pos": pe.expand(n, -1, -1, -1),
                   "obj_ptr": torch.randn(n, 256, generator=g)}
    cond, non_cond = {t_cond: rec()}, {}
    tr = P._GraphedTrack(vp, n)
    sig = lambda m, p: sorted((round(float(m[i].sum()), 2), round(float(p[i].sum()), 2)) for i in range(m.shape[0]))
    for t in range(t_cond + 1, T):
        frame_out[t] = rec()
        if tr.next_t != t or t == t_cond + 9:          # also re-bind in the middle of the video (a tracker picked up from stored records)
            tr.bind("owner", T, t_cond, cond[t_cond], {tt: non_cond[tt] for tt in range(max(t_cond + 1, t - 15), t)}, t)
        cap["t"], tr.warm = t, False                    # the eager body every time: there is no graph on the CPU
        tr.step(t, torch.zeros(1, 256, 64, 64), [torch.zeros(1, 32, 256, 256), torch.zeros(1, 64, 128, 128)])
        non_cond[t] = frame_out[t]
        mem, mpos, ptrs, opos = H.assemble_memory(w, t, cond, {k: v for k, v in non_cond.items() if k < t}, T, cache={})
        L, Pn = int(cap["counts"][0, 0]), int(cap["counts"][1, 0])
        assert (L, Pn) == (mem.shape[0], ptrs.shape[0]), (t, L, Pn, mem.shape, ptrs.shape)
        assert torch.equal(cap["counts"][:, 0], cap["counts"][:, 1])
        assert sig(cap["mem"][:L], cap["pos"][:L]) == sig(mem, mpos), t
        assert sig(cap["ptr"][:Pn], cap["ppos"][:Pn]) == sig(ptrs, opos), t
