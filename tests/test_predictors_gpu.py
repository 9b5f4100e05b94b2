"""GPU: the standalone predictors (reference API, engine underneath) against the reference's own predictor outputs
(golden fixtures made by the unmodified reference) and the fp32 oracle.  Gate (BASELINE.json north_star): mask IoU >= 0.99."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def iou_bits(a_bool, b_bool):
    u = (a_bool | b_bool).sum()
    return 1.0 if u == 0 else float((a_bool & b_bool).sum()) / float(u)


@pytest.fixture(scope="module")
def weights():
    from oracle import sam2_oracle as O
    return O.make_weights(0)


def test_config1_set_image_predict_iou(weights, golden):
    from oracle.synth import synth_image_u8
    from sam2_opt_b200.predictor import SAM2ImagePredictor
    p = SAM2ImagePredictor(weights)
    p.set_image(synth_image_u8(0))
    masks, iou, low = p.predict(point_coords=np.array([[512, 512]]), point_labels=np.array([1]), multimask_output=True, return_logits=True)
    ref_bits = np.unpackbits(golden["c1_low_res_bits"]).reshape(3, 256, 256).astype(bool)
    ious = [iou_bits(low[i] > 0, ref_bits[i]) for i in range(3)]
    print("config1 low-res mask IoU vs reference:", [f"{v:.4f}" for v in ious], "iou_pred", iou, "ref", golden["c1_iou"])
    assert min(ious) >= 0.99
    assert np.abs(iou - golden["c1_iou"]).max() < 2e-2
    area = (masks > 0).sum(axis=(1, 2))
    assert np.all(np.abs(area - golden["c1_mask_area"][0]) <= 0.02 * golden["c1_mask_area"][0] + 50)
    with pytest.raises(Exception, match="unsupported runtime backend"):
        p.speedup("tensorrt")
    p.speedup("b200")          # re-snapshot works
    p.release(); p.release()   # idempotent


def test_config3_video_vs_reference_golden(weights, golden):
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    vp = SAM2VideoPredictor(weights)
    st = vp.init_state(synth_video(4, seed=1))
    vp.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
    ious = []
    for t, ids, m in vp.propagate_in_video(st):
        low = (st["cond"][t][1] if t in st["cond"] else st["non_cond"][1][t])["pred_masks"][0, 0]
        ref = np.unpackbits(golden["c3_low_res_bits"][t]).reshape(256, 256).astype(bool)
        ious.append(iou_bits((low > 0).cpu().numpy(), ref))
        assert m.shape == (1, 1, 1024, 1024) and ids == [1]
    print("config3 per-frame IoU vs reference:", [f"{v:.4f}" for v in ious])
    assert min(ious) >= 0.99
    vp.release()


def test_video_full_bank_vs_oracle_and_object_batching(weights):
    """9 frames (the 7-frame bank fills), two objects batched in one engine call per frame; object 1 must match the fp32
    oracle's single-object run frame by frame, and batching must not change an object's result."""
    from oracle import sam2_oracle as O
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    T = 9
    frames = synth_video(T, seed=2)
    vp = SAM2VideoPredictor(weights)
    st = vp.init_state(frames)
    vp.add_new_points_or_box(st, 0, 1, points=np.array([[512, 512]], dtype=np.float32), labels=np.array([1], np.int32))
    vp.add_new_points_or_box(st, 0, 2, points=np.array([[300, 700]], dtype=np.float32), labels=np.array([1], np.int32))
    got = {t: m for t, _, m in vp.propagate_in_video(st)}
    st1 = vp.init_state(frames)
    vp.add_new_points_or_box(st1, 0, 2, points=np.array([[300, 700]], dtype=np.float32), labels=np.array([1], np.int32))
    alone = {t: m for t, _, m in vp.propagate_in_video(st1)}
    with torch.no_grad():
        ref = {t: low for t, low in O.video_propagate(weights, frames)}
    ious, ious_b = [], []
    for t in range(T):
        low = (st["cond"][t][1] if t in st["cond"] else st["non_cond"][1][t])["pred_masks"][0, 0].cpu()
        ious.append(O.mask_iou(low, ref[t][0, 0]))
        ious_b.append(iou_bits((got[t][1, 0] > 0).cpu().numpy(), (alone[t][0, 0] > 0).cpu().numpy()))
    print("9-frame IoU vs oracle:", [f"{v:.4f}" for v in ious])
    print("batched-vs-alone IoU:", [f"{v:.4f}" for v in ious_b])
    assert min(ious) >= 0.99
    assert min(ious_b) >= 0.995
    vp.release()


def test_video_fill_hole_area():
    """fill_hole_area post-processes the stored / returned low-res scores (reference
    sam2_video_predictor_official.py:889-894).  The prompted frame, which depends on nothing else, must equal the plain
    run's frame passed through the oracle's hole filling bit for bit; every later frame (whose memory was encoded from
    the filled prompted frame, as in the reference's preflight) must be hole-free up to that area."""
    import numpy as np
    from oracle import postprocess_oracle as PO
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    from sam2_opt_b200.synth import synth_video
    from sam2_opt_b200.synth_weights import make_weights
    w = make_weights(0)
    frames = synth_video(4, seed=2)
    outs = {}
    for area in (0, 8):
        vp = SAM2VideoPredictor(w, fill_hole_area=area)
        st = vp.init_state(frames)
        vp.add_new_points_or_box(st, 0, 1, points=[[512, 512]], labels=[1])
        list(vp.propagate_in_video(st))
        lows = [st["cond"][0][1]["pred_masks"]] + [st["non_cond"][1][t]["pred_masks"] for t in (1, 2, 3)]
        outs[area] = [x.cpu().numpy() for x in lows]
        vp.release()
    want0 = PO.fill_holes_in_mask_scores(outs[0][0], 8)
    assert np.array_equal(outs[8][0], want0)
    print("pixels changed by hole filling on the prompted frame:", int((want0 != outs[0][0]).sum()))
    for filled in outs[8]:
        assert np.array_equal(PO.fill_holes_in_mask_scores(filled, 8), filled)


def test_video_batched_bank_and_graphed_heads_equal_the_plain_path(weights):
    """The per-frame batched memory-bank assembly (one group of objects sharing their history) and the CUDA-graph replay
    of the SAM heads are pure scheduling changes: three objects must give bit-identical low-res masks with both switched
    off (per-object assembly + eager heads).  Also covers the fallback when objects do NOT share their history (one
    object is added after the first propagation, so the groups differ) - it must run and track every object."""
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    T = 10
    frames = synth_video(T, seed=4)
    pts = {1: [[512, 512]], 2: [[300, 700]], 3: [[700, 300]]}

    def run(plain, ring=False):
        vp = SAM2VideoPredictor(weights, graph_heads=not plain, bank_ring=ring)
        st = vp.init_state(frames)
        for o, p in pts.items():
            vp.add_new_points_or_box(st, 0, o, points=np.array(p, dtype=np.float32), labels=np.array([1], np.int32))
        if plain:
            class _Never(dict):                      # never remembers a batched frame -> always the per-object path
                def setdefault(self, k, d=None):
                    return {}
                def get(self, k, d=None):
                    return {}
            st["non_cond_b"] = _Never()
        out = {t: m.clone() for t, _, m in vp.propagate_in_video(st)}
        return vp, st, out

    vp_a, st_a, a = run(plain=False)
    vp_b, st_b, b = run(plain=True)
    for t in range(T):
        assert torch.equal(a[t], b[t]), f"frame {t}: batched/graphed path differs from the plain path"
    # the device-resident bank ring (default) holds the same values in a rotating slot order: the attention sums the keys in a
    # different order, so the logits agree to fp32 / bf16 round-off instead of bit for bit
    vp_r, st_r, r = run(plain=False, ring=True)
    assert st_r["rings"] and all(len(ring.frames) == min(T - 1, 7) for ring in st_r["rings"].values())
    for t in range(T):
        for i in range(3):
            v = iou_bits((r[t][i, 0] > 0).cpu().numpy(), (a[t][i, 0] > 0).cpu().numpy())
            assert v >= 0.999, f"frame {t} object {i + 1}: IoU {v:.4f} between the ring and the re-assembled bank"
    vp_r.release()
    assert len(st_a["non_cond_b"]) == T - 1 and all(tuple(k) == (1, 2, 3) for d in st_a["non_cond_b"].values() for k in d)
    vp_a.release(); vp_b.release()
    # mixed history (7 frames, so that two conditioning frames + recent frames stay within the 7-slot bank of the E3
    # boundary): tracking stops after frame 2, then object 3 gets a prompt on the not yet tracked frame 3, which becomes a
    # second conditioning frame for it (a click on an already tracked frame would be a correction stored as a tracked
    # output, see test_refinement_and_correction_clicks_match_the_reference).  Its conditioning frames now differ from
    # objects 1-2, the frame groups split ((1, 2) and (3,)) and no batched record matches: the per-object fallback runs.
    # Objects 1-2 keep their history, so their masks stay the same up to the summation order of a batch of 2 instead of 3.
    T2 = 7
    vp_a = SAM2VideoPredictor(weights)
    st_f = vp_a.init_state(frames[:T2])
    for o, p in pts.items():
        vp_a.add_new_points_or_box(st_f, 0, o, points=np.array(p, dtype=np.float32), labels=np.array([1], np.int32))
    first = {t: m.clone() for t, _, m in vp_a.propagate_in_video(st_f)}
    st_a = vp_a.init_state(frames[:T2])
    for o, p in pts.items():
        vp_a.add_new_points_or_box(st_a, 0, o, points=np.array(p, dtype=np.float32), labels=np.array([1], np.int32))
    assert [t for t, _, _ in vp_a.propagate_in_video(st_a, max_frame_num_to_track=2)] == [0, 1, 2]
    vp_a.add_new_points_or_box(st_a, 3, 3, points=np.array([[705, 305]], dtype=np.float32), labels=np.array([1], np.int32))
    assert 3 in st_a["cond"] and list(st_a["cond"][3]) == [3]
    again = {t: m for t, _, m in vp_a.propagate_in_video(st_a, start_frame_idx=3)}
    assert sorted(again) == list(range(3, T2)) and all(m.shape[0] == 3 for m in again.values())
    for t in range(4, T2):
        for i in range(2):
            v = iou_bits((again[t][i, 0] > 0).cpu().numpy(), (first[t][i, 0] > 0).cpu().numpy())
            assert v >= 0.995, f"frame {t} object {i + 1}: IoU {v:.4f} after re-prompting object 3"
    assert all({(1, 2), (3,)} <= set(d) for t, d in st_a["non_cond_b"].items() if t >= 4)
    vp_a.release()


def test_refinement_and_correction_clicks_match_the_reference(weights):
    """Interactive use beyond the first click (reference add_new_points_or_box, sam2_video_predictor_official.py:267-399):
    (1) a second click on the prompted frame with clear_old_points=False feeds the previous mask logits (clamped to +-32)
        back into the SAM decoder as its mask prompt;
    (2) a click on an already TRACKED frame corrects the tracked mask with memory-conditioned features and is stored as a
        non-conditioning output, whose memory is encoded by the next preflight.
    Both against the unmodified reference's torch path in strict fp32 on the same device, step by step."""
    from oracle import reference_loader as rl
    if not rl.available():
        pytest.skip("reference sources not present")
    import ref_gpu
    from oracle.synth import synth_video
    from sam2_opt_b200.predictor import SAM2VideoPredictor
    torch.set_grad_enabled(False)
    T = 6
    frames = synth_video(T, seed=5)
    f32, i32 = np.float32, np.int32
    script = [
        ("click", 0, dict(points=np.array([[512, 512]], f32), labels=np.array([1], i32))),
        ("click", 0, dict(points=np.array([[600, 430]], f32), labels=np.array([1], i32), clear_old_points=False)),
        ("propagate", None, {}),
        ("click", 3, dict(points=np.array([[400, 640]], f32), labels=np.array([0], i32))),
        ("propagate", 3, {}),
    ]

    def run(pred, st, is_ref):
        outs = []
        for kind, t, kw in script:
            if kind == "click":
                _, _, m = pred.add_new_points_or_box(st, t, 1, **kw)
                outs.append([(m[0, 0] > 0).cpu().numpy()])
            else:
                outs.append([(m[0, 0] > 0).cpu().numpy() for _, _, m in pred.propagate_in_video(st, start_frame_idx=t)])
        return outs

    ref_m = ref_gpu.build("video", weights)
    with ref_gpu.synthetic_video(frames):
        ref_st = ref_m.init_state("synthetic")
    want = run(ref_m, ref_st, True)
    # where the reference keeps the correction: a tracked (non-conditioning) output of frame 3
    assert 3 in ref_st["output_dict_per_obj"][0]["non_cond_frame_outputs"] and 3 not in ref_st["output_dict_per_obj"][0]["cond_frame_outputs"]
    del ref_m
    torch.cuda.empty_cache()
    vp = SAM2VideoPredictor(weights)
    st = vp.init_state(frames)
    got = run(vp, st, False)
    assert 3 in st["non_cond"][1] and 3 not in st["cond"]
    for step, (g, w_) in enumerate(zip(got, want)):
        assert len(g) == len(w_)
        ious = [iou_bits(a, b) for a, b in zip(g, w_)]
        print(f"step {step} ({script[step][0]} @ {script[step][1]}): IoU vs reference", [f"{v:.4f}" for v in ious])
        assert min(ious) >= 0.99, (step, ious)
    # the refinement click really changed the mask (otherwise step 1 would not test the mask-prompt path)
    assert iou_bits(want[0][0], want[1][0]) < 0.999
    vp.release()
