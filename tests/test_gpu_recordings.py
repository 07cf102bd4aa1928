"""GPU tests on whole stretches of the reference's example recordings and on a long synthetic sequence: the velocity
OUTPUT of the drop-in pipeline against (a) the reference's own functions run on the same frames, (b) the encoder log
the 116.4 mm/min recording ships, (c) the exact motion of a synthetic sequence.

Fixtures (tests/golden/clips/<name>/, made by tools/make_clip_fixtures.py + tools/make_golden_clips.py in the build
container, where /root/reference exists):
  clip.h264 / clip_tstamps.txt   2040 consecutive frames of the stretch where the filament moves, cut losslessly at an
                                 IDR frame, with their timecode lines
  reference_series.npz           rows / positions of the reference's own end_process (estimate_vel_vid.py:493-583) on
                                 those frames after CLAHE + non-local means (its default order, :703-709)
  reference_pairs.npz            the reference's calc_ECC_transform (:306-369) on every 5th consecutive cleaned pair
  encoder_series.npz (116.4)     position vs time from the reference's encoder_velocity (:417-487) on the full log
tests/golden/synth_subpixel*.npz the reference's end_process on two synthetic sequences (tools/make_golden_synth.py)

What the numbers mean is written up in DESIGN.md section 1 ("Agreement with the reference on recordings").
"""
import json
import os
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLIPS = os.path.join(ROOT, "tests", "golden", "clips")
MANIFEST = json.load(open(os.path.join(CLIPS, "MANIFEST.json")))
NAMES = ("29.1", "58.2", "116.4")

_cache = {}


def clip_frames(name):
    """(raw, clean) frames of a clip: decoded here, CLAHE + NLM on the GPU (the drop-in's default pre-processing)."""
    if name not in _cache:
        from motion_estimation_b200 import api, ingest
        m = MANIFEST[name]
        raw = ingest.read_frames(os.path.join(CLIPS, name, "clip.h264"), 0, m["frames"])
        if raw.shape[0] != m["frames"] or zlib.crc32(raw.tobytes()) != m["crc32_decoded"]:
            pytest.skip(f"this box's H.264 decoder does not reproduce the frames the golden series were made from ({name})")
        clean = api.denoise_hf5(api.adaptive_histogram_equalization(raw), n_list=len(raw) + 1)
        api.close_engines()
        _cache[name] = (raw, clean)
    return _cache[name]


@pytest.mark.parametrize("name", NAMES)
def test_preprocessing_reproduces_the_frames_the_reference_ran_on(me, name):
    """CLAHE + non-local means of 2040 real frames on the GPU = the frames the golden series were computed from (those
    were checked against cv2's own CLAHE / fastNlMeansDenoising[Multi] at five positions of each clip when the golden
    was made; tests/test_gpu_clahe.py and test_gpu_nlm.py compare with cv2 directly)."""
    _, clean = clip_frames(name)
    assert zlib.crc32(clean.tobytes()) == MANIFEST[name]["crc32_clean"]


# share of the reference's converged pairs the GPU search must match within 0.3 half-resolution pixels, and the median
# |difference| allowed: measured 0.993 / 0.929 / 0.805 and 0.061 / 0.083 / 0.114 px at the drop-in's default configuration
# (0.965 / 0.838 / 0.755 with the library's texture gate of 2 grey levels).  The misses are pairs across a frame-drop gap
# (40+ px of motion): non-local means has smeared those frames over 9 neighbours, the per-block minima scatter and the
# median falls short of the global ECC fit.
PAIR_BARS = {"29.1": (0.98, 0.08), "58.2": (0.91, 0.10), "116.4": (0.79, 0.13)}


@pytest.mark.parametrize("name", NAMES)
def test_pair_displacements_vs_reference_ecc_on_cleaned_frames(me, name):
    _, clean = clip_frames(name)
    g = np.load(os.path.join(CLIPS, name, "reference_pairs.npz"))
    ks = g["k"]
    from motion_estimation_b200 import api
    cfg = api.default_config                                                 # the drop-in default
    with me.MotionEstimator(clean.shape[1], clean.shape[2], cfg, max_frames=len(clean), max_pairs=len(ks)) as est:
        est.upload(clean)
        rec = est.search([(int(k), int(k) + 1) for k in ks])
    conv = g["flag"] == 1
    diff = np.abs(rec["pdx"] * np.float32(0.5) - g["pdx"])[conv]
    share, med = float(np.mean(diff < 0.3)), float(np.median(diff))
    print(f"{name}: {conv.sum()} converged reference pairs, {share:.3f} within 0.3 px, median |diff| {med:.3f} px")
    assert conv.sum() >= 0.9 * len(ks)
    assert share >= PAIR_BARS[name][0] and med <= PAIR_BARS[name][1], (name, share, med)
    assert rec["flag"].all()


@pytest.mark.parametrize("name", NAMES)
def test_end_process_vs_reference_end_process(me, name):
    """The whole pair loop on 2040 cleaned frames against the reference's own end_process on the same frames, in the
    reference's unit.  The reference's ECC fails on 98 / 114 / 138 of these pairs and still accumulates the garbage
    (rows from -295 to +286 mm/min, a 26 mm 'travel' on the 58.2 stretch), so the comparison is on robust statistics
    of the velocity rows and on the travel over the reference's plausible rows (0 <= vx < 200 mm/min)."""
    from motion_estimation_b200 import api
    _, clean = clip_frames(name)
    ref = np.load(os.path.join(CLIPS, name, "reference_series.npz"))
    res = api.end_process(clean, imagePath=os.path.join(CLIPS, name), firstFrame=0, numberList=list(range(len(clean) + 1)),
                          return_details=True)
    api.close_engines()
    assert res.failed_estimates == 0
    ours_v = np.array([r[0] for r in res.out_information])
    ref_v = ref["out_information"][:, 0]
    assert 0.85 * len(ref_v) <= len(ours_v) <= 1.15 * len(ref_v)                       # about the same pairs get accepted
    q_ours, q_ref = np.percentile(ours_v, [25, 50, 75]), np.percentile(ref_v, [25, 50, 75])
    print(f"{name}: velocity rows 25/50/75 % ours {np.round(q_ours, 2)} reference {np.round(q_ref, 2)}")
    assert np.all(np.abs(q_ours - q_ref) <= 0.16 * q_ref), (name, q_ours, q_ref)       # measured: within 14 / 4 / 3 %
    ok = (ref_v >= 0) & (ref_v < 200)
    ref_travel = float(np.diff(np.concatenate([[0.0], ref["motion"]]))[ok].sum())
    ours_travel = res.motion_list_opt[-1]
    print(f"{name}: travel ours {ours_travel:.3f} mm, reference over its plausible rows {ref_travel:.3f} mm")
    # measured -5 % / -13 % / -9 %: the block median under-reads the smeared frames of the fast stretches (see above)
    assert 0.84 * ref_travel <= ours_travel <= 1.05 * ref_travel, (name, ours_travel, ref_travel)
    # no diverged rows: nothing negative, nothing beyond what a 64 px jump in one 5.5 ms frame interval can give (the
    # reference's rows on these stretches run from -295 to +286 mm/min)
    assert ours_v.min() >= 0 and ours_v.max() < (32 / 611) / (5.5e-3 / 60)


def test_end_process_vs_encoder_ground_truth(me):
    """The 116.4 mm/min recording ships the encoder log of the filament drive.  Travel over the excerpt's 28.3 s: the
    encoder says 54.2 mm.  Optical travel needs the frames the camera DROPPED (37 % of the time lies inside gaps of
    50-450 ms, in which the filament moves further than a +-64 px window can see), so it is lower by construction; what
    is pinned here is that the drop-in on the decoded frames (fast path, full-resolution unit) recovers more of it than
    the reference's pipeline does, and that a wider search recovers more still."""
    from motion_estimation_b200 import api
    raw, clean = clip_frames("116.4")
    d = os.path.join(CLIPS, "116.4")
    enc = np.load(os.path.join(d, "encoder_series.npz"))
    ref = np.load(os.path.join(d, "reference_series.npz"))
    travel = {}
    for label, frames, rng in (("raw64", raw, 64), ("raw128", raw, 128), ("clean64", clean, 64)):
        res = api.end_process(frames, imagePath=d, firstFrame=0, numberList=list(range(len(frames) + 1)),
                              config=me.SearchConfig(16, rng), units="fullres", return_details=True)
        t = np.array(res.tsList)
        travel[label] = res.motion_list_opt[-1] - res.motion_list_opt[0]
        enc_travel = float(np.interp(t[-1], enc["t_s"], enc["position_mm"]) - np.interp(t[0], enc["t_s"], enc["position_mm"]))
    api.close_engines()
    ref_travel_fullres = 2.0 * float(ref["motion"][-1])                                # the reference's unit is half-resolution px
    print(f"116.4: encoder {enc_travel:.1f} mm; optical, full-resolution unit: {travel}; reference x2: {ref_travel_fullres:.1f} mm")
    assert 50.0 < enc_travel < 58.0
    assert travel["raw64"] > travel["clean64"] and travel["raw128"] >= travel["raw64"]
    assert travel["raw64"] > ref_travel_fullres


def _synthetic_run(me, golden_file, rng):
    """GPU end_process on a synthetic sequence; returns (relative error of the sequence-mean velocity vs the exact
    motion, the same for the reference's own end_process or None, per-pair errors in half-resolution px)."""
    from motion_estimation_b200 import api, synth
    g = np.load(os.path.join(ROOT, "tests", "golden", golden_file))
    n, H, W, dt = int(g["n"]), int(g["h"]), int(g["w"]), float(g["dt_ms"])
    kw = dict(step_lo=float(g["step_lo"]), step_hi=float(g["step_hi"])) if "step_lo" in g.files else {}
    frames, x, _ = synth.subpixel_sequence(n, H, W, int(g["seed"]), **kw)
    stamps = [dt * k for k in range(n + 3)]
    res = api.end_process(frames, firstFrame=0, timestamps_ms=stamps, config=me.SearchConfig(16, rng), return_details=True)
    api.close_engines()
    ap, d = np.array(res.accepted_pairs), np.array(res.pair_displacements)
    err = d[:, 0] - (x[ap[:, 1]] - x[ap[:, 0]]) / 2
    ts, pos = np.array(res.tsList), np.array(res.motion_list_opt)
    v = (pos[-1] - pos[0]) / ((ts[-1] - ts[0]) / 60)
    v_true = ((x[ap[-1, 1]] - x[ap[0, 1]]) / 2 / 611) / ((ts[-1] - ts[0]) / 60)
    rel_ref = None
    if zlib.crc32(frames.tobytes()) == int(g["crc32_frames"]):                           # the golden was made from these very frames
        assert int(g["ref_failed"]) == 0
        rts, rpos = g["ref_ts"], g["ref_motion"]
        v_ref = (rpos[-1] - rpos[0]) / ((rts[-1] - rts[0]) / 60)
        k0, k1 = int(round(rts[0] * 1000 / dt)) - 1, int(round(rts[-1] * 1000 / dt)) - 1
        rel_ref = v_ref / (((x[k1] - x[k0]) / 2 / 611) / ((rts[-1] - rts[0]) / 60)) - 1
    return v / v_true - 1, rel_ref, err


def test_velocity_accuracy_synthetic_every_pair_accepted(me):
    """north_star's velocity figure (<= 0.1 %), where it can be checked: 400 clean 1280x128 frames translating 4.5-9.5 px
    per frame, so that every consecutive pair passes the reference's gate and the sub-pixel phases of the measured
    displacements are uniform.  Sequence-mean velocity of the GPU end_process against the exact motion: 0.1 %."""
    rel, rel_ref, err = _synthetic_run(me, "synth_subpixel_fast.npz", 16)
    print(f"synthetic, every pair accepted: {len(err)} pairs, per-pair error mean {err.mean():+.4f} std {err.std():.4f} max "
          f"{np.abs(err).max():.4f} half-res px; sequence-mean velocity vs exact {rel:+.5f}; reference's end_process {rel_ref}")
    assert np.abs(err).max() < 0.12 and abs(rel) <= 1e-3
    if rel_ref is not None:
        assert abs((1 + rel) / (1 + rel_ref) - 1) <= 2e-3


def test_velocity_accuracy_synthetic_gated(me):
    """The same with slow motion (0.6-2.4 px per frame, 1000 frames): now the reference's gate (accept the first frame whose
    displacement reaches 2 half-resolution px) picks displacements just above an integer, where a 3-point fit on a SAD
    surface leans towards that integer -- a selection effect, not noise: measured -0.22 % for the block search and +0.14 % for
    the reference's own ECC loop on the same frames.  Stated tolerance here: 0.3 % against the exact motion, 0.5 % between
    the two (north_star's 0.1 % is met only without the selection effect, see the test above)."""
    rel, rel_ref, err = _synthetic_run(me, "synth_subpixel.npz", 16)
    print(f"synthetic, gated: {len(err)} pairs, per-pair error mean {err.mean():+.4f} std {err.std():.4f} half-res px; "
          f"sequence-mean velocity vs exact {rel:+.5f}; reference's end_process {rel_ref}")
    assert np.abs(err).max() < 0.12 and abs(rel) <= 3e-3
    if rel_ref is not None:
        assert abs((1 + rel) / (1 + rel_ref) - 1) <= 5e-3


def test_dense_field_mean_vs_reference_farneback(me):
    """a5 / f3: the frame mean of the dense per-block field (part of every pair record) against the reference's
    calculate_dense_flow (Farneback mean, estimate_vel_vid.py:385-391) on synthetic pairs where Farneback converges."""
    from motion_estimation_b200 import api, synth
    g = np.load(os.path.join(ROOT, "tests", "golden", "synth_subpixel.npz"))
    n, H, W = int(g["n"]), int(g["h"]), int(g["w"])
    frames, x, y = synth.subpixel_sequence(n, H, W, int(g["seed"]))
    if zlib.crc32(frames.tobytes()) != int(g["crc32_frames"]):
        pytest.skip("synthetic frames differ from the ones the golden was made from (cv2 build differs)")
    worst = 0.0
    for (a, b), (fx, fy) in zip(g["flow_pairs"], g["ref_flow"]):
        mx, my = api.calculate_dense_flow(frames[a], frames[b], config=me.SearchConfig(16, 16))
        true_dx = x[b] - x[a]
        worst = max(worst, abs(mx - fx))
        assert abs(mx - true_dx) < 0.2 and abs(my - (y[b] - y[a])) < 0.2, (a, b, mx, my, true_dx)     # measured <= 0.17 px
        assert abs(mx - fx) < 0.25 and abs(my - fy) < 0.25, (a, b, mx, fx, my, fy)
    print(f"dense field mean vs Farneback mean: worst |difference| {worst:.3f} px over {len(g['flow_pairs'])} pairs")
    api.close_engines()
