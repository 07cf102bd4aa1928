"""GPU parity tests: the CUDA path, called through the C ABI (ctypes), against the C oracle.

Integer displacements, SADs and flags must be bit-exact; the refined displacements and the pair
medians are float32 computed with the same correctly-rounded operations in the same order, so they
are compared for exact equality as well (tolerance 0; north_star allows 1e-3 px).
At the sizes BASELINE.json names (1080p, 4K) the oracle is too slow to run on every block, so those
cases use size-independent properties: known integer shifts recovered, batch-layout invariance, and
oracle equality on a sample of blocks via the cost surface."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import sad as osad   # noqa: E402  (checker only)

KEYS = ("mv", "sad", "sub", "flags")


def check_pair(est, idx, prev, cur, cfg, rec):
    o = osad.search(prev, cur, cfg.block, cfg.range, min_range=cfg.min_range, subpixel=cfg.subpixel,
                    min_valid=cfg.min_valid, min_curv=cfg.min_curv, range_y=cfg.range_y)
    g = est.fetch_blocks(idx)
    for k in KEYS:
        np.testing.assert_array_equal(g[k], o[k], err_msg=f"{k} (pair {idx})")
    assert (rec["pdx"], rec["pdy"], rec["flag"], rec["nvalid_x"], rec["nvalid_y"], rec["n_textured"], rec["mean_dx"], rec["mean_dy"]) == \
           (o["pdx"], o["pdy"], o["flag"], o["nvalid_x"], o["nvalid_y"], o["n_textured"], o["mean_dx"], o["mean_dy"])


CASES = [
    # H, W, B, R   (ragged frames, every block size, ranges with and without 16-byte aligned origins)
    (64, 96, 16, 8), (64, 640, 16, 32), (70, 100, 16, 8), (72, 200, 8, 12), (96, 160, 32, 16),
    (270, 480, 16, 32), (136, 256, 8, 64), (48, 64, 16, 1), (33, 47, 16, 5), (64, 128, 32, 40),
    (80, 144, 16, 50), (16, 16, 16, 4), (40, 336, 8, 7), (128, 128, 16, 96),
    (160, 192, 16, 100), (144, 304, 16, 127), (72, 136, 8, 90), (200, 168, 32, 72),   # ranges that need candidate slabs
    (144, 320, 16, 128), (80, 304, 8, 128),                                          # 17-bit candidate rank: 64-bit keys
]


@pytest.mark.parametrize("H,W,B,R", CASES)
def test_search_equals_oracle(me, H, W, B, R):
    from motion_estimation_b200 import synth
    n = 4
    frames, _ = synth.translating_frames(n, H, W, min(R, 24))
    cfg = me.SearchConfig(block=B, range=R)
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=n + 1) as est:
        est.upload(frames)
        pairs = [(0, 1), (1, 2), (2, 3), (3, 0), (2, 2)]
        res = est.search(pairs)
        for i, (a, c) in enumerate(pairs):
            check_pair(est, i, frames[a], frames[c], cfg, res[i])


RESIDENT_CASES = [(64, 640, 16, 32), (270, 480, 16, 32), (136, 256, 8, 64), (96, 160, 32, 16), (64, 128, 32, 40), (128, 128, 16, 96),
                  (160, 192, 16, 100), (72, 136, 8, 90), (200, 168, 32, 72), (144, 320, 16, 128), (80, 304, 8, 128), (40, 336, 8, 7)]


@pytest.mark.parametrize("H,W,B,R", RESIDENT_CASES)
def test_resident_kernel_equals_oracle(me, monkeypatch, H, W, B, R):
    """The warp-specialised one-CTA-per-SM kernel forced on small launches: single-slab, slabbed and 64-bit-key plans,
    every block size, shifted and unshifted window origins; twice, so tickets, queues and counters must be left clean."""
    from motion_estimation_b200 import synth
    monkeypatch.setenv("ME_PERSISTENT", "2")
    monkeypatch.setenv("ME_TMR", "1")                   # the resident kernel takes one block row per unit
    n = 4
    frames, _ = synth.translating_frames(n, H, W, min(R, 24))
    cfg = me.SearchConfig(block=B, range=R)
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=n + 1) as est:
        if not est.geometry["resident_kernel"]:
            pytest.skip("two stages of this plan do not fit in shared memory")
        est.upload(frames)
        pairs = [(0, 1), (1, 2), (2, 3), (3, 0), (2, 2)]
        for _ in range(2):
            res = est.search(pairs)
            for i, (a, c) in enumerate(pairs):
                check_pair(est, i, frames[a], frames[c], cfg, res[i])


@pytest.mark.parametrize("kw", [dict(subpixel=False), dict(min_range=60), dict(min_curv=400), dict(min_valid=10 ** 6)])
def test_gates_and_modes_equal_oracle(me, kw):
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(3, 96, 160, 10)
    cfg = me.SearchConfig(block=16, range=12, **kw)
    with me.MotionEstimator(96, 160, cfg, max_frames=3, max_pairs=2) as est:
        est.upload(frames)
        res = est.search([(0, 1), (1, 2)])
        for i in range(2):
            check_pair(est, i, frames[i], frames[i + 1], cfg, res[i])
    if "min_valid" in kw:
        assert res[0]["flag"] == 0 and res[0]["nvalid_x"] > 0


def test_ties_and_flat_frames_equal_oracle(me):
    """Tie-break order is where parallel reductions usually go wrong: constant and periodic frames."""
    flat = np.full((64, 96), 9, np.uint8)
    per = np.tile(np.arange(8, dtype=np.uint8) * 20, (64, 12))
    dark = np.zeros((64, 96), np.uint8)
    dark[20:30, 40:50] = 200
    frames = np.stack([flat, per, dark, np.roll(dark, 3, axis=1)])
    cfg = me.SearchConfig(block=16, range=12, min_range=0)
    with me.MotionEstimator(64, 96, cfg, max_frames=4, max_pairs=4) as est:
        est.upload(frames)
        pairs = [(0, 0), (1, 1), (2, 3), (0, 1)]
        res = est.search(pairs)
        for i, (a, c) in enumerate(pairs):
            check_pair(est, i, frames[a], frames[c], cfg, res[i])


def test_plain_loader_equals_tma_loader(me, monkeypatch):
    """ME_NO_TMA=1 stages the window with ordinary loads; both stagings must give identical bytes."""
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(3, 80, 176, 16)
    cfg = me.SearchConfig(block=16, range=20)
    out = []
    for flag in ("0", "1"):
        monkeypatch.setenv("ME_NO_TMA", flag)
        with me.MotionEstimator(80, 176, cfg, max_frames=3, max_pairs=2) as est:
            assert est.geometry["uses_tma"] == (flag == "0")
            est.upload(frames)
            res = est.search([(0, 1), (1, 2)])
            out.append((res.tobytes(), est.fetch_blocks(1)))
    assert out[0][0] == out[1][0]
    for k in KEYS:
        np.testing.assert_array_equal(out[0][1][k], out[1][1][k])


@pytest.mark.parametrize("nsx,nsy", [(1, 2), (2, 1), (2, 3), (1, 7)])
def test_candidate_slabs_do_not_change_results(me, monkeypatch, nsx, nsy):
    """Force the candidate square to be cut into several CTAs per tile (cross-CTA argmin through global keys)."""
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(3, 112, 240, 30)
    cfg = me.SearchConfig(block=16, range=36)
    monkeypatch.setenv("ME_NSX", str(nsx))
    monkeypatch.setenv("ME_NSY", str(nsy))
    with me.MotionEstimator(112, 240, cfg, max_frames=3, max_pairs=3) as est:
        assert (est.geometry["slabs_x"], est.geometry["slabs_y"]) == (nsx, nsy)
        est.upload(frames)
        pairs = [(0, 1), (1, 2), (2, 0)]
        for _ in range(2):                                  # twice: the global keys and tickets must be left clean
            res = est.search(pairs)
            for i, (a, c) in enumerate(pairs):
                check_pair(est, i, frames[a], frames[c], cfg, res[i])


@pytest.mark.parametrize("tbx", [1, 2, 4, 8])
def test_tile_width_does_not_change_results(me, monkeypatch, tbx):
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(2, 64, 208, 12)      # 13 blocks wide: ragged last tile for every tbx > 1
    cfg = me.SearchConfig(block=16, range=16)
    monkeypatch.setenv("ME_TBX", str(tbx))
    with me.MotionEstimator(64, 208, cfg, max_frames=2, max_pairs=1) as est:
        assert est.geometry["tile_blocks_x"] == tbx
        est.upload(frames)
        res = est.search([(0, 1)])
        check_pair(est, 0, frames[0], frames[1], cfg, res[0])


@pytest.mark.parametrize("B,R,tmr,H,W", [(16, 8, 2, 112, 240), (16, 8, 3, 112, 240), (16, 6, 4, 120, 208), (8, 5, 4, 88, 200),
                                         (8, 12, 5, 88, 200), (32, 9, 2, 160, 224), (16, 16, 7, 112, 144), (16, 20, 2, 64, 640)])
def test_tile_height_does_not_change_results(me, monkeypatch, B, R, tmr, H, W):
    """Tiles of several block rows (the plan for small ranges): rows at the top and bottom of the frame have
    their own, smaller dy range; the last tile row is ragged for most of these shapes."""
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(3, H, W, min(R, 12))
    cfg = me.SearchConfig(block=B, range=R)
    monkeypatch.setenv("ME_TMR", str(tmr))
    with me.MotionEstimator(H, W, cfg, max_frames=3, max_pairs=3) as est:
        assert est.geometry["tile_blocks_y"] == tmr
        est.upload(frames)
        pairs = [(0, 1), (1, 2), (2, 0)]
        for _ in range(2):
            res = est.search(pairs)
            for i, (a, c) in enumerate(pairs):
                check_pair(est, i, frames[a], frames[c], cfg, res[i])


def test_example_recordings_equal_oracle(me, golden_clips):
    """Every example sequence: integer displacements bit-exact against the oracle (north_star)."""
    cfg = me.SearchConfig(block=16, range=32)
    for name, g in golden_clips.items():
        f = g["frames"]
        with me.MotionEstimator(f.shape[1], f.shape[2], cfg, max_frames=len(f), max_pairs=64) as est:
            est.upload(f)
            pairs = [tuple(p) for p in g["pairs"]] + [(k, k + 1) for k in range(12, len(f) - 1)]
            res = est.search(pairs)
            for i, (a, c) in enumerate(pairs):
                check_pair(est, i, f[a], f[c], cfg, res[i])
            # and against the vectors committed with the excerpts
            for (a, c), want, r in zip(g["pairs"], g["sad_pairs"], res):
                assert [float(r["pdx"]), float(r["pdy"]), r["flag"], r["nvalid_x"], r["nvalid_y"]] == list(want)


def test_estimate_pair_and_bgr_ingest(me, golden_clips):
    f = golden_clips["58.2"]["frames"]
    cfg = me.SearchConfig(block=8, range=12)
    with me.MotionEstimator(64, 640, cfg, max_frames=4, max_pairs=2) as est:
        r = est.estimate_pair(f[8], f[9])
        o = osad.search(f[8], f[9], 8, 12)
        assert (r["pdx"], r["pdy"], r["flag"], r["nvalid_x"]) == (o["pdx"], o["pdy"], o["flag"], o["nvalid_x"])
        # BGR ingest: conversion on the GPU must equal cv.cvtColor, hence the same search result
        import cv2
        rng = np.random.default_rng(0)
        bgr = rng.integers(0, 256, (2, 64, 640, 3), dtype=np.uint8)
        bgr = np.stack([cv2.GaussianBlur(b, (0, 0), 2.0) for b in bgr])
        gray = np.stack([cv2.cvtColor(b, cv2.COLOR_BGR2GRAY) for b in bgr])
        est.upload_bgr(bgr, first_slot=2)
        a = est.search([(2, 3)])
        est.upload(gray, first_slot=0)
        b = est.search([(0, 1)])
        assert a.tobytes() == b.tobytes()
        np.testing.assert_array_equal(gray[0], osad.bgr2gray(bgr[0]))


def test_search_device_on_a_torch_tensor(me):
    import torch
    from motion_estimation_b200 import synth
    frames, _ = synth.translating_frames(3, 64, 128, 10)
    cfg = me.SearchConfig(block=16, range=12)
    t = torch.from_numpy(frames).cuda().contiguous()
    with me.MotionEstimator(64, 128, cfg, max_frames=2, max_pairs=2) as est:
        n = est.search_device(t.data_ptr(), t.stride(1), t.shape[0], [(0, 1), (1, 2)],
                              stream=torch.cuda.current_stream().cuda_stream)
        res = est.fetch_pairs(n)
        for i in range(2):
            check_pair(est, i, frames[i], frames[i + 1], cfg, res[i])


def test_errors_are_reported_not_swallowed(me):
    with me.MotionEstimator(64, 128, me.SearchConfig(16, 8), max_frames=2, max_pairs=2) as est:
        with pytest.raises(me.MEError):
            est.fetch_pairs(1)                              # fetch before search
        with pytest.raises(me.MEError):
            est.launch([(0, 5)])                            # slot out of range
        with pytest.raises(ValueError):
            est.upload(np.zeros((1, 32, 128), np.uint8))    # wrong frame height
    with pytest.raises(me.MEError) as e:
        me.MotionEstimator(64, 128, me.SearchConfig(16, 129))
    assert e.value.code == -2 and "range" in str(e.value)   # ranges up to 128 are built


# ----------------------------------------------------------------------------- BASELINE sizes
def _sample_blocks_equal_oracle(est, idx, prev, cur, B, R, blocks):
    g = est.fetch_blocks(idx)
    for bx, by in blocks:
        cs = osad.cost_surface(prev, cur, B, R, bx, by).astype(np.int64)
        rank = int(np.argmin(cs))
        U = 2 * R + 1
        assert tuple(g["mv"][by, bx]) == (rank % U - R, rank // U - R), (bx, by)
        assert g["sad"][by, bx] == cs.flat[rank]


def test_full_hd_properties(me):
    """configs[2]: 1920x1080, B=16, R=32.  Known integer shifts recovered exactly (integer mode),
    sampled blocks equal the oracle, and results do not depend on batch layout."""
    from motion_estimation_b200 import synth
    n = 5
    frames, shifts = synth.translating_frames(n, 1080, 1920, 32)
    cfg = me.SearchConfig(block=16, range=32, subpixel=False)
    with me.MotionEstimator(1080, 1920, cfg, max_frames=n, max_pairs=n) as est:
        est.upload(frames)
        pairs = [(k, k + 1) for k in range(n - 1)]
        res = est.search(pairs)
        keep = [est.fetch_blocks(i) for i in range(len(pairs))]
        for i, (dx, dy) in enumerate(shifts):
            assert res[i]["flag"] == 1 and res[i]["pdx"] == dx and res[i]["pdy"] == dy
            inner = keep[i]["mv"][3:-3, 3:-3]
            assert (inner[..., 0] == dx).all() and (inner[..., 1] == dy).all()
            assert (keep[i]["sad"][3:-3, 3:-3] == 0).all()
        _sample_blocks_equal_oracle(est, 0, frames[0], frames[1], 16, 32,
                                    [(0, 0), (119, 66), (1, 65), (60, 33), (118, 0), (0, 66), (59, 1)])
        # same pairs one at a time, in reverse order: identical records and blocks
        for i in reversed(range(len(pairs))):
            one = est.search([pairs[i]])
            assert one[0].tobytes() == res[i].tobytes()
            b = est.fetch_blocks(0)
            for k in KEYS:
                np.testing.assert_array_equal(b[k], keep[i][k])


def test_4k_properties(me):
    """configs[3]: 3840x2160, B=8, R=64."""
    from motion_estimation_b200 import synth
    frames, shifts = synth.translating_frames(3, 2160, 3840, 64)
    cfg = me.SearchConfig(block=8, range=64, subpixel=False)
    with me.MotionEstimator(2160, 3840, cfg, max_frames=3, max_pairs=2) as est:
        est.upload(frames)
        res = est.search([(0, 1), (1, 2)])
        for i, (dx, dy) in enumerate(shifts):
            assert res[i]["flag"] == 1 and res[i]["pdx"] == dx and res[i]["pdy"] == dy
        _sample_blocks_equal_oracle(est, 1, frames[1], frames[2], 8, 64, [(0, 0), (479, 269), (240, 135), (7, 262)])


@pytest.mark.parametrize("H,W,B,R,Ry", [(64, 640, 16, 64, 8), (64, 640, 16, 32, 4), (270, 480, 16, 32, 3), (136, 256, 8, 64, 10),
                                        (144, 320, 16, 128, 16), (96, 160, 32, 16, 15), (72, 200, 8, 12, 1), (720, 1280, 16, 8, 2),
                                        (160, 192, 16, 100, 40)])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_rectangular_range_equals_oracle(me, monkeypatch, H, W, B, R, Ry, mode):
    """range_y < range: only |dy| <= Ry is searched (the recordings move along x).  Same spec otherwise; default plan,
    resident kernel forced, one CTA per tile forced."""
    from motion_estimation_b200 import synth
    if mode:
        monkeypatch.setenv("ME_PERSISTENT", "2" if mode == 1 else "0")
    frames, _ = synth.translating_frames(4, H, W, min(2 * Ry, R, 20), seed_texture=H + Ry, seed_motion=W + R)
    cfg = me.SearchConfig(block=B, range=R, range_y=Ry)
    pairs = [(0, 1), (1, 2), (2, 3), (3, 0), (1, 1)]
    with me.MotionEstimator(H, W, cfg, max_frames=4, max_pairs=len(pairs)) as est:
        est.upload(frames)
        res = est.search(pairs)
        for i in (0, 3, 4) if H >= 700 else range(len(pairs)):
            check_pair(est, i, frames[pairs[i][0]], frames[pairs[i][1]], cfg, res[i])
        one = est.search([pairs[3]])                      # a one-pair launch takes the small-batch plan
        assert one[0].tobytes() == res[3].tobytes()


def test_searches_in_flight_and_ticketed_fetch(me):
    """Several searches queued before any records are fetched: every ticket returns the records of its own search, whatever is
    launched after it (four ring slots), and the plain fetch still means the last search."""
    from motion_estimation_b200 import synth
    H, W, n = 144, 320, 7
    frames, _ = synth.translating_frames(n, H, W, 12, seed_texture=51, seed_motion=52)
    cfg = me.SearchConfig(block=16, range=16)
    batches = [[(0, 1), (1, 2)], [(2, 3)], [(3, 4), (4, 5), (5, 6)], [(6, 0), (5, 1)], [(1, 0)], [(2, 5), (3, 3)]]
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=3) as est:
        est.upload(frames)
        want = [est.search(b).copy() for b in batches]                      # one at a time
        tickets = []
        for i, b in enumerate(batches):
            est.launch(b)
            tickets.append(est.last_ticket())
            if i >= 2:                                                       # keep three in flight, fetch the oldest
                j = i - 2
                assert est.fetch_pairs(len(batches[j]), ticket=tickets[j]).tobytes() == want[j].tobytes(), j
        for j in (len(batches) - 2, len(batches) - 1):
            assert est.fetch_pairs(len(batches[j]), ticket=tickets[j]).tobytes() == want[j].tobytes(), j
        assert est.fetch_pairs(len(batches[-1])).tobytes() == want[-1].tobytes()
        with pytest.raises(me.MEError):
            est.fetch_pairs(3, ticket=tickets[-1])                           # that search had two pairs


def test_resident_search_overlapping_denoise_on_the_copy_stream(me):
    """Pre-processing kernels run on the copy stream while a resident (one CTA per SM) search runs on the compute stream:
    NLM CTAs hold SMs, so not all resident CTAs need to be co-resident from the start.  Results and the kernel's error flag
    must be unaffected (fetch_pairs raises if the resident kernel's wait guard tripped)."""
    from motion_estimation_b200 import synth
    H, W, n = 360, 640, 8
    frames, shifts = synth.translating_frames(n + 12, H, W, 24, seed_texture=41, seed_motion=42)
    cfg = me.SearchConfig(block=16, range=32)
    with me.MotionEstimator(H, W, cfg, max_frames=n + 12, max_pairs=n) as est:
        assert est.geometry["resident_kernel"] == 1
        est.upload(frames)
        pairs = [(k, k + 1) for k in range(n)]
        want = est.search(pairs).copy()
        for rep in range(3):
            est.launch(pairs)                                             # compute stream: resident kernel over slots 0..n
            est.denoise(n + 1, [1] * 11, 15.0, 19, 41)                    # copy stream: NLM on slots the search does not read
            got = est.fetch_pairs(n)
            assert got.tobytes() == want.tobytes(), rep
        check_pair(est, 2, frames[2], frames[3], cfg, got[2])


def _noisy_translating_frames(n, H, W, max_shift, seed):
    """Translating texture plus independent per-frame noise: the SAD minima are not exact zeros, so the sub-pixel fit,
    the peak gates and the median all have real work to do."""
    from motion_estimation_b200 import synth
    frames, shifts = synth.translating_frames(n, H, W, max_shift, seed_texture=seed, seed_motion=seed + 1)
    noise = np.random.default_rng(seed + 2).integers(-6, 7, size=frames.shape)
    return np.clip(frames.astype(np.int16) + noise, 0, 255).astype(np.uint8), shifts


def test_full_hd_every_block_and_the_fused_median_with_subpixel(me):
    """configs[2] at full size with the sub-pixel path on: ALL 8040 blocks of two 1920x1080 pairs (vectors, SADs, refined
    vectors, flags) and the pair records (3-level radix-select median, mean, counts) equal the oracle, through the resident
    kernel (1005 units per pair) and, for the second pair alone, through one CTA per tile."""
    frames, shifts = _noisy_translating_frames(3, 1080, 1920, 32, seed=21)
    cfg = me.SearchConfig(block=16, range=32)
    with me.MotionEstimator(1080, 1920, cfg, max_frames=3, max_pairs=2) as est:
        est.upload(frames)
        res = est.search([(0, 1), (1, 2)])
        for i in range(2):
            check_pair(est, i, frames[i], frames[i + 1], cfg, res[i])
            assert res[i]["nvalid_x"] > 7000 and abs(res[i]["pdx"] - shifts[i][0]) < 0.2 and abs(res[i]["pdy"] - shifts[i][1]) < 0.2


def test_4k_every_block_with_subpixel(me):
    """configs[3] at full size: all 129 600 blocks of a 3840x2160 pair, 8x8 blocks, +-64 px (slabbed candidate square,
    resident kernel), sub-pixel path on, against the oracle (about a second of host time on the box's cores)."""
    frames, shifts = _noisy_translating_frames(2, 2160, 3840, 64, seed=22)
    cfg = me.SearchConfig(block=8, range=64)
    with me.MotionEstimator(2160, 3840, cfg, max_frames=2, max_pairs=1) as est:
        est.upload(frames)
        res = est.search([(0, 1)])
        check_pair(est, 0, frames[0], frames[1], cfg, res[0])
        assert res[0]["nvalid_x"] > 100000 and abs(res[0]["pdx"] - shifts[0][0]) < 0.2


def _fuzz_case(seed):
    """A seeded random geometry: frame size (ragged against every block size), block, range (aligned and unaligned window
    origins, slabbed and 64-bit-key ranges now and then), batch size (small and large launches take different plans)."""
    rng = np.random.default_rng(1000 + seed)
    B = int(rng.choice([8, 16, 32]))
    R = int(rng.choice([1, 3, 6, 8, 11, 16, 17, 24, 28, 32, 33, 40, 48, 57, 64, 80, 100, 128]))
    H = int(rng.integers(B, 7 * B + 1)) if R > 48 else int(rng.integers(B, 12 * B + 1))
    W = int(rng.integers(B, 10 * B + 1)) if R > 48 else int(rng.integers(B, 40 * B + 1))
    n_pairs = int(rng.choice([1, 2, 5, 24]))
    mode = int(rng.integers(0, 3))          # 0 default plan, 1 resident kernel forced, 2 one CTA per tile forced
    kw = dict(subpixel=bool(rng.integers(0, 2)), min_range=int(rng.choice([0, 2, 30])), min_curv=int(rng.choice([1, 50])))
    kw["range_y"] = int(rng.choice([0, 0, 1, max(1, R // 4), max(1, R - 1)]))       # rectangular search windows now and then
    return H, W, B, R, n_pairs, mode, kw


@pytest.mark.parametrize("seed", range(40))
def test_random_geometries_equal_oracle(me, monkeypatch, seed):
    from motion_estimation_b200 import synth
    H, W, B, R, n_pairs, mode, kw = _fuzz_case(seed)
    if mode:
        monkeypatch.setenv("ME_PERSISTENT", "2" if mode == 1 else "0")
    n = 4
    frames, _ = synth.translating_frames(n, H, W, min(R, 20), seed_texture=seed, seed_motion=seed + 1)
    cfg = me.SearchConfig(block=B, range=R, **kw)
    rng = np.random.default_rng(seed)
    pairs = [tuple(int(v) for v in rng.integers(0, n, 2)) for _ in range(n_pairs)]
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=max(n_pairs, 2)) as est:
        est.upload(frames)
        res = est.search(pairs)
        for i in sorted(set(int(v) for v in rng.integers(0, n_pairs, 3))):      # the oracle runs on up to three pairs of the batch
            a, c = pairs[i]
            check_pair(est, i, frames[a], frames[c], cfg, res[i])
