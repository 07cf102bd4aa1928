"""CPU tests of the oracle itself: the C restatement against the independent NumPy restatement,
against the committed golden vectors, and against properties the spec implies.
(Parity of the SAD path with the reference is unpinned -- the reference has no SAD search; what the
reference does pin, its ECC function, is checked through oracle/ecc_port.py below.)"""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import ecc_port, sad, sad_np

KEYS = ("mv", "sad", "sub", "flags")


def assert_same(a, b):
    for k in KEYS:
        np.testing.assert_array_equal(a[k], b[k], err_msg=k)
    for k in ("pdx", "pdy", "flag", "nvalid_x", "nvalid_y", "n_textured", "mean_dx", "mean_dy"):
        assert a[k] == b[k], k


def textured(rng, H, W, blur=2.0):
    import cv2
    a = rng.integers(0, 256, (H + 64, W + 64), dtype=np.uint8)
    return cv2.GaussianBlur(a, (0, 0), blur)


@pytest.mark.parametrize("H,W,B,R,mc", [(64, 96, 16, 8, 1), (40, 52, 8, 5, 3), (33, 47, 4, 3, 1),
                                        (48, 80, 16, 40, 1), (64, 64, 32, 9, 2), (17, 130, 16, 20, 1)])
def test_c_oracle_equals_numpy_restatement(H, W, B, R, mc):
    rng = np.random.default_rng(H * 1000 + W)
    a = textured(rng, H, W)
    prev, cur = a[32:32 + H, 32:32 + W], a[30:30 + H, 29:29 + W]      # content moves (+3, +2)
    assert_same(sad.search(prev, cur, B, R, min_range=4, min_curv=mc), sad_np.search(prev, cur, B, R, min_range=4, min_curv=mc))


@pytest.mark.parametrize("H,W,B,R,Ry", [(64, 160, 16, 24, 3), (64, 640, 16, 64, 8), (40, 96, 8, 12, 1), (96, 96, 32, 20, 20)])
def test_rectangular_range_c_oracle_equals_numpy_restatement(H, W, B, R, Ry):
    """range_y: |dy| <= Ry < R (the recordings move along x); same spec otherwise, same raster-order tie-break."""
    rng = np.random.default_rng(H + W + Ry)
    a = textured(rng, H, W)
    prev, cur = a[32:32 + H, 32:32 + W], a[31:31 + H, 27:27 + W]      # content moves (+5, +1)
    c, n = sad.search(prev, cur, B, R, range_y=Ry), sad_np.search(prev, cur, B, R, range_y=Ry)
    assert_same(c, n)
    assert np.abs(c["mv"][..., 1]).max() <= Ry
    if Ry == R:
        assert_same(c, sad.search(prev, cur, B, R))


def test_integer_only_mode():
    rng = np.random.default_rng(5)
    a = textured(rng, 48, 64)
    prev, cur = a[32:80, 32:96], a[33:81, 30:94]
    r = sad.search(prev, cur, 16, 6, subpixel=False)
    assert_same(r, sad_np.search(prev, cur, 16, 6, subpixel=False))
    np.testing.assert_array_equal(r["sub"], r["mv"].astype(np.float32))


def test_tie_break_is_first_minimum_in_raster_order():
    # constant frames: every in-frame candidate has SAD 0, so the argmin is the first in-domain
    # candidate in (dy outer, dx inner) order = the most negative in-frame (dy, dx)
    z = np.full((48, 64), 7, np.uint8)
    r = sad.search(z, z, 16, 5, min_range=0)
    for by in range(3):
        for bx in range(4):
            dx, dy = r["mv"][by, bx]
            assert dx == max(-5, -bx * 16) and dy == max(-5, -by * 16)
    assert (r["sad"] == 0).all()
    assert r["nvalid_x"] == 0 and r["flag"] == 0           # flat cost surface: den == 0, never peaked
    assert_same(r, sad_np.search(z, z, 16, 5, min_range=0))
    # a periodic pattern produces exact ties at several displacements; both restatements must agree
    p = np.tile(np.arange(8, dtype=np.uint8) * 20, (48, 8))
    assert_same(sad.search(p, p, 16, 12), sad_np.search(p, p, 16, 12))


def test_cost_surface_matches_volume():
    rng = np.random.default_rng(2)
    a = textured(rng, 40, 56)
    prev, cur = a[32:72, 32:88], a[31:71, 34:90]
    vol = sad_np.cost_volume(prev, cur, 8, 6)
    for bx, by in [(0, 0), (3, 2), (6, 4)]:
        cs = sad.cost_surface(prev, cur, 8, 6, bx, by).astype(np.int64)
        v = vol[by, bx].copy()
        v[v == sad_np._INF] = 2 ** 32 - 1
        np.testing.assert_array_equal(cs, v)


def test_empty_and_degenerate_inputs():
    z = np.zeros((8, 8), np.uint8)
    with pytest.raises(ValueError):
        sad.search(z, z, 16, 4)                            # frame smaller than a block
    with pytest.raises(ValueError):
        sad.search(np.zeros((16, 16), np.uint8), np.zeros((16, 32), np.uint8), 16, 4)
    r = sad.search(np.zeros((16, 16), np.uint8), np.zeros((16, 16), np.uint8), 16, 4)   # one block, domain {(0,0)}
    assert tuple(r["mv"][0, 0]) == (0, 0) and r["flags"][0, 0] == 0 and r["flag"] == 0


@settings(max_examples=25, deadline=None)
@given(dx=st.integers(-6, 6), dy=st.integers(-6, 6), seed=st.integers(0, 1000))
def test_integer_translation_is_recovered_exactly(dx, dy, seed):
    rng = np.random.default_rng(seed)
    a = textured(rng, 64, 96, blur=1.5)
    prev = a[32:96, 32:128]
    cur = a[32 - dy:96 - dy, 32 - dx:128 - dx]            # content moves by (+dx, +dy)
    r = sad.search(prev, cur, 16, 8)
    # the parabola is not symmetric around an exact integer minimum, so the refined median is only close
    assert r["flag"] == 1 and abs(r["pdx"] - dx) < 0.1 and abs(r["pdy"] - dy) < 0.1
    ri = sad.search(prev, cur, 16, 8, subpixel=False)
    assert ri["pdx"] == dx and ri["pdy"] == dy
    inner = r["mv"][1:-1, 1:-1]
    assert (inner[..., 0] == dx).all() and (inner[..., 1] == dy).all()
    assert (r["sad"][1:-1, 1:-1] == 0).all()


def test_golden_sad_pins(golden_clips):
    """The C oracle still produces the vectors committed with the example excerpts."""
    for name, g in golden_clips.items():
        f = g["frames"]
        o0 = sad.search(f[0], f[4], 16, 32)
        o1 = sad.search(f[8], f[9], 8, 12)
        for k in KEYS:
            np.testing.assert_array_equal(o0[k], g[f"sad_b16r32_{k}"], err_msg=f"{name} {k}")
            np.testing.assert_array_equal(o1[k], g[f"sad_b8r12_{k}"], err_msg=f"{name} {k}")
        for (a, b), want in zip(g["pairs"], g["sad_pairs"]):
            o = sad.search(f[a], f[b], 16, 32)
            got = [float(o["pdx"]), float(o["pdy"]), o["flag"], o["nvalid_x"], o["nvalid_y"]]
            assert got == list(want), (name, a, b)


def test_search_pairs_batch_equals_single(golden_clips):
    f = golden_clips["29.1"]["frames"]
    pairs = [(0, 1), (3, 7), (10, 5)]
    pd, pi = sad.search_pairs(f, pairs, 16, 16)
    for (a, b), d, i in zip(pairs, pd, pi):
        o = sad.search(f[a], f[b], 16, 16)
        assert (d[0], d[1], i[0], i[1], i[2]) == (o["pdx"], o["pdy"], o["flag"], o["nvalid_x"], o["nvalid_y"])


def test_ecc_port_reproduces_reference_outputs(golden_clips):
    """oracle/ecc_port.py against numbers the reference's own calc_ECC_transform / calculate_dense_flow
    produced in the build container (tools/make_golden.py); same cv2 wheel on both machines."""
    for name, g in golden_clips.items():
        f = g["frames"]
        for (a, b), want, wflow in list(zip(g["pairs"], g["ecc_ref"], g["flow_ref"]))[::3]:
            pdx, pdy, flag = ecc_port.ecc_pair_video(f[a], f[b])
            assert flag == int(want[2]), (name, a, b)
            if flag:
                np.testing.assert_allclose([pdx, pdy], want[:2], rtol=0, atol=1e-4)
            np.testing.assert_allclose(ecc_port.dense_flow_mean(f[a], f[b]), wflow, rtol=0, atol=1e-4)


def test_sad_agrees_with_reference_ecc_where_both_measure(golden_clips):
    """Soft link to the reference: where its ECC converged and the motion is inside the search range,
    the block search sees the same displacement (ECC is in half-resolution pixels, Finding B)."""
    n = 0
    for name, g in golden_clips.items():
        for want, s in zip(g["ecc_ref"], g["sad_pairs"]):
            if int(want[2]) == 1 and abs(2 * want[0]) < 28 and s[3] > 100:
                assert abs(2 * want[0] - s[0]) < 0.6 and abs(2 * want[1] - s[1]) < 0.6, (name, want, s)
                n += 1
    assert n > 40


def test_bgr2gray_matches_cv2():
    import cv2
    rng = np.random.default_rng(3)
    bgr = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    np.testing.assert_array_equal(sad.bgr2gray(bgr), cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY))


@pytest.mark.parametrize("shape,clip,tiles", [((64, 640), 2.0, (8, 8)), ((70, 100), 2.0, (8, 8)), ((64, 100), 2.0, (8, 8)),
                                             ((33, 47), 2.0, (8, 8)), ((120, 90), 4.0, (4, 6)), ((96, 96), 0.0, (8, 8)),
                                             ((540, 960), 2.0, (8, 8))])
def test_clahe_restatement_equals_cv2(shape, clip, tiles):
    """oracle/clahe_np.py against the library call the reference makes (estimate_vel_vid.py:203-205)."""
    import cv2
    from oracle import clahe_np
    rng = np.random.default_rng(shape[0] * 7 + shape[1])
    ref = cv2.createCLAHE(clipLimit=clip, tileGridSize=tiles)
    for img in (rng.integers(0, 256, shape, dtype=np.uint8),
                cv2.GaussianBlur(rng.integers(0, 256, shape, dtype=np.uint8), (0, 0), 3),
                (cv2.GaussianBlur(rng.integers(0, 256, shape, dtype=np.uint8), (0, 0), 2) // 12).astype(np.uint8),
                np.full(shape, 9, np.uint8)):
        np.testing.assert_array_equal(clahe_np.apply(img, clip, tiles), ref.apply(img))


def test_clahe_restatement_on_example_frames(golden_clips):
    import cv2
    from oracle import clahe_np
    ref = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
    for g in golden_clips.values():
        for k in (0, 17):
            np.testing.assert_array_equal(clahe_np.apply(g["frames"][k]), ref.apply(g["frames"][k]))


@pytest.mark.parametrize("shape,T,k,h,tw,sw", [((20, 33), 1, 0, 15, 19, 41), ((20, 33), 5, 2, 15, 19, 41), ((20, 33), 9, 4, 15, 19, 41),
                                               ((37, 50), 3, 1, 10, 7, 21), ((37, 50), 1, 2, 3, 7, 21), ((30, 30), 3, 2, 30, 5, 11)])
def test_nlm_restatement_equals_cv2(shape, T, k, h, tw, sw):
    """oracle/nlm_np.py against the library calls the reference makes (estimate_vel_vid.py:226-232, 251-262)."""
    import cv2
    from oracle import nlm_np
    rng = np.random.default_rng(sum(shape) + T)
    base = cv2.GaussianBlur(rng.integers(0, 256, shape).astype(np.uint8), (0, 0), 2).astype(np.float64) * 1.5 - 60
    frames = np.clip(base[None] + rng.normal(0, 10, (max(2 * k + 1, T + k),) + shape), 0, 255).astype(np.uint8)
    if T == 1:
        ref = cv2.fastNlMeansDenoising(src=frames[k], h=h, templateWindowSize=tw, searchWindowSize=sw)
    else:
        ref = cv2.fastNlMeansDenoisingMulti(srcImgs=frames, imgToDenoiseIndex=k, temporalWindowSize=T, h=h,
                                            templateWindowSize=tw, searchWindowSize=sw)
    got = nlm_np.denoise(frames, k, T, h, tw, sw)
    assert h < 5 or (ref != frames[k]).mean() > 0.5               # the filter did something (h=3 is below the noise: identity)
    np.testing.assert_array_equal(got, ref)


def test_nlm_reference_window_rule():
    """Which call denoise_hf5 makes for frame k (estimate_vel_vid.py:250-262)."""
    from oracle import nlm_np
    assert nlm_np.reference_windows(12) == [1, 1, 5, 5, 5, 9, 9, 5, 5, 1, 1, 1]
    assert nlm_np.reference_windows(11, 12) == [1, 1, 5, 5, 5, 9, 9, 5, 5, 1, 1]     # len(numberList) = frames + 1 in the scripts
    with pytest.raises(ValueError):
        nlm_np.denoise(np.zeros((3, 8, 8), np.uint8), 0, 3)
