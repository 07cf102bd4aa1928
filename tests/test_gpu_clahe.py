"""GPU tests of the CLAHE kernels (estimate_vel_vid.py:201-219 on the device): bit-exact against
oracle/clahe_np.py, which tests/test_oracle.py pins against cv2."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import clahe_np   # noqa: E402  (checker only)


@pytest.mark.parametrize("H,W,clip,tiles", [(64, 640, 2.0, (8, 8)), (70, 100, 2.0, (8, 8)), (64, 100, 2.0, (8, 8)),
                                           (33, 47, 2.0, (8, 8)), (120, 90, 4.0, (4, 6)), (96, 96, 0.0, (8, 8)),
                                           (270, 480, 3.0, (16, 8))])
def test_clahe_equals_oracle(me, H, W, clip, tiles):
    import cv2
    rng = np.random.default_rng(H + W)
    frames = np.stack([rng.integers(0, 256, (H, W), dtype=np.uint8),
                       cv2.GaussianBlur(rng.integers(0, 256, (H, W), dtype=np.uint8), (0, 0), 3),
                       (cv2.GaussianBlur(rng.integers(0, 256, (H, W), dtype=np.uint8), (0, 0), 2) // 12).astype(np.uint8),
                       np.full((H, W), 200, np.uint8)])
    with me.MotionEstimator(max(H, 16), max(W, 16), me.SearchConfig(16, 4), max_frames=6, max_pairs=1) as est:
        est.upload(frames, first_slot=1)
        est.clahe(1, 4, clip, tiles)
        got = est.download(1, 4)
        for k in range(4):
            np.testing.assert_array_equal(got[k], clahe_np.apply(frames[k], clip, tiles), err_msg=f"frame {k}")
        # untouched slots stay untouched, and a second pass is a new CLAHE of the already equalised frame
        est.clahe(2, 1, clip, tiles)
        np.testing.assert_array_equal(est.download(2, 1)[0], clahe_np.apply(got[1], clip, tiles))
        np.testing.assert_array_equal(est.download(1, 1)[0], got[0])


def test_adaptive_histogram_equalization_drop_in(me, golden_clips):
    """Same frames out as the reference's function produces with cv2 (checked against cv2 directly here)."""
    import cv2
    from motion_estimation_b200 import api
    f = golden_clips["29.1"]["frames"][:10]
    eq = api.adaptive_histogram_equalization(f, chunk=4)
    ref = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
    assert eq.shape == f.shape and eq.dtype == np.uint8
    for k in range(len(f)):
        np.testing.assert_array_equal(eq[k], ref.apply(f[k]))
    api.close_engines()


def test_search_after_clahe_sees_the_equalised_frames(me, golden_clips):
    from oracle import sad as osad
    f = golden_clips["58.2"]["frames"][:3]
    cfg = me.SearchConfig(16, 16)
    with me.MotionEstimator(64, 640, cfg, max_frames=3, max_pairs=2) as est:
        est.upload(f)
        est.clahe(0, 3)
        res = est.search([(0, 1), (1, 2)])
        eq = np.stack([clahe_np.apply(x) for x in f])
        for i in range(2):
            o = osad.search(eq[i], eq[i + 1], 16, 16)
            assert (res[i]["pdx"], res[i]["pdy"], res[i]["nvalid_x"]) == (o["pdx"], o["pdy"], o["nvalid_x"])
