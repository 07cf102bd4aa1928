"""GPU tests of the non-local-means kernel (estimate_vel_vid.py:226-232, 246-262 on the device): bit-exact against
cv2's fastNlMeansDenoising / fastNlMeansDenoisingMulti (the third-party code the reference calls; installed here
and on the GPU box) and against oracle/nlm_np.py, which tests/test_oracle.py pins against cv2."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from oracle import nlm_np   # noqa: E402  (checker only)


def _noisy(n, H, W, seed, sigma=10.0):
    import cv2
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (H, W)).astype(np.uint8)
    base = cv2.GaussianBlur(base, (0, 0), 2.5).astype(np.float64) * 1.6 - 70
    return np.clip(base[None] + rng.normal(0, sigma, (n, H, W)), 0, 255).astype(np.uint8)


def _cv(frames, k, T, h=15, tw=19, sw=41):
    import cv2
    if T == 1:
        return cv2.fastNlMeansDenoising(src=frames[k], h=h, templateWindowSize=tw, searchWindowSize=sw)
    return cv2.fastNlMeansDenoisingMulti(srcImgs=frames, imgToDenoiseIndex=k, temporalWindowSize=T, h=h,
                                         templateWindowSize=tw, searchWindowSize=sw)


@pytest.mark.parametrize("H,W,split", [(64, 640, None), (64, 640, 1), (40, 150, None), (70, 129, 3), (20, 33, None), (130, 260, 1)])
def test_reference_parameters_equal_cv2(me, monkeypatch, H, W, split):
    """h=15, template 19, search 41 with the reference's mix of single / 5-frame / 9-frame windows."""
    if split is not None:
        monkeypatch.setenv("ME_NLM_SPLIT", str(split))
    n = 11
    frames = _noisy(n, H, W, H * 1000 + W)
    windows = nlm_np.reference_windows(n)
    assert windows == [1, 1, 5, 5, 5, 9, 5, 5, 1, 1, 1]
    with me.MotionEstimator(max(H, 16), max(W, 16), me.SearchConfig(16, 4), max_frames=n + 1, max_pairs=1) as est:
        est.upload(frames, first_slot=1)
        est.denoise(1, windows)
        got = est.download(1, n)
    for k, T in enumerate(windows):
        np.testing.assert_array_equal(got[k], _cv(frames, k, T), err_msg=f"frame {k} window {T}")


@pytest.mark.parametrize("h,tw,sw", [(10, 7, 21), (3, 7, 21), (30, 7, 11), (15, 19, 21), (15, 18, 40), (0, 7, 21)])
def test_other_parameters_equal_cv2(me, h, tw, sw):
    frames = _noisy(4, 45, 70, 5)
    windows = [1, 3, 3, 0] if tw % 2 and sw % 2 else [1, 1, 0, 1]     # cv2's Multi rejects even window sizes, the single call rounds them
    with me.MotionEstimator(45, 70, me.SearchConfig(16, 4), max_frames=4, max_pairs=1) as est:
        est.upload(frames)
        est.denoise(0, windows, h, tw, sw)
        got = est.download(0, 4)
    for k, T in enumerate(windows):
        want = frames[k] if T == 0 else _cv(frames, k, T, h, tw, sw)
        np.testing.assert_array_equal(got[k], want, err_msg=f"frame {k} window {T}")


def test_equals_oracle_small(me):
    frames = _noisy(3, 24, 40, 9)
    with me.MotionEstimator(24, 40, me.SearchConfig(8, 4), max_frames=3, max_pairs=1) as est:
        est.upload(frames)
        est.denoise(0, [1, 3, 1])
        got = est.download(0, 3)
    for k, T in enumerate([1, 3, 1]):
        np.testing.assert_array_equal(got[k], nlm_np.denoise(frames, k, T))


def test_rejects_what_cv2_rejects(me):
    with me.MotionEstimator(32, 32, me.SearchConfig(8, 4), max_frames=4, max_pairs=1) as est:
        est.upload(np.zeros((4, 32, 32), np.uint8))
        for bad in ([5, 1, 1, 1], [1, 1, 1, 3], [1, 4, 1, 1], [1, -1, 1, 1]):
            with pytest.raises(me.MEError):
                est.denoise(0, bad)
        with pytest.raises(me.MEError):
            est.denoise(0, [1, 3, 1], template_window=18)   # 'All windows sizes should be odd' for the temporal call
        with pytest.raises(me.MEError):
            est.denoise(0, [1, 1], template_window=11)      # not built
        with pytest.raises(me.MEError):
            est.denoise(2, [1, 1, 1])                       # leaves the ring


def test_denoise_hf5_drop_in(me, golden_clips):
    """The reference's pre-processing order on a real excerpt: CLAHE then denoise_hf5, against cv2 doing the same."""
    import cv2
    from motion_estimation_b200 import api
    f = golden_clips["116.4"]["frames"][:12]
    eq = api.adaptive_histogram_equalization(f)
    clean = api.denoise_hf5(eq, chunk=5)                    # chunks smaller than the sequence: halo frames travel with each chunk
    assert clean.shape == eq.shape and clean.dtype == np.uint8
    windows = api.reference_windows(len(eq))
    clahe = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
    eq_ref = np.stack([clahe.apply(z) for z in f])
    for k, T in enumerate(windows):
        np.testing.assert_array_equal(clean[k], _cv(eq_ref, k, T), err_msg=f"frame {k} window {T}")
    # one frame through the reference's per-frame helper
    np.testing.assert_array_equal(api.denoising(eq, 6, 9), _cv(eq_ref, 6, 9))
    with pytest.raises(ValueError):
        api.denoising(eq, 1, 5)
    # a video shorter than numberList: the reference stops at the first window that leaves the array
    short = api.denoise_hf5(eq, n_list=len(eq) + 4)
    assert len(short) == len(eq) - 4
    np.testing.assert_array_equal(short[7], _cv(eq_ref, 7, 9))
    api.close_engines()


@pytest.mark.parametrize("seed", range(12))
def test_random_shapes_equal_cv2(me, monkeypatch, seed):
    """Seeded random frame sizes (below, around and above one 64x128 tile), windows, strengths and CTA splits."""
    rng = np.random.default_rng(500 + seed)
    H, W = int(rng.integers(16, 150)), int(rng.integers(16, 300))
    tw, sw = int(rng.choice([7, 19])), int(rng.choice([11, 21, 41]))
    h = float(rng.choice([4, 10, 15, 40]))
    n = 5
    windows = [1, int(rng.choice([1, 3])), int(rng.choice([3, 5])), int(rng.choice([0, 3])), 1]
    split = int(rng.choice([0, 1, 2, 5]))
    if split:
        monkeypatch.setenv("ME_NLM_SPLIT", str(split))
    frames = _noisy(n, H, W, seed, sigma=float(rng.choice([3, 10, 25])))
    with me.MotionEstimator(max(H, 16), max(W, 16), me.SearchConfig(8, 4), max_frames=n, max_pairs=1) as est:
        est.upload(frames)
        est.denoise(0, windows, h, tw, sw)
        got = est.download(0, n)
    for k, T in enumerate(windows):
        want = frames[k] if T == 0 else _cv(frames, k, T, h, tw, sw)
        np.testing.assert_array_equal(got[k], want, err_msg=f"frame {k} window {T} ({H}x{W}, h={h}, {tw}/{sw}, split {split})")
