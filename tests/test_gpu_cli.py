"""GPU tests of the two command-line entry points (same argv as the reference's scripts) on a small
synthetic recording written as an image sequence plus a timecode file."""
import csv
import importlib.util
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture()
def recording(tmp_path):
    import cv2
    from motion_estimation_b200 import synth
    n, H, W = 24, 64, 320
    canvas = synth.texture(H + 8, W + 8 * n + 16, seed=3)
    frames = np.stack([canvas[4:4 + H, 8 * n - 7 * k: 8 * n - 7 * k + W] for k in range(n)])   # content moves +7 px per frame
    d = tmp_path / "rec.d"
    d.mkdir()
    for k, f in enumerate(frames):
        cv2.imwrite(str(d / f"out.{k:04d}.png"), f)
    with open(d / "rec_tstamps.txt", "w") as fh:
        fh.write("# timecode format v2\n" + "\n".join(f"{5.531 * k:.3f}" for k in range(n + 2)) + "\n")
    with open(d / "tstamps.csv", "w") as fh:
        fh.write("\n".join(f"{int(5531 * (k + 1))},0" for k in range(n + 2)) + "\n")
    return str(d), frames


def test_estimate_vel_vid_cli(recording, tmp_path):
    d, frames = recording
    out = tmp_path / "out"
    res = _load("estimate_vel_vid").main(["0", "20", "-p", d, "--out", str(out), "--units", "fullres", "--range", "16", "--no-preprocess"])
    rows = list(csv.reader(open(out / "velocity_estimates.csv")))
    assert rows[0] == ["Timestamp [s]", "mm/min X-axis optical", "mm/min Y-axis optical"]
    assert len(rows) - 1 == len(res.out_information) == 19                 # every pair moves 7 px >= the gate of 2
    vx = np.array([float(r[0]) for r in rows[1:]])
    want = (7.0 / 611) / (5.531e-3 / 60)                                    # 7 px per 5.531 ms
    np.testing.assert_allclose(vx[1:], want, rtol=0.01)                     # row 0 is measured from t = 0 like the reference
    ref_units = _load("estimate_vel_vid").main(["0", "20", "-p", d, "--out", str(out), "--range", "16", "--no-preprocess"])
    np.testing.assert_allclose([r[0] for r in ref_units.out_information][1:], want / 2, rtol=0.01)   # half-resolution unit


def test_estimate_vel_vid_cli_default_is_the_reference_pipeline(recording, tmp_path):
    """The plain command line does what the reference's main() does (its lines 697-709): CLAHE, non-local means, then the
    pair loop.  The cache it leaves (images.h5 ['clean_images'] with h5py, else images.npy) holds what cv2 produces from
    the same frames, -c reuses it, and the explicit --preprocess of earlier versions gives the same CSV."""
    import cv2
    d, frames = recording
    out = tmp_path / "out3"
    mod = _load("estimate_vel_vid")
    res = mod.main(["0", "14", "-p", d, "--out", str(out), "--units", "fullres", "--range", "16"])
    clean = mod.load_cache(os.path.join(d, "images.h5"), os.path.join(d, "images.npy"))
    assert clean.shape == (14, 64, 320)
    clahe = cv2.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
    eq = np.stack([clahe.apply(z) for z in frames[:14]])
    for k, T in ((0, 1), (3, 5), (6, 9), (10, 5), (13, 1)):
        want = cv2.fastNlMeansDenoising(src=eq[k], h=15, templateWindowSize=19, searchWindowSize=41) if T == 1 else \
            cv2.fastNlMeansDenoisingMulti(srcImgs=eq, imgToDenoiseIndex=k, temporalWindowSize=T, h=15, templateWindowSize=19, searchWindowSize=41)
        np.testing.assert_array_equal(clean[k], want, err_msg=f"frame {k}")
    want_v = (7.0 / 611) / (5.531e-3 / 60)
    np.testing.assert_allclose([r[0] for r in res.out_information][1:], want_v, rtol=0.02)
    csv_default = open(out / "velocity_estimates.csv").read()
    again = mod.main(["0", "14", "-p", d, "--out", str(out), "--units", "fullres", "--range", "16", "-c"])
    assert [tuple(r) for r in again.out_information] == [tuple(r) for r in res.out_information]
    mod.main(["0", "14", "-p", d, "--out", str(tmp_path / "out4"), "--units", "fullres", "--range", "16", "--preprocess"])
    assert open(tmp_path / "out4" / "velocity_estimates.csv").read() == csv_default
    fast = mod.main(["0", "14", "-p", d, "--out", str(tmp_path / "out5"), "--units", "fullres", "--range", "16", "--no-preprocess"])
    assert len(fast.out_information) == len(res.out_information)


def test_estimate_vel_cli(recording, tmp_path):
    d, frames = recording
    out = tmp_path / "out2"
    res = _load("estimate_vel").main(["1", "12", "-p", d, "--out", str(out), "--range", "16"])
    rows = list(csv.reader(open(out / "velocity_estimates.csv")))
    assert rows[0] == ["mm/min X-axis", "mm/min Y-axis", "Timestamp [s]"] and len(rows) - 1 == 11
    k = 5
    t = int(5531 * (1 + k + 1)) / 10E+6                                      # the reference's divisor (estimate_vel.py:354)
    assert abs(res.out_information[k][2] - t) < 1e-12
    assert abs(res.out_information[k][0] - (7.0 / 611) / (t / 60)) / ((7.0 / 611) / (t / 60)) < 0.01
