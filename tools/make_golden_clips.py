#!/usr/bin/env python3
"""Whole-stretch golden series from the REFERENCE'S OWN functions (run in the BUILD container only).

Input: the clip excerpts of tools/make_clip_fixtures.py (tests/golden/clips/<name>/clip.h264, 2040 moving
frames each) after the reference's pre-processing order -- CLAHE, then non-local means
(estimate_vel_vid.py:703-709).  The cleaned frames come from the GPU kernels (tools/gpu_clean_clips.py on the
B200 box -> gpurun_out/clean_<name>.xz), which are bit-exact with the cv2 calls they replace; a sample of frames is
re-computed here with cv2 itself to confirm that before anything is derived from them.

Then, per clip, the reference's `end_process` (estimate_vel_vid.py:493-583, with its `calc_ECC_transform`
:306-369 and `calculate_dense_flow` :385-391, extracted with `ast` and exec'd against the installed cv2) runs over
all 2040 frames; its rows (vxm, vym, t), cumulative position, and the two totals it prints (:581-582) are saved to
tests/golden/clips/<name>/reference_series.npz together with the crc32 of the cleaned frames.

For the 116.4 recording the reference's `encoder_velocity` (:417-487) runs on the full encoder log and timecode
file; its (time, position) series is saved as tests/golden/clips/116.4/encoder_series.npz.
"""
import glob
import io
import json
import lzma
import os
import sys
import time
import zlib
from pathlib import Path

import cv2 as cv
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_golden import REF, reference_functions   # noqa: E402
from make_clip_fixtures import decode              # noqa: E402

CLIPS = os.path.join(ROOT, "tests", "golden", "clips")
H, W = 64, 640


def cv2_clean_frame(eq, k, n_list):
    """What denoise_hf5 computes for frame k (estimate_vel_vid.py:250-262)."""
    if k <= 1 or k >= n_list - 3:
        return cv.fastNlMeansDenoising(src=eq[k], h=15, templateWindowSize=19, searchWindowSize=41)
    t = 5 if (k <= 4 or k >= n_list - 5) else 9
    return cv.fastNlMeansDenoisingMulti(srcImgs=list(eq), imgToDenoiseIndex=k, temporalWindowSize=t, h=15,
                                        templateWindowSize=19, searchWindowSize=41)


def load_clean(name, n):
    clean = np.frombuffer(lzma.decompress(open(os.path.join(ROOT, "gpurun_out", f"clean_{name}.xz"), "rb").read()),
                          np.uint8).reshape(-1, H, W)
    assert len(clean) == n, (name, len(clean))
    return clean


def reference_pairs(man):
    """The reference's calc_ECC_transform on every 5th consecutive pair (k, k+1) of the cleaned frames: the per-pair
    numbers the GPU search is compared with directly (tests/test_gpu_recordings.py)."""
    ref = reference_functions({"calc_ECC_transform"}, {})
    for name, m in man.items():
        clean = load_clean(name, m["frames"])
        assert zlib.crc32(clean.tobytes()) == m["crc32_clean"]
        ks = np.arange(0, m["frames"] - 1, 5)
        ecc = np.array([[float(v) for v in ref["calc_ECC_transform"](clean[k], clean[k + 1])] for k in ks])
        np.savez_compressed(os.path.join(CLIPS, name, "reference_pairs.npz"), k=ks.astype(np.int32), pdx=ecc[:, 0].astype(np.float32),
                            pdy=ecc[:, 1].astype(np.float32), flag=ecc[:, 2].astype(np.int32))
        print(name, "reference pairs:", len(ks), "converged:", int(ecc[:, 2].sum()), flush=True)


def main():
    man = json.load(open(os.path.join(CLIPS, "MANIFEST.json")))
    if "--pairs-only" in sys.argv:
        return reference_pairs(man)
    for name, m in man.items():
        if "--encoder-only" in sys.argv:
            break
        n = m["frames"]
        raw = decode(os.path.join(CLIPS, name, "clip.h264"), 0, n)
        assert zlib.crc32(raw.tobytes()) == m["crc32_decoded"]
        clean = load_clean(name, n)
        clahe = cv.createCLAHE(clipLimit=2.0, tileGridSize=(8, 8))
        eq = np.stack([clahe.apply(z) for z in raw])
        checked = []
        for k in (0, 3, 777, n - 5, n - 1):                   # single-image, 5-frame, 9-frame, 5-frame, single-image calls
            want = cv2_clean_frame(eq, k, n + 1)
            assert np.array_equal(want, clean[k]), f"{name}: GPU-cleaned frame {k} differs from cv2"
            checked.append(k)
        printed = io.StringIO()
        g = {"firstFrame": 0, "numberList": list(range(0, n + 1)), "imagePath": os.path.join(CLIPS, name),
             "print": lambda *a, **k: printed.write(" ".join(str(x) for x in a) + "\n")}
        ns = reference_functions({"calc_ECC_transform", "calculate_dense_flow", "FB_opt_flow_params", "end_process",
                                  "max_filament_speed", "max_filament_speed_sec", "pixels_per_mm", "velocity_list_x",
                                  "velocity_list_y", "motion_list_opt", "out_information", "tsList"}, g)
        t0 = time.time()
        out_information, vx, vy, ts_list, motion = ns["end_process"](clean)
        dt = time.time() - t0
        tail = printed.getvalue().strip().split("\n")[-3:]
        failed = int(tail[0].split(":")[1].split("of")[0])
        final_ecc = float(tail[1].split(":")[1])
        final_dense = float(tail[2].split(":")[1])
        np.savez_compressed(os.path.join(CLIPS, name, "reference_series.npz"),
                            out_information=np.array(out_information, np.float64).reshape(-1, 3),
                            motion=np.array(motion, np.float64), ts=np.array(ts_list, np.float64),
                            failed_estimates=failed, final_position_ecc=final_ecc, final_position_dense=final_dense,
                            crc32_clean=zlib.crc32(clean.tobytes()), frames_checked_against_cv2=np.array(checked))
        m.update(crc32_clean=zlib.crc32(clean.tobytes()), reference_rows=len(out_information), reference_failed=failed,
                 reference_final_position_mm=final_ecc, reference_final_dense_mm=final_dense, reference_seconds=round(dt, 1))
        print(name, {k: m[k] for k in m if k.startswith("reference") or k == "crc32_clean"}, flush=True)

    if "--encoder-only" not in sys.argv:
        reference_pairs(man)

    # ---- encoder ground truth of the 116.4 recording: the reference's own parser on the full log
    from scipy import interpolate
    src_dir = os.path.join(REF, "examples/116.4_mm_min.d")
    ts_lines = open(glob.glob(os.path.join(src_dir, "*tstamps.txt"))[0]).read().split("\n")
    # the shipped timecode file is cut off in the middle of its last line: index the last COMPLETE line
    last_frame = max(i + 1 for i, v in enumerate(ts_lines[:-1]) if v.strip() and not v.startswith("#")) - 1
    ns = reference_functions({"encoder_velocity", "encoder_out_info_list", "enc_pos_list", "enc_ts_list"},
                             {"imagePath": src_dir, "lastFrame": last_frame, "interpolate": interpolate})
    pos, t = ns["encoder_velocity"]()
    np.savez_compressed(os.path.join(CLIPS, "116.4", "encoder_series.npz"), t_s=np.array(t, np.float32),
                        position_mm=np.array(pos, np.float32), last_frame=last_frame)
    man["116.4"].update(encoder_rows=len(pos), encoder_last_frame=last_frame)
    json.dump(man, open(os.path.join(CLIPS, "MANIFEST.json"), "w"), indent=1)
    print("encoder rows", len(pos), "t_end", t[-1], "travel", pos[-1] - pos[0])


if __name__ == "__main__":
    main()
