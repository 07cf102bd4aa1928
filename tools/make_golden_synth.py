#!/usr/bin/env python3
"""Golden vectors for the velocity-accuracy test on a synthetic sub-pixel sequence (BUILD container only).

1000 frames of 1280x128, clean translating texture, 0.6..2.4 px per frame (motion_estimation_b200.synth.
subpixel_sequence, seed 5).  The REFERENCE'S OWN end_process (estimate_vel_vid.py:493-583 with its ECC :306-369 and
Farneback :385-391, ast-extracted like tools/make_golden.py) runs over the sequence; its rows, cumulative position and
the dense-flow total it prints are stored in tests/golden/synth_subpixel.npz with the crc32 of the frames.  The
reference's calculate_dense_flow on a sample of fixed pairs is stored too (for the dense-field comparison)."""
import io
import os
import sys
import tempfile
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from make_golden import reference_functions          # noqa: E402
from motion_estimation_b200 import synth             # noqa: E402

N, H, W, SEED, DT_MS = 1000, 128, 1280, 5, 5.531


FAST = dict(n=400, lo=4.5, hi=9.5)      # second sequence: every consecutive pair passes the gate, sub-pixel phases uniform


def run_reference(frames):
    d = tempfile.mkdtemp()
    n = len(frames)
    lines = ["# timecode format v2"] + [repr(DT_MS * k) for k in range(n + 3)]
    open(os.path.join(d, "s_tstamps.txt"), "w").write("\n".join(lines) + "\n")
    printed = io.StringIO()
    g = {"firstFrame": 0, "numberList": list(range(n + 1)), "imagePath": d,
         "print": lambda *a, **k: printed.write(" ".join(str(v) for v in a) + "\n")}
    ns = reference_functions({"calc_ECC_transform", "calculate_dense_flow", "FB_opt_flow_params", "end_process",
                              "max_filament_speed", "max_filament_speed_sec", "pixels_per_mm", "velocity_list_x",
                              "velocity_list_y", "motion_list_opt", "out_information", "tsList"}, g)
    info, vx, vy, ts, motion = ns["end_process"](frames)
    tail = printed.getvalue().strip().split("\n")[-3:]
    return info, ts, motion, int(tail[0].split(":")[1].split("of")[0]), float(tail[2].split(":")[1])


def main_fast():
    frames, x, y = synth.subpixel_sequence(FAST["n"], H, W, SEED, step_lo=FAST["lo"], step_hi=FAST["hi"])
    info, ts, motion, failed, dense = run_reference(frames)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "synth_subpixel_fast.npz"), n=FAST["n"], h=H, w=W, seed=SEED,
                        dt_ms=DT_MS, step_lo=FAST["lo"], step_hi=FAST["hi"], crc32_frames=zlib.crc32(frames.tobytes()),
                        ref_out_information=np.array(info, np.float64).reshape(-1, 3), ref_motion=np.array(motion, np.float64),
                        ref_ts=np.array(ts, np.float64), ref_failed=failed, ref_final_dense_mm=dense)
    print(f"fast sequence: reference end_process {len(info)} rows, {failed} failed, final {motion[-1]:.4f} mm")


def main():
    if "--fast-only" in sys.argv:
        return main_fast()
    frames, x, y = synth.subpixel_sequence(N, H, W, SEED)
    d = tempfile.mkdtemp()
    lines = ["# timecode format v2"] + [repr(DT_MS * k) for k in range(N + 3)]
    open(os.path.join(d, "s_tstamps.txt"), "w").write("\n".join(lines) + "\n")
    printed = io.StringIO()
    g = {"firstFrame": 0, "numberList": list(range(N + 1)), "imagePath": d,
         "print": lambda *a, **k: printed.write(" ".join(str(v) for v in a) + "\n")}
    ns = reference_functions({"calc_ECC_transform", "calculate_dense_flow", "FB_opt_flow_params", "end_process",
                              "max_filament_speed", "max_filament_speed_sec", "pixels_per_mm", "velocity_list_x",
                              "velocity_list_y", "motion_list_opt", "out_information", "tsList"}, g)
    t0 = time.time()
    info, vx, vy, ts, motion = ns["end_process"](frames)
    tail = printed.getvalue().strip().split("\n")[-3:]
    failed, final_dense = int(tail[0].split(":")[1].split("of")[0]), float(tail[2].split(":")[1])
    pairs = [(k, k + g_) for k in range(10, N - 8, 97) for g_ in (1, 3)]
    flows = [ns["calculate_dense_flow"](frames[a], frames[b]) for a, b in pairs]
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "synth_subpixel.npz"), n=N, h=H, w=W, seed=SEED, dt_ms=DT_MS,
                        crc32_frames=zlib.crc32(frames.tobytes()), x=x, y=y,
                        ref_out_information=np.array(info, np.float64).reshape(-1, 3), ref_motion=np.array(motion, np.float64),
                        ref_ts=np.array(ts, np.float64), ref_failed=failed, ref_final_dense_mm=final_dense,
                        flow_pairs=np.array(pairs, np.int32), ref_flow=np.array(flows, np.float64))
    print(f"reference end_process: {len(info)} rows, {failed} failed, {time.time() - t0:.0f} s; final {motion[-1]:.4f} mm, dense {final_dense:.4f} mm")


if __name__ == "__main__":
    main()
    if "--fast-only" not in sys.argv:
        main_fast()
