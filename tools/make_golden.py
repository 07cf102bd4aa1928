#!/usr/bin/env python3
"""Generate tests/golden/* from the reference (run in the BUILD container only).

/root/reference does not exist on the GPU box, so everything the tests need from it is frozen here:

* a 40-frame excerpt of each example recording (decoded with cv.VideoCapture like
  estimate_vel_vid.py:165-178) with its timestamp lines;
* the outputs of the REFERENCE'S OWN functions on those frames -- ``calc_ECC_transform``,
  ``calculate_dense_flow`` and ``end_process`` are extracted from /root/reference/estimate_vel_vid.py
  with ``ast`` (the script cannot be imported: it parses argv and needs h5py / h264decoder /
  matplotlib at import time) and exec'd against the installed cv2;
* a check that oracle/ecc_port.py (the restated drivers) returns bit-identical numbers;
* the C SAD oracle's results on the same frames (regression pins for the spec).

Nothing is copied from the reference into the repo: only numbers it computed.
"""
import ast
import glob
import json
import linecache
import os
import sys
from pathlib import Path

import cv2 as cv
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ecc_port, sad as osad   # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")
CLIPS = {   # name -> (directory, first frame of the excerpt)
    "29.1": ("examples/29.1_mm_min.d", 1480),
    "58.2": ("examples/58.2_mm_mind", 1420),
    "116.4": ("examples/116.4_mm_min.d", 1000),
}
N_EXCERPT = 40


def reference_functions(names, extra_globals):
    """exec the named top-level functions / assignments of the reference video script."""
    src = open(os.path.join(REF, "estimate_vel_vid.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body
            if (isinstance(n, ast.FunctionDef) and n.name in names)
            or (isinstance(n, ast.Assign) and any(isinstance(t, ast.Name) and t.id in names for t in n.targets))]
    ns = {"cv": cv, "np": np, "Path": Path, "linecache": linecache, "print": lambda *a, **k: None}
    ns.update(extra_globals)
    exec(compile(ast.Module(body=keep, type_ignores=[]), "reference:estimate_vel_vid.py", "exec"), ns)
    return ns


def decode(directory, first, n):
    path = sorted(glob.glob(os.path.join(REF, directory, "*.h264")))[0]
    cap = cv.VideoCapture(path)
    frames, k = [], 0
    while len(frames) < n:
        ok, f = cap.read()
        if not ok:
            break
        if k >= first:
            frames.append(cv.cvtColor(f, cv.COLOR_BGR2GRAY))
        k += 1
    cap.release()
    return np.stack(frames)


def encoder_golden():
    """Trimmed copy of the 116.4 recording's encoder log + timecodes (data, not code) as a fixture
    directory, and what the reference's encoder_velocity / dataset_correlation return on it."""
    from scipy import interpolate
    src_dir = os.path.join(REF, "examples/116.4_mm_min.d")
    fix = os.path.join(OUT, "encoder_116.4")
    os.makedirs(fix, exist_ok=True)
    enc = open(glob.glob(os.path.join(src_dir, "*encoder.csv"))[0]).read().split("\n")[:9000]
    open(os.path.join(fix, "excerpt_encoder.csv"), "w").write("\n".join(enc) + "\n")
    ts = open(glob.glob(os.path.join(src_dir, "*tstamps.txt"))[0]).read().split("\n")[:700]
    open(os.path.join(fix, "excerpt_tstamps.txt"), "w").write("\n".join(ts) + "\n")
    last_frame = 600
    ns = reference_functions({"encoder_velocity", "dataset_correlation", "encoder_out_info_list", "enc_pos_list", "enc_ts_list"},
                             {"imagePath": fix, "lastFrame": last_frame, "interpolate": interpolate})
    pos, t = ns["encoder_velocity"]()
    opt_t = list(np.linspace(0.01, t[-1] * 0.9, 40))
    opt_p = list(np.cumsum(np.linspace(0.05, 0.3, 40)))
    r = ns["dataset_correlation"](opt_p, opt_t, pos, t)
    np.savez_compressed(os.path.join(OUT, "encoder_116.4.npz"), last_frame=last_frame, enc_pos=np.array(pos), enc_ts=np.array(t),
                        opt_t=np.array(opt_t), opt_p=np.array(opt_p), corr_opt=np.asarray(r[0], float), corr_enc=np.asarray(r[1], float),
                        corr_grid=np.asarray(r[2], float), corr_vel_enc=np.asarray(r[3], float), corr_vel_opt=np.asarray(r[4], float))
    return dict(rows=len(pos), last_frame=last_frame)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = reference_functions({"calc_ECC_transform", "calculate_dense_flow", "FB_opt_flow_params"}, {})
    summary = {}
    for name, (directory, first) in CLIPS.items():
        frames = decode(directory, first, N_EXCERPT)
        ts_file = sorted(glob.glob(os.path.join(REF, directory, "*tstamps.txt")))[0]
        ts_lines = open(ts_file).read().split("\n")
        # line L (1-based) of the timecode file holds frame L-2; keep the header and the excerpt's lines
        stamps = np.array([float(ts_lines[first + 1 + i]) for i in range(N_EXCERPT)])

        # --- reference per-pair functions vs the port, consecutive and 4-frame gaps
        pairs = [(i, i + 1) for i in range(0, 12)] + [(i, i + 4) for i in range(0, 32, 4)]
        ecc_ref, ecc_port_out, flow_ref, flow_port = [], [], [], []
        for a, b in pairs:
            r = ref["calc_ECC_transform"](frames[a], frames[b])
            p = ecc_port.ecc_pair_video(frames[a], frames[b])
            ecc_ref.append([float(r[0]), float(r[1]), int(r[2])])
            ecc_port_out.append([float(p[0]), float(p[1]), int(p[2])])
            fr = ref["calculate_dense_flow"](frames[a], frames[b])
            fp = ecc_port.dense_flow_mean(frames[a], frames[b])
            flow_ref.append([float(fr[0]), float(fr[1])])
            flow_port.append([float(fp[0]), float(fp[1])])
        ecc_ref, ecc_port_out = np.array(ecc_ref), np.array(ecc_port_out)
        same = np.array_equal(ecc_ref, ecc_port_out) and np.array_equal(np.array(flow_ref), np.array(flow_port))

        # --- the reference's own end_process on the excerpt (anchor rule, gate, timestamps, mm/min)
        number_list = list(range(first, first + N_EXCERPT + 1))
        g = {"firstFrame": first, "numberList": number_list, "imagePath": os.path.join(REF, directory)}
        ns = reference_functions({"calc_ECC_transform", "calculate_dense_flow", "FB_opt_flow_params", "end_process",
                                  "max_filament_speed", "max_filament_speed_sec", "pixels_per_mm", "velocity_list_x",
                                  "velocity_list_y", "motion_list_opt", "out_information", "tsList"}, g)
        out_information, vx, vy, ts_list, motion = ns["end_process"](frames)

        # --- C SAD oracle pins (spec regression): per-pair records for every listed pair, blocks for two
        sad_rec = []
        for a, b in pairs:
            o = osad.search(frames[a], frames[b], 16, 32)
            sad_rec.append([float(o["pdx"]), float(o["pdy"]), o["flag"], o["nvalid_x"], o["nvalid_y"]])
        o0 = osad.search(frames[0], frames[4], 16, 32)
        o1 = osad.search(frames[8], frames[9], 8, 12)

        np.savez_compressed(
            os.path.join(OUT, f"clip_{name}.npz"), frames=frames, first=first, stamps_ms=stamps,
            pairs=np.array(pairs, np.int32), ecc_ref=ecc_ref, flow_ref=np.array(flow_ref),
            endproc_out_information=np.array(out_information, np.float64).reshape(-1, 3),
            endproc_motion=np.array(motion, np.float64), endproc_ts=np.array(ts_list, np.float64),
            sad_pairs=np.array(sad_rec, np.float64),
            sad_b16r32_mv=o0["mv"], sad_b16r32_sad=o0["sad"], sad_b16r32_sub=o0["sub"], sad_b16r32_flags=o0["flags"],
            sad_b8r12_mv=o1["mv"], sad_b8r12_sad=o1["sad"], sad_b8r12_sub=o1["sub"], sad_b8r12_flags=o1["flags"])
        summary[name] = dict(first=first, n=int(len(frames)), port_matches_reference=bool(same),
                             ecc_flags=[int(v) for v in ecc_ref[:, 2]], endproc_rows=len(out_information),
                             cv2=cv.__version__)
        print(name, summary[name])
    summary["encoder"] = encoder_golden()
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as f:
        json.dump(dict(generated_by="tools/make_golden.py", reference="/root/reference (ruzko/motion-estimation)",
                       clips=summary), f, indent=1)


if __name__ == "__main__":
    main()
