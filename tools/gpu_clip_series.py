#!/usr/bin/env python3
"""GPU box: the drop-in pipeline (decode -> CLAHE -> NLM -> end_process) on the committed clip excerpts; writes the
series to gpurun_out/series_<name>.npz for analysis in the build container (tools/analyse_clips.py)."""
import json
import os
import sys
import time
import zlib

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import api, ingest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLIPS = os.path.join(ROOT, "tests", "golden", "clips")
OUT = os.path.join(ROOT, "gpurun_out")
man = json.load(open(os.path.join(CLIPS, "MANIFEST.json")))
for name, m in man.items():
    d = os.path.join(CLIPS, name)
    raw = ingest.read_frames(os.path.join(d, "clip.h264"), 0, m["frames"])
    clean = api.denoise_hf5(api.adaptive_histogram_equalization(raw), n_list=len(raw) + 1)
    print(name, "clean crc ok:", zlib.crc32(clean.tobytes()) == m.get("crc32_clean"), flush=True)
    out = {}
    for label, frames in (("clean", clean), ("raw", raw)):
        for rng in (32, 64, 128):
            t0 = time.time()
            res = api.end_process(frames, imagePath=d, firstFrame=0, numberList=list(range(len(frames) + 1)),
                                  config=me.SearchConfig(16, rng), return_details=True)
            dt = time.time() - t0
            key = f"{label}_r{rng}"
            out[key + "_info"] = np.array(res.out_information, np.float64).reshape(-1, 3)
            out[key + "_motion"] = np.array(res.motion_list_opt, np.float64)
            out[key + "_pairs"] = np.array(res.accepted_pairs, np.int32).reshape(-1, 2)
            out[key + "_disp"] = np.array(res.pair_displacements, np.float64).reshape(-1, 3)
            out[key + "_dense"] = np.array([res.final_position_dense])
            print(f"  {key}: {len(res.out_information)} rows, travel {res.motion_list_opt[-1] if res.motion_list_opt else 0:.3f} mm, "
                  f"dense {res.final_position_dense:.3f} mm, failed {res.failed_estimates}, {res.searches} searches / {res.launches} launches, {dt:.3f}s",
                  flush=True)
    np.savez_compressed(os.path.join(OUT, f"series_{name}.npz"), **out)
api.close_engines()
