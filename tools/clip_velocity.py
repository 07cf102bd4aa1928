"""Whole example recordings through the drop-in pipeline on the GPU box; prints velocity statistics
against the nominal feed rate in the directory name.  (The recordings are copied to tools/_clips by
hand in the build container; they are data of the reference and are not committed.)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import api, ingest

root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_clips")
for name, rng in (("29.1", 32), ("58.2", 32), ("116.4", 64)):
    d = os.path.join(root, name)
    if not os.path.isdir(d):
        continue
    t0 = time.time()
    frames = ingest.read_frames(d)
    t_dec = time.time() - t0
    lines = open(ingest.find_timestamps(d)).read().split("\n")
    n = min(len(frames), sum(1 for l in lines[1:] if l.strip()) - 1)
    frames = frames[:n]
    if len(sys.argv) > 1 and sys.argv[1] == "preprocess":          # the reference's default pipeline: CLAHE, then non-local means
        t0 = time.time(); frames = api.adaptive_histogram_equalization(frames); t_eq = time.time() - t0
        t0 = time.time(); frames = api.denoise_hf5(frames, n_list=len(frames) + 1); t_nlm = time.time() - t0
        print(f"{name}: CLAHE of {len(frames)} frames {t_eq:.2f}s, NLM (h=15, 19/41, 1/5/9 frames) {t_nlm:.2f}s "
              f"= {t_nlm / len(frames) * 1e3:.2f} ms/frame incl. host copies", flush=True)
    for units in (("fullres",) if len(sys.argv) > 1 else ("reference", "fullres")):
        t0 = time.time()
        res = api.end_process(frames, imagePath=d, firstFrame=0, config=me.SearchConfig(block=16, range=rng),
                              units=units, return_details=True)
        dt = time.time() - t0
        ts = np.array(res.tsList); pos = np.array(res.motion_list_opt)
        vx = np.array(res.velocity_list_x)
        # mean velocity over the moving part: total travel / time between first and last accepted pair
        moving = pos[-1] / ((ts[-1] - ts[0]) / 60.0) if len(ts) > 1 else float("nan")
        # robust: slope over the central 80 % of the travel
        lo, hi = np.searchsorted(pos, [0.1 * pos[-1], 0.9 * pos[-1]])
        slope = (pos[hi - 1] - pos[lo]) / ((ts[hi - 1] - ts[lo]) / 60.0) if hi - 1 > lo else float("nan")
        print(f"{name} mm/min clip: {n} frames decoded in {t_dec:.2f}s; units={units}: end_process {dt:.2f}s "
              f"({res.searches} searches in {res.launches} launches, {len(res.out_information)} accepted pairs, "
              f"{res.failed_estimates} failed); travel {pos[-1]:.2f} mm; mean velocity over the run {moving:.2f} mm/min, "
              f"central-80% slope {slope:.2f} mm/min (nominal {name}); median accepted vx {np.median(vx):.1f}", flush=True)
    api.close_engines()
