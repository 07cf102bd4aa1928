"""Run one search configuration a few times (for ncu captures): python tools/run_one.py H W B R P [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
H, W, B, R, P = map(int, sys.argv[1:6])
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 3
frames, _ = synth.translating_frames(P + 1, H, W, min(R, 32))
with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R), max_frames=P + 1, max_pairs=P) as est:
    est.upload(frames)
    pairs = [(k, k + 1) for k in range(P)]
    ms = []
    for _ in range(reps):
        est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
    g = est.geometry
    print(f"{W}x{H} B={B} R={R} P={P}: ms={min(ms):.3f} frac={g['cand_px_per_pair_valid'] * P / min(ms) / 1e9 / 74.4:.3f} {g}")
