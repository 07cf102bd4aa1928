"""One search shape, a few launches (for ncu captures): python tools/one_shape.py H W B R PAIRS [range_y]. GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
H, W, B, R, P = (int(v) for v in sys.argv[1:6])
ry = int(sys.argv[6]) if len(sys.argv) > 6 else 0
frames, _ = synth.translating_frames(P + 1, H, W, min(R, 12))
with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R, range_y=ry), max_frames=P + 1, max_pairs=P) as est:
    est.upload(frames)
    pairs = [(k, k + 1) for k in range(P)]
    for _ in range(4):
        est.launch(pairs)
        est.sync()
    g = est.geometry
    ms = est.last_search_ms()
    print(f"{W}x{H} B={B} R={R} ry={ry} P={P}: {ms:.3f} ms, frac {g['cand_px_per_pair_valid'] * P / ms / 1e9 / 74.4:.3f}, geometry {g}")
