"""Repeat launches of a few plans and compare every record with the first launch (and the first pair with the
oracle): catches races in the resident kernel's stage hand-off, the tickets and the cross-slab keys.
    python tools/stress.py [launches]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
from oracle import sad as osad          # checker only

n_launch = int(sys.argv[1]) if len(sys.argv) > 1 else 200
bad = 0
for (H, W, B, R, P) in ((1080, 1920, 16, 32, 64), (1080, 1920, 32, 32, 24), (720, 1280, 16, 48, 24), (1080, 1920, 16, 8, 64),
                        (1080, 1920, 8, 64, 4), (1080, 1920, 16, 100, 2), (64, 640, 16, 32, 64)):
    frames, _ = synth.translating_frames(P + 1, H, W, min(R, 24))
    cfg = me.SearchConfig(block=B, range=R)
    with me.MotionEstimator(H, W, cfg, max_frames=P + 1, max_pairs=P) as est:
        est.upload(frames)
        pairs = [(k, k + 1) for k in range(P)]
        first = est.search(pairs).copy()
        blocks0 = {k: v.copy() for k, v in est.fetch_blocks(P - 1).items()}
        o = osad.search(frames[0], frames[1], B, R)
        ok0 = (first[0]["pdx"] == o["pdx"] and first[0]["pdy"] == o["pdy"] and first[0]["nvalid_x"] == o["nvalid_x"])
        diffs = 0
        for it in range(n_launch):
            res = est.search(pairs)
            if res.tobytes() != first.tobytes():
                diffs += 1
            elif it % 16 == 0:
                b = est.fetch_blocks(P - 1)
                if any(not np.array_equal(b[k], blocks0[k]) for k in blocks0):
                    diffs += 1
        g = est.geometry
        print(f"{W}x{H} B={B} R={R} P={P} resident={g['resident_kernel']} rows={g['tile_blocks_y']} slabs={g['slabs_x']}x{g['slabs_y']}: "
              f"{n_launch} launches, {diffs} differ from the first, pair 0 vs oracle {'ok' if ok0 else 'MISMATCH'}", flush=True)
        bad += diffs + (0 if ok0 else 1)
print("STRESS", "FAILED" if bad else "OK")
sys.exit(1 if bad else 0)
