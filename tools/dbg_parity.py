import os, sys
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth
from oracle import sad as osad
def parity(H, W, B, R, n=4):
    frames, shifts = synth.translating_frames(n, H, W, min(R, 24))
    cfg = me.SearchConfig(block=B, range=R)
    with me.MotionEstimator(H, W, cfg, max_frames=n, max_pairs=n + 1) as est:
        est.upload(frames)
        pairs = [(0, 1), (1, 2), (2, 3), (3, 0), (2, 2)]
        res = est.search(pairs)
        print(est.geometry)
        for i, (a, c) in enumerate(pairs):
            o = osad.search(frames[a], frames[c], B, R)
            g = est.fetch_blocks(i)
            for k in ("mv", "sad", "sub", "flags"):
                if not np.array_equal(o[k], g[k]):
                    d = np.argwhere(o[k] != g[k])
                    print(f"  MISMATCH {k} pair {i}: {len(d)} of {o[k].size}, first {d[:4].tolist()} oracle {o[k][tuple(d[0])]} gpu {g[k][tuple(d[0])]}")
            r = res[i]
            print(i, r, o['pdx'], o['pdy'], o['nvalid_x'], o['nvalid_y'])
parity(*map(int, sys.argv[1:5]))
