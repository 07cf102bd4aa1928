"""Tile-shape sweep for the small-range / small-block shapes (ME_TBX x ME_TMR x ME_PITCH overrides). GPU box only."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motion_estimation_b200 as me
from motion_estimation_b200 import synth

KEYS = ("ME_TBX", "ME_TMR", "ME_PITCH", "ME_PERSISTENT", "ME_THREADS")

def timing(frames, H, W, B, R, P, env):
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R), max_frames=P + 1, max_pairs=P) as est:
            est.upload(frames)
            pairs = [(k, k + 1) for k in range(P)]
            est.search(pairs)
            ms = []
            for _ in range(3):
                est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
            g = est.geometry
            best = min(ms)
            return g["cand_px_per_pair_valid"] * P / best / 1e9 / 74.4, best
    except me.MEError as e:
        return None, str(e)[:60]

for (H, W, B, R, P) in ((1080, 1920, 16, 8, 64), (1080, 1920, 16, 16, 64), (1080, 1920, 8, 8, 64), (1080, 1920, 8, 16, 64), (1080, 1920, 32, 16, 64)):
    frames, _ = synth.translating_frames(P + 1, H, W, min(R, 8))
    f0, ms0 = timing(frames, H, W, B, R, P, {})
    print(f"{W}x{H} B={B} R={R}: default frac {f0:.3f} ({ms0:.3f} ms)", flush=True)
    res = []
    nsub = 2 if B == 8 else 1
    for pitch in (144, 208, 240):
        for tbx in (2, 3, 4, 5, 6, 7, 8, 10, 12, 13):
            for tmr in (1, 2, 3, 4, 6, 8):
                if tbx * nsub * tmr > 128:
                    continue
                f, ms = timing(frames, H, W, B, R, P, {"ME_TBX": tbx, "ME_TMR": tmr, "ME_PITCH": pitch})
                if f is not None:
                    res.append((f, pitch, tbx, tmr, ms))
    res.sort(reverse=True)
    for r in res[:6]:
        print("   frac %.3f pitch %d tbx(macro-blocks) %d tmr %d ms %.3f" % r, flush=True)
