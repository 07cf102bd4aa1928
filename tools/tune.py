"""Kernel timing under different launch-shape overrides (ME_TBX, ME_THREADS, ME_SMEM_CAP). GPU box only."""
import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth

def timing(H, W, B, R, P, env, reps=3):
    for k in ("ME_TBX", "ME_THREADS", "ME_SMEM_CAP", "ME_THREADS_MAX"):
        os.environ.pop(k, None)
    os.environ.update({k: str(v) for k, v in env.items()})
    frames, _ = synth.translating_frames(P + 1, H, W, R)
    try:
        with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R), max_frames=P + 1, max_pairs=P) as est:
            est.upload(frames)
            pairs = [(k, k + 1) for k in range(P)]
            est.search(pairs)
            ms = []
            for _ in range(reps):
                est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
            g = est.geometry
            best = min(ms)
            print(f"{W}x{H} B={B} R={R} P={P} env={env}: tbx={g['tile_blocks_x']} thr={g['threads_per_cta']} smem={g['smem_bytes']} "
                  f"ms={best:.3f} pairs/s={P / best * 1e3:.1f} Tcand.px/s valid={g['cand_px_per_pair_valid'] * P / best / 1e9:.2f} "
                  f"frac={g['cand_px_per_pair_valid'] * P / best / 1e9 / 74.4:.3f}", flush=True)
    except me.MEError as e:
        print(f"env={env}: {e}", flush=True)

if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "hd"
    if which == "hd":
        for tbx in (0, 2, 4, 5, 8, 11, 12):
            for thr in (0, 128, 224, 256, 448):
                env = {}
                if tbx: env["ME_TBX"] = tbx
                if thr: env["ME_THREADS"] = thr
                timing(1080, 1920, 16, 32, 32, env)
    elif which == "4k":
        for tbx in (0, 2, 4, 6, 8, 12):
            for thr in (0, 256, 448):
                env = {}
                if tbx: env["ME_TBX"] = tbx
                if thr: env["ME_THREADS"] = thr
                timing(2160, 3840, 8, 64, 2, env, reps=2)
    elif which == "misc":
        timing(64, 640, 16, 32, 64, {})
        timing(720, 1280, 16, 16, 32, {})
        timing(1080, 1920, 32, 32, 16, {})
        timing(1080, 1920, 16, 64, 8, {})
        timing(1080, 1920, 16, 8, 64, {})
