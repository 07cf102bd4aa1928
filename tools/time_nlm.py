"""Time me_denoise on a batch of frames: python tools/time_nlm.py H W N [T] [smooth]   (T = temporal window for inner frames;
"smooth" = low-noise frames where most patches are similar and every weight lookup is taken)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
H, W, N = map(int, sys.argv[1:4])
T = int(sys.argv[4]) if len(sys.argv) > 4 else 9
rng = np.random.default_rng(0)
frames = rng.integers(0, 256, (N, H, W), dtype=np.uint8)
if len(sys.argv) > 5:
    frames = (128 + rng.normal(0, 4, (N, H, W))).astype(np.uint8)
half = T // 2
windows = [T if half <= k < N - half else 1 for k in range(N)]
with me.MotionEstimator(max(H, 16), max(W, 16), me.SearchConfig(16, 4), max_frames=N, max_pairs=1) as est:
    est.upload(frames)
    best = 1e9
    for _ in range(3):
        est.upload(frames); est.download(0, 1)
        t = time.perf_counter(); est.denoise(0, windows); est.download(0, 1); dt = time.perf_counter() - t
        best = min(best, dt)
    offs = sum(w * 41 * 41 for w in windows) * H * W
    print(f"{W}x{H} N={N} T={T}: {best*1e3:.2f} ms, {best/N*1e3:.3f} ms/frame, {offs/best/1e12:.3f} T offset*px/s")
