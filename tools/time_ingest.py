"""Device time of the ingest-side kernels (BGR->luma, CLAHE) on 1080p frames; GPU box only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
H, W, N = 1080, 1920, 64
rng = np.random.default_rng(0)
frames = rng.integers(0, 256, (N, H, W), dtype=np.uint8)
with me.MotionEstimator(H, W, me.SearchConfig(16, 32), max_frames=N, max_pairs=1) as est:
    est.upload(frames); est.sync()
    for rep in range(3):
        t0 = time.perf_counter(); est.clahe(0, N); est.sync(); dt = time.perf_counter() - t0
    print(f"clahe: {N} frames {W}x{H} in {dt * 1e3:.3f} ms (host clock incl. launch) = {N / dt:.0f} frames/s, "
          f"{3 * N * H * W / dt / 1e9:.1f} GB/s of 3 B/px algorithmic traffic")
    bgr = rng.integers(0, 256, (8, H, W, 3), dtype=np.uint8)
    for rep in range(3):
        t0 = time.perf_counter(); est.upload_bgr(bgr, 0); est.sync(); dt = time.perf_counter() - t0
    print(f"upload_bgr: 8 frames in {dt * 1e3:.3f} ms (pageable host source, staging + H2D + kernel) = {8 * H * W * 3 / dt / 1e9:.2f} GB/s")
