"""Kernel-only throughput over the shapes of BASELINE.json's sweep config (one GPU): resolution x block x range."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import motion_estimation_b200 as me
from motion_estimation_b200 import synth

PEAK = 74.4e12
rows = []
for (H, W) in ((720, 1280), (1080, 1920), (2160, 3840)):
    base, _ = synth.translating_frames(65, H, W, 8)
    for B in (8, 16, 32):
        for R in (8, 16, 32, 64, 128):
            P = int(max(2, min(64, round(2e11 / ((W // B) * (H // B) * B * B * (2 * R + 1) ** 2)))))   # ~3 ms launches
            try:
                with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R), max_frames=P + 1, max_pairs=P) as est:
                    est.upload(base[:P + 1])
                    pairs = [(k, k + 1) for k in range(P)]
                    est.search(pairs)
                    ms = []
                    for _ in range(2):
                        est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
                    g = est.geometry
                    t = min(ms) * 1e-3
                    rows.append((W, H, B, R, P / t, g["cand_px_per_pair_valid"] * P / t / 1e12, g["cand_px_per_pair_valid"] * P / t / PEAK,
                                 P, g["tile_blocks_x"], g["tile_blocks_y"], g["resident_kernel"], g["threads_per_cta"], g["smem_bytes"], g["slabs_x"], g["slabs_y"], g["window_pitch"]))
            except me.MEError as e:
                rows.append((W, H, B, R, 0, 0, 0, str(e)))
            r = rows[-1]
            print(",".join(f"{v:.3f}" if isinstance(v, float) else str(v) for v in r), flush=True)
