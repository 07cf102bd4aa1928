"""BASELINE.json configs[4] on one GPU: 64 concurrent recordings of one shape, one frame pair per recording and
step, all 64 pairs in ONE launch (the ring holds the previous and the current frame of every recording), for
resolution x block x range.  Prints kernel-only and end-to-end (host frames in, host records out) rates.

    python tools/sweep.py [--recordings 64] [--quick]

With several GPUs the recordings shard over the ranks exactly like bench.py's pairs (no data-path collective);
this tool measures one rank's share."""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import motion_estimation_b200 as me
from motion_estimation_b200 import synth

ap = argparse.ArgumentParser()
ap.add_argument("--recordings", type=int, default=64)
ap.add_argument("--quick", action="store_true", help="1080p only")
a = ap.parse_args()
PEAK = 74.4e12
N = a.recordings
print("W,H,B,R,recordings,pairs/s kernel,Tcand*px/s,frac of 74.4T roofline,pairs/s e2e,tile_blocks_x,tile_blocks_y,resident_kernel,"
      "threads,smem,slabs_x,slabs_y,pitch", flush=True)
for (H, W) in (((1080, 1920),) if a.quick else ((720, 1280), (1080, 1920), (2160, 3840))):
    # recording r = frames (2r, 2r+1) of a translating texture: every recording moves by its own shift
    base, _ = synth.translating_frames(2 * N, H, W, 8)
    try:
        import torch
        pinned = torch.from_numpy(base).pin_memory().numpy()
    except Exception:
        pinned = base
    for B in (8, 16, 32):
        for R in (8, 16, 32, 64, 128):
            pairs = [(2 * r, 2 * r + 1) for r in range(N)]
            try:
                with me.MotionEstimator(H, W, me.SearchConfig(block=B, range=R), max_frames=2 * N, max_pairs=N) as est:
                    est.upload(pinned)
                    est.search(pairs)
                    ms = []
                    for _ in range(2):
                        est.launch(pairs); est.sync(); ms.append(est.last_search_ms())
                    t = min(ms) * 1e-3
                    e2e = []
                    for _ in range(2):
                        t0 = time.perf_counter()
                        est.upload(pinned); est.search(pairs)
                        e2e.append(time.perf_counter() - t0)
                    g = est.geometry
                    print(",".join(str(v) for v in (W, H, B, R, N, f"{N / t:.1f}", f"{g['cand_px_per_pair_valid'] * N / t / 1e12:.3f}",
                          f"{g['cand_px_per_pair_valid'] * N / t / PEAK:.3f}", f"{N / min(e2e):.1f}", g["tile_blocks_x"], g["tile_blocks_y"],
                          g["resident_kernel"], g["threads_per_cta"], g["smem_bytes"], g["slabs_x"], g["slabs_y"], g["window_pitch"])), flush=True)
            except me.MEError as e:
                print(f"{W},{H},{B},{R},{N},error,{e}", flush=True)
