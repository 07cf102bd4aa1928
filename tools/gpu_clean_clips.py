#!/usr/bin/env python3
"""GPU box: the reference's pre-processing order (CLAHE, then non-local means; estimate_vel_vid.py:703-709)
on the committed clip excerpts, with the GPU kernels (bit-exact with cv2, tests/test_gpu_clahe.py /
test_gpu_nlm.py).  Writes the cleaned frames, lzma-compressed, to gpurun_out/ so that the build container
can run the reference's own end_process on them (tools/make_golden_clips.py)."""
import lzma
import os
import sys
import time
import zlib

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from motion_estimation_b200 import api, ingest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLIPS = os.path.join(ROOT, "tests", "golden", "clips")
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
import json
man = json.load(open(os.path.join(CLIPS, "MANIFEST.json")))
for name, m in man.items():
    frames = ingest.read_frames(os.path.join(CLIPS, name, "clip.h264"), 0, m["frames"])
    ok = zlib.crc32(frames.tobytes()) == m["crc32_decoded"]
    t0 = time.time()
    eq = api.adaptive_histogram_equalization(frames)
    t1 = time.time()
    clean = api.denoise_hf5(eq, n_list=len(eq) + 1)
    t2 = time.time()
    blob = lzma.compress(clean.tobytes(), preset=6)
    open(os.path.join(OUT, f"clean_{name}.xz"), "wb").write(blob)
    print(f"{name}: {frames.shape} decoded (crc ok: {ok}), CLAHE {t1 - t0:.2f}s, NLM {t2 - t1:.2f}s, {len(clean)} clean frames, "
          f"{len(blob) / 1e6:.1f} MB, crc32 {zlib.crc32(clean.tobytes())}", flush=True)
api.close_engines()
