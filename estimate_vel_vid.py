#!/usr/bin/env python3
"""B200 drop-in for the reference's video pipeline entry point.

Same command line as the reference script of this name (its argparse block is at lines 34-69):

    estimate_vel_vid.py [first last] [-f] [-p DIR] [-c] [-h264]  (+ the options added below)

DIR holds the recording (``video.mkv``, else the first ``*.mkv`` / ``*.h264``), the first
``*tstamps.txt`` and optionally ``*encoder.csv``.  Output is ``velocity_estimates.csv`` with the
reference's header and row layout (its lines 410, 572, 633-636).  What changed is the engine: the
per-pair displacement is the CUDA block search instead of cv.findTransformECC, and the recording is
decoded sequentially into one array.  The reference's CLAHE / non-local-means pre-processing (its lines
201-299, 99 % of its run time) runs on the GPU, bit-exact with the cv2 calls it replaces, when --clahe /
--denoise / --preprocess (both) are given; by default the search sees the decoded frames.
"""
import argparse
import os
import sys
from pathlib import Path

import numpy as np


def parse(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("fileIndex", type=int, nargs="*", help="first and last frame index")
    ap.add_argument("-f", "--full", nargs="?", dest="useWholeVideo", default="False", const="True")
    ap.add_argument("-p", "--path", nargs="?", type=Path, dest="srcDirectory", default="/dev/shm/")
    ap.add_argument("-c", "--continue", nargs="?", dest="continuation", default="False", const="True",
                    help="reuse DIR/images.npy written by a previous run instead of decoding")
    ap.add_argument("-h264", nargs="?", dest="h264", default="False", const="True",
                    help="read the first *.h264 in DIR instead of the .mkv")
    # additions (the reference has no search parameters)
    ap.add_argument("--block", type=int, default=16)
    ap.add_argument("--range", type=int, default=32, dest="search_range")
    ap.add_argument("--units", choices=("reference", "fullres"), default="reference",
                    help="'reference' halves displacements like the reference's half-resolution ECC result")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--out", type=Path, default=None, help="directory for the CSV (default: DIR, else the cwd)")
    ap.add_argument("--clahe", action="store_true", help="apply CLAHE(2.0,(8,8)) per frame like reference :201-219 (GPU kernels, bit-exact with cv2)")
    ap.add_argument("--denoise", action="store_true", help="non-local-means denoise like reference :246-299 (h=15, 19/41 windows, 1/5/9 frames; GPU kernel, bit-exact with cv2)")
    ap.add_argument("--preprocess", action="store_true", help="--clahe and --denoise: the reference's full default pipeline before the pair loop")
    a = ap.parse_args(argv)
    whole = a.useWholeVideo == "True"
    if not whole and len(a.fileIndex) != 2:
        ap.error("give FIRST LAST frame indices, or -f for the whole video")
    return a, whole


def main(argv=None):
    a, whole = parse(argv)
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api, ingest

    src = str(a.srcDirectory)
    first, last = (0, None) if whole else a.fileIndex
    cache = os.path.join(src, "images.npy")
    if a.continuation == "True":
        if not os.path.exists(cache):
            sys.exit(f"-c given but {cache} does not exist")
        frames = np.load(cache)
    else:
        target = src
        if a.h264 == "True":
            hits = sorted(Path(src).glob("*.h264"))
            if not hits:
                sys.exit(f"no *.h264 in {src}")
            target = str(hits[0])
        frames = ingest.read_frames(target, first, last)
        if a.clahe or a.preprocess:                   # reference lines 201-219, on the GPU, bit-exact with cv2's CLAHE
            api.set_defaults(device=a.device)
            frames = api.adaptive_histogram_equalization(frames)
        if a.denoise or a.preprocess:                 # reference lines 246-299: non-local means, bit-exact with cv2's
            api.set_defaults(device=a.device)
            frames = api.denoise_hf5(frames, n_list=len(frames) + 1)      # len(numberList) = frames + 1 (its line 73)
            np.save(cache, frames) if os.access(src, os.W_OK) else None   # the -c cache (the reference keeps images.h5)
    if len(frames) < 2:
        sys.exit("need at least two frames")
    number_list = list(range(first, first + len(frames) + 1))
    cfg = me.SearchConfig(block=a.block, range=a.search_range)
    res = api.end_process(frames, imagePath=src, firstFrame=first, numberList=number_list, config=cfg,
                          device=a.device, units=a.units, return_details=True)
    out_dir = str(a.out) if a.out else (src if os.access(src, os.W_OK) else os.getcwd())
    os.makedirs(out_dir, exist_ok=True)
    from motion_estimation_b200 import postprocess
    postprocess.write_velocity_csv(out_dir, res.out_information)        # reference header + (vx, vy, t) rows (its lines 410, 572)
    # encoder ground truth and the resampled position series, when the recording ships an encoder log (reference 417-487, 589-642)
    import glob
    if glob.glob(os.path.join(src, "*encoder.csv")) and len(res.tsList) >= 3:
        # the reference indexes the timecode file with lastFrame (its line 429); the example files are
        # truncated, so fall back to the last line that still holds a value
        stamp_lines = open(ingest.find_timestamps(src)).read().split("\n")
        last_line = min(first + len(frames), max(i + 1 for i, v in enumerate(stamp_lines) if v.strip() and not v.startswith("#")))
        enc_pos, enc_ts = postprocess.encoder_velocity(src, last_line)
        opt, enc, grid, _, _ = postprocess.dataset_correlation(res.motion_list_opt, res.tsList, enc_pos, enc_ts)
        postprocess.write_positions_csv(out_dir, grid, opt, enc)
        print(f"wrote {os.path.join(out_dir, 'correlated_positions.csv')}")
    pos = res.motion_list_opt[-1] if res.motion_list_opt else 0.0
    print(f"\nfailed motion estimates: {res.failed_estimates} of {len(number_list)}")
    print(f"final position according to block search: {pos}")
    print(f"accepted pairs: {len(res.accepted_pairs)}; searches run: {res.searches} in {res.launches} launches")
    print(f"wrote {os.path.join(out_dir, 'velocity_estimates.csv')}")
    return res


if __name__ == "__main__":
    main()
