#!/usr/bin/env python3
"""B200 drop-in for the reference's image-sequence entry point.

    estimate_vel.py first last [-p DIR] [-d]

The reference script of this name demosaics raspiraw ``out.NNNN.raw`` files (its lines 64-128; needs
rawpy and a sensor header that do not ship) and then compares strictly consecutive frames (lines
333-378).  This entry point keeps the second half: frames first..last of the recording in DIR (an
image sequence, or the video an example directory holds), consecutive pairs through the CUDA block
search, timestamps from ``tstamps.csv`` (column 0, microseconds, reference line 352-354) or, when
only a ``*tstamps.txt`` timecode file exists, from that, and ``velocity_estimates.csv`` with the
reference's columns (lines 319, 374, 382-385).
"""
import argparse
import csv
import os
import sys
from pathlib import Path


def parse(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("fileIndex", type=int, nargs=2, help="first and last image index")
    ap.add_argument("-p", "--path", nargs="?", type=Path, dest="srcDirectory", default="/dev/shm/")
    ap.add_argument("-d", nargs="?", dest="doLineDoubling", default="False", const="True",
                    help="accepted for compatibility; unused here as in the reference")
    ap.add_argument("--block", type=int, default=16)
    ap.add_argument("--range", type=int, default=64, dest="search_range")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--out", type=Path, default=None)
    return ap.parse_args(argv)


def load_timestamps_us(src, first, n):
    """Microsecond stamp per frame index first..first+n-1, read the way reference line 352 indexes them."""
    p = os.path.join(src, "tstamps.csv")
    if os.path.exists(p):
        lines = open(p).read().split("\n")
        return [int(lines[k - 1].split(",")[0]) for k in range(first, first + n)]   # linecache is 1-based
    from motion_estimation_b200 import ingest
    t = ingest.find_timestamps(src)
    if t is None:
        sys.exit(f"neither tstamps.csv nor *tstamps.txt in {src}")
    vals = [float(v) for v in open(t).read().split("\n")[1:] if v.strip()]
    return [int(round(vals[k] * 1000.0)) for k in range(first, first + n)]          # ms -> us


def main(argv=None):
    a = parse(argv)
    import motion_estimation_b200 as me
    from motion_estimation_b200 import api, ingest

    src = str(a.srcDirectory)
    first, last = a.fileIndex
    frames = ingest.read_frames(src, first, last + 1)
    if len(frames) < 2:
        sys.exit("need at least two frames")
    stamps = load_timestamps_us(src, first, len(frames))
    cfg = me.SearchConfig(block=a.block, range=a.search_range)
    res = api.end_process_raw(frames, stamps, config=cfg, device=a.device, return_details=True)
    out_dir = str(a.out) if a.out else (src if os.access(src, os.W_OK) else os.getcwd())
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, "velocity_estimates.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["mm/min X-axis", "mm/min Y-axis", "Timestamp [s]"])
        w.writerows(res.out_information)
    print(f"{len(res.out_information)} pairs -> {os.path.join(out_dir, 'velocity_estimates.csv')}")
    return res


if __name__ == "__main__":
    main()
