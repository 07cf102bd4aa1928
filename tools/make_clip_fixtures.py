#!/usr/bin/env python3
"""Cut >= 2000-frame excerpts out of the reference's example recordings (run in the BUILD container only).

The recordings are H.264 Annex-B elementary streams with one NAL unit per frame and an IDR frame every
60 frames, so an excerpt that starts on an IDR frame is a byte slice: SPS + PPS of the stream head, then
the NAL units of frames [first, first + n).  Nothing is re-encoded; decoding the excerpt gives the same
pixels as decoding the whole file and skipping to `first` (checked below).  The timecode excerpt keeps
the header line and the lines of the same frames, so that running the pair loop on the excerpt with
firstFrame = 0 reads the timestamps the whole recording would give it with firstFrame = first
(estimate_vel_vid.py:498,511,540: line L of the file belongs to frame L - 2).

Output: tests/golden/clips/<name>/clip.h264, clip_tstamps.txt (data of the reference, not code).
"""
import glob
import json
import os
import re
import sys
import zlib

import cv2 as cv
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/examples"
OUT = os.path.join(ROOT, "tests", "golden", "clips")
# name -> (directory, first frame (multiple of 60 = IDR), frames): the stretch where the filament moves
CLIPS = {"29.1": ("29.1_mm_min.d", 600, 2040), "58.2": ("58.2_mm_mind", 600, 2040), "116.4": ("116.4_mm_min.d", 240, 2040)}


def decode(path, first=0, n=None):
    cap = cv.VideoCapture(path)
    out, k = [], 0
    while n is None or len(out) < n:
        ok, f = cap.read()
        if not ok:
            break
        if k >= first:
            out.append(cv.cvtColor(f, cv.COLOR_BGR2GRAY))
        k += 1
    cap.release()
    return np.stack(out)


def main():
    manifest = {}
    for name, (d, first, n) in CLIPS.items():
        src = sorted(glob.glob(os.path.join(REF, d, "*.h264")))[0]
        b = open(src, "rb").read()
        pos = [m.start() for m in re.finditer(b"\x00\x00\x00\x01", b)] + [len(b)]
        types = [b[p + 4] & 31 for p in pos[:-1]]
        assert types[:3] == [7, 8, 5] and types[2 + first] == 5, "excerpt must start on an IDR frame"
        tail = 8                                   # a few more frames than asked for: the decoder holds the last ones back
        cut = b[:pos[2]] + b[pos[2 + first]:pos[2 + first + n + tail]]
        od = os.path.join(OUT, name)
        os.makedirs(od, exist_ok=True)
        open(os.path.join(od, "clip.h264"), "wb").write(cut)
        ts = open(sorted(glob.glob(os.path.join(REF, d, "*tstamps.txt")))[0]).read().split("\n")
        lines = [ts[0]] + ts[first + 1:first + 1 + n + 2]
        open(os.path.join(od, "clip_tstamps.txt"), "w").write("\n".join(lines) + "\n")
        got = decode(os.path.join(od, "clip.h264"), 0, n)
        want = decode(src, first, n)
        assert got.shape == want.shape and np.array_equal(got, want), f"{name}: the excerpt does not decode to the same frames"
        manifest[name] = dict(source=os.path.relpath(src, "/root/reference"), first=first, frames=n, bytes=len(cut),
                              crc32_decoded=zlib.crc32(got.tobytes()), cv2=cv.__version__)
        print(name, manifest[name])
    json.dump(manifest, open(os.path.join(OUT, "MANIFEST.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
