"""Frame ingest: recording on disk -> (N, H, W) uint8 luma (or BGR for GPU-side conversion).

Reference: estimate_vel_vid.py:78-195 (``get_meta``, ``read_vid_mkv``, the broken ``read_vid_h264``).
Differences that matter for speed, not for pixels: frames are decoded sequentially (the reference
seeks before every read, :171), written into one preallocated array (the reference re-allocates the
whole stack per frame with np.append, :189), and the raw ``.h264`` elementary stream is opened
directly through OpenCV's FFMPEG backend instead of the ``h264decoder`` module (:14, :128-160).
"""
from __future__ import annotations

import glob
import os
from typing import Optional, Tuple

import numpy as np

VIDEO_PATTERNS = ("video.mkv", "*.mkv", "*.h264", "*.mp4", "*.avi")
IMAGE_PATTERNS = ("*.png", "*.tif", "*.tiff", "*.jpg", "*.jpeg", "*.bmp")


def find_recording(path: str) -> Tuple[str, str]:
    """('video', file) or ('images', directory) for a file or an example directory such as
    examples/29.1_mm_min.d (video.mkv preferred like estimate_vel_vid.py:166, else the .h264)."""
    if os.path.isfile(path):
        return "video", path
    if not os.path.isdir(path):
        raise FileNotFoundError(path)
    for pat in VIDEO_PATTERNS:
        hits = sorted(glob.glob(os.path.join(path, pat)))
        if hits:
            return "video", hits[0]
    for pat in IMAGE_PATTERNS:
        if glob.glob(os.path.join(path, pat)):
            return "images", path
    raise FileNotFoundError(f"no video or image files in {path}")


def get_meta(path: str):
    """(width, height, totalFrames) as estimate_vel_vid.py:78-90 reads them (frame count may be 0 or
    wrong for a raw .h264 stream, which carries no index)."""
    import cv2 as cv
    kind, target = find_recording(path)
    if kind == "images":
        files = list_images(target)
        im = cv.imread(files[0], cv.IMREAD_GRAYSCALE)
        return im.shape[1], im.shape[0], len(files)
    cap = cv.VideoCapture(target)
    try:
        return int(cap.get(3)), int(cap.get(4)), int(cap.get(7))
    finally:
        cap.release()


def list_images(directory: str):
    files = []
    for pat in IMAGE_PATTERNS:
        files.extend(glob.glob(os.path.join(directory, pat)))
    return sorted(files)


def read_frames(path: str, first: int = 0, last: Optional[int] = None, bgr: bool = False) -> np.ndarray:
    """Frames [first, last) of a recording as (N,H,W) uint8 luma, or (N,H,W,3) BGR when ``bgr`` (for
    ``MotionEstimator.upload_bgr``, which applies cv.cvtColor's BGR2GRAY arithmetic on the GPU)."""
    import cv2 as cv
    kind, target = find_recording(path)
    if kind == "images":
        files = list_images(target)[first:last]
        flag = cv.IMREAD_COLOR if bgr else cv.IMREAD_GRAYSCALE
        return np.stack([cv.imread(f, flag) for f in files])
    cap = cv.VideoCapture(target)
    if not cap.isOpened():
        raise IOError(f"cannot open {target}")
    try:
        W, H = int(cap.get(3)), int(cap.get(4))
        count = int(cap.get(7))
        cap_n = None if last is None else max(last - first, 0)
        guess = cap_n if cap_n is not None else (count - first if count > first else 1024)
        shape = (max(guess, 1), H, W, 3) if bgr else (max(guess, 1), H, W)
        out = np.empty(shape, np.uint8)
        n = 0
        idx = 0
        while cap_n is None or n < cap_n:
            ok, frame = cap.read()
            if not ok:
                break
            if idx >= first:
                if n == out.shape[0]:                      # unknown length: grow geometrically
                    out = np.concatenate([out, np.empty_like(out)], axis=0)
                if bgr:
                    out[n] = frame
                else:
                    cv.cvtColor(frame, cv.COLOR_BGR2GRAY, dst=out[n])   # estimate_vel_vid.py:178
                n += 1
            idx += 1
        return out[:n]
    finally:
        cap.release()


def find_timestamps(directory: str) -> Optional[str]:
    hits = sorted(glob.glob(os.path.join(directory, "*tstamps.txt")))   # estimate_vel_vid.py:516
    return hits[0] if hits else None
