"""Synthetic translating-texture frames for tests and bench.py.

uint8 luma, multi-octave (1/f-like) texture: the sum over pitches s in {1,2,4,...,64} of
sqrt(s) * (a bicubic up-sampling of an N(0,1) grid of pitch s), min-max scaled to 0..255 with
``numpy.random.default_rng(0)``.  Frames are crops of one canvas at per-frame origins, so that
frame k+1 is frame k translated by a known integer vector with |d| <= max_shift on each axis
(``default_rng(1)``).  Blurred white noise is avoided on purpose: the reference's ECC diverges on it.
"""
from __future__ import annotations

import numpy as np


def _upsample(grid, shape):
    try:
        import cv2
        return cv2.resize(grid, (shape[1], shape[0]), interpolation=cv2.INTER_CUBIC)
    except Exception:  # pragma: no cover - cv2 is part of the image
        from scipy import ndimage
        z = ndimage.zoom(grid, (shape[0] / grid.shape[0], shape[1] / grid.shape[1]), order=3)
        return z[:shape[0], :shape[1]]


def texture(height, width, seed=0, octaves=(1, 2, 4, 8, 16, 32, 64)):
    """(height, width) uint8 multi-octave texture."""
    rng = np.random.default_rng(seed)
    acc = np.zeros((height, width), np.float32)
    for s in octaves:
        gh, gw = -(-height // s) + 1, -(-width // s) + 1
        g = rng.standard_normal((gh, gw)).astype(np.float32)
        acc += np.sqrt(np.float32(s)) * (g if s == 1 else _upsample(g, (gh * s, gw * s)))[:height, :width]
    lo, hi = float(acc.min()), float(acc.max())
    return np.clip(np.rint((acc - lo) * (255.0 / (hi - lo))), 0, 255).astype(np.uint8)


def translating_frames(n_frames, height, width, max_shift, seed_texture=0, seed_motion=1):
    """(frames (N,H,W) uint8, shifts (N-1,2) int32) with shifts[k] = (dx,dy) of frame k+1 relative to frame k.

    +dx means the content moved to the right, which is the sign the reference's ECC reports
    (estimate_vel_vid.py:358 on a right-moving filament gives pdx > 0).
    """
    half = max(int(max_shift) // 2, 0)
    canvas = texture(height + 2 * half + 2, width + 2 * half + 2, seed_texture)
    rng = np.random.default_rng(seed_motion)
    e = rng.integers(-half, half + 1, size=(n_frames, 2))
    frames = np.empty((n_frames, height, width), np.uint8)
    for k in range(n_frames):
        ox, oy = half + 1 + int(e[k, 0]), half + 1 + int(e[k, 1])
        frames[k] = canvas[oy:oy + height, ox:ox + width]
    shifts = (e[:-1] - e[1:]).astype(np.int32)
    return frames, shifts


def subpixel_sequence(n_frames, height, width, seed=5, step_lo=0.6, step_hi=2.4, y_amplitude=0.3):
    """(frames (N,H,W) uint8, x (N,) float64, y (N,) float64): a clean texture translating with SUB-PIXEL motion.

    Frame k shows one canvas displaced by (x[k], y[k]) full-resolution pixels (+x = content moves right), every frame
    resampled from the canvas itself with a Lanczos kernel (no accumulated blur).  x advances by uniform random steps
    in [step_lo, step_hi] px per frame, so the sub-pixel phases of the measured displacements are spread evenly; y is a
    slow sinusoid of +-y_amplitude px.  Used by the velocity-accuracy tests: the true displacement between any two frames
    is x[b] - x[a] exactly.
    """
    import cv2
    rng = np.random.default_rng(seed)
    x = np.concatenate([[0.0], np.cumsum(rng.uniform(step_lo, step_hi, n_frames - 1))])
    y = y_amplitude * np.sin(np.arange(n_frames) * 0.21)
    margin = int(x[-1]) + 16
    canvas = texture(height + 16, width + margin + 16, seed).astype(np.float32)
    frames = np.empty((n_frames, height, width), np.uint8)
    for k in range(n_frames):
        m = np.float32([[1, 0, margin + 8 - x[k]], [0, 1, 8 - y[k]]])          # dst(u, v) = canvas(u + m02, v + m12)
        f = cv2.warpAffine(canvas, m, (width, height), flags=cv2.INTER_LANCZOS4 | cv2.WARP_INVERSE_MAP,
                           borderMode=cv2.BORDER_REFLECT101)
        frames[k] = np.clip(np.rint(f), 0, 255).astype(np.uint8)
    return frames, x, y
