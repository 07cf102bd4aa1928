"""NumPy restatement of cv.createCLAHE(clipLimit, tileGridSize).apply for 8-bit images (test infrastructure).

Reference call site: estimate_vel_vid.py:201-219 (``adaptive_histogram_equalization``:
``cv.createCLAHE(clipLimit=2.0, tileGridSize=(8,8)).apply(z)`` per frame).  The arithmetic lives in
OpenCV (third party, not vendored; pinned opencv_contrib_python==4.6.0.66, installed 4.13): per-tile
256-bin histogram, clip at ``max(int(clip * tile_area / 256), 1)``, excess redistributed in a batch plus
a strided residual, LUT = round(cumsum * 255 / tile_area), then bilinear blending of the four
surrounding tile LUTs in float32 with one rounding at the end.  tests/test_oracle.py pins this file
against the installed cv2 bit for bit; the CUDA kernels are then compared with this file.
"""
from __future__ import annotations

import numpy as np


def _reflect101(i, n):
    i = np.where(i >= n, 2 * (n - 1) - i, i)
    return np.abs(i)


def tile_luts(src, clip_limit=2.0, tiles=(8, 8)):
    """(tiles_y, tiles_x, 256) uint8 lookup tables and the (tile_h, tile_w) used."""
    src = np.asarray(src, np.uint8)
    tx, ty = tiles
    H, W = src.shape
    He, We = H, W
    if H % ty or W % tx:                                   # OpenCV pads BOTH axes by (tiles - size % tiles), a whole extra
        He, We = H + ty - H % ty, W + tx - W % tx          # tile count on an axis that was already divisible
    if (He, We) != (H, W):                                 # copyMakeBorder(..., BORDER_REFLECT_101) on the right / bottom
        ext = src[_reflect101(np.arange(He), H)[:, None], _reflect101(np.arange(We), W)[None, :]]
    else:
        ext = src
    th, tw = He // ty, We // tx
    area = tw * th
    lut_scale = np.float32(255.0) / np.float32(area)
    cl = max(int(clip_limit * area / 256), 1) if clip_limit > 0 else 0
    luts = np.zeros((ty, tx, 256), np.uint8)
    for j in range(ty):
        for i in range(tx):
            hist = np.bincount(ext[j * th:(j + 1) * th, i * tw:(i + 1) * tw].ravel(), minlength=256).astype(np.int64)
            if cl > 0:
                clipped = int(np.maximum(hist - cl, 0).sum())
                hist = np.minimum(hist, cl)
                batch = clipped // 256
                resid = clipped - batch * 256
                hist += batch
                if resid:
                    step = max(256 // resid, 1)
                    k = np.arange(256)
                    hist += ((k % step == 0) & (k // step < resid)).astype(np.int64)
            cs = np.cumsum(hist).astype(np.float32) * lut_scale
            luts[j, i] = np.clip(np.rint(cs), 0, 255).astype(np.uint8)
    return luts, (th, tw)


def apply(src, clip_limit=2.0, tiles=(8, 8)):
    src = np.asarray(src, np.uint8)
    tx, ty = tiles
    H, W = src.shape
    luts, (th, tw) = tile_luts(src, clip_limit, tiles)
    one = np.float32(1)
    ys = np.arange(H, dtype=np.float32) * (one / np.float32(th)) - np.float32(0.5)
    xs = np.arange(W, dtype=np.float32) * (one / np.float32(tw)) - np.float32(0.5)
    y1 = np.floor(ys).astype(np.int32)
    x1 = np.floor(xs).astype(np.int32)
    ya = (ys - y1.astype(np.float32)).astype(np.float32)
    xa = (xs - x1.astype(np.float32)).astype(np.float32)
    ya1, xa1 = (one - ya).astype(np.float32), (one - xa).astype(np.float32)
    y2, x2 = np.minimum(y1 + 1, ty - 1), np.minimum(x1 + 1, tx - 1)
    y1, x1 = np.maximum(y1, 0), np.maximum(x1, 0)
    f = lambda a: a.astype(np.float32)
    top = (f(luts[y1[:, None], x1[None, :], src]) * xa1[None, :]).astype(np.float32) + \
          (f(luts[y1[:, None], x2[None, :], src]) * xa[None, :]).astype(np.float32)
    bot = (f(luts[y2[:, None], x1[None, :], src]) * xa1[None, :]).astype(np.float32) + \
          (f(luts[y2[:, None], x2[None, :], src]) * xa[None, :]).astype(np.float32)
    res = (top * ya1[:, None]).astype(np.float32) + (bot * ya[:, None]).astype(np.float32)
    return np.clip(np.rint(res), 0, 255).astype(np.uint8)
