"""Independent NumPy restatement of the spec in oracle/sad_oracle.c (test infrastructure).

Written array-first (whole cost volumes, numpy.argmin, numpy.sort) so that it shares no
control flow with the C oracle: the two are cross-checked in tests/test_oracle.py, and the
CUDA path is then compared with the C one.  Use on small frames only (memory ~ nb*(2R+1)^2).
"""
from __future__ import annotations

import numpy as np

FLAG_TEXTURED = 1
FLAG_X_INTERIOR = 2
FLAG_Y_INTERIOR = 4
FLAG_X_PEAKED = 8
FLAG_Y_PEAKED = 16
_INF = np.iinfo(np.int64).max


def cost_volume(prev, cur, block, rng, range_y=0):
    """(nby, nbx, 2R+1, 2R+1) int64 SAD volume indexed [by, bx, dy+R, dx+R]; out-of-domain = _INF."""
    prev = np.asarray(prev, np.uint8).astype(np.int64)
    cur = np.asarray(cur, np.uint8).astype(np.int64)
    H, W = prev.shape
    B, R = block, rng
    U = 2 * R + 1
    nbx, nby = W // B, H // B
    vol = np.full((nby, nbx, U, U), _INF, np.int64)
    ys = np.arange(nby) * B
    xs = np.arange(nbx) * B
    ref = prev[: nby * B, : nbx * B].reshape(nby, B, nbx, B)
    Ry = range_y if range_y > 0 else R
    for dy in range(-Ry, Ry + 1):
        oky = (ys + dy >= 0) & (ys + dy + B <= H)
        if not oky.any():
            continue
        for dx in range(-R, R + 1):
            okx = (xs + dx >= 0) & (xs + dx + B <= W)
            if not okx.any():
                continue
            # shifted view of cur, zero where the shifted block would leave the frame (masked below)
            sh = np.zeros((nby * B, nbx * B), np.int64)
            y0, y1 = max(0, -dy), min(nby * B, H - dy)
            x0, x1 = max(0, -dx), min(nbx * B, W - dx)
            sh[y0:y1, x0:x1] = cur[y0 + dy:y1 + dy, x0 + dx:x1 + dx]
            sad = np.abs(sh.reshape(nby, B, nbx, B) - ref).sum(axis=(1, 3))
            m = oky[:, None] & okx[None, :]
            vol[:, :, dy + R, dx + R] = np.where(m, sad, _INF)
    return vol


def _median_spec(v):
    s = np.sort(np.asarray(v, np.float32))
    n = s.size
    return np.float32(0.5) * np.float32(s[(n - 1) // 2] + s[n // 2])


def search(prev, cur, block=16, rng=32, min_range=2, subpixel=True, min_valid=1, min_curv=1, range_y=0):
    prev = np.asarray(prev, np.uint8)
    vol = cost_volume(prev, cur, block, rng, range_y)
    nby, nbx, U, _ = vol.shape
    B, R = block, rng
    min_curv = max(int(min_curv), 1)
    flat = vol.reshape(nby, nbx, U * U)
    rank = np.argmin(flat, axis=2)                      # first minimum, C order == raster (dy outer)
    iy, ix = rank // U, rank % U
    best = np.take_along_axis(flat, rank[..., None], 2)[..., 0]
    mv = np.stack([ix - R, iy - R], -1).astype(np.int16)

    def at(dyi, dxi):
        ok = (dyi >= 0) & (dyi < U) & (dxi >= 0) & (dxi < U)
        c = vol[np.arange(nby)[:, None], np.arange(nbx)[None, :], np.clip(dyi, 0, U - 1), np.clip(dxi, 0, U - 1)]
        return np.where(ok, c, _INF)

    blk = prev[: nby * B, : nbx * B].reshape(nby, B, nbx, B)
    spread = blk.max(axis=(1, 3)).astype(np.int32) - blk.min(axis=(1, 3)).astype(np.int32)
    textured = spread >= min_range
    flags = (textured * FLAG_TEXTURED).astype(np.uint8)
    sub = mv.astype(np.float32)
    valid = []
    for axis, (cm, cp, f_int, f_peak) in enumerate((
            (at(iy, ix - 1), at(iy, ix + 1), FLAG_X_INTERIOR, FLAG_X_PEAKED),
            (at(iy - 1, ix), at(iy + 1, ix), FLAG_Y_INTERIOR, FLAG_Y_PEAKED))):
        interior = (cm != _INF) & (cp != _INF)
        cm_ = np.where(interior, cm, 0)
        cp_ = np.where(interior, cp, 0)
        den = np.where(interior, cm_ - 2 * best + cp_, 0)
        num = np.where(interior, cm_ - cp_, 0)
        peaked = interior & (den >= min_curv)
        flags |= (interior * f_int + peaked * f_peak).astype(np.uint8)
        if subpixel:
            q = num.astype(np.float32) / np.where(den > 0, den, 1).astype(np.float32)
            off = np.where(den > 0, np.float32(0.5) * q, np.float32(0)).astype(np.float32)
            sub[..., axis] = (sub[..., axis] + off).astype(np.float32)
        valid.append(textured & peaked)

    nx, ny = int(valid[0].sum()), int(valid[1].sum())
    pdx = _median_spec(sub[..., 0][valid[0]]) if nx else np.float32(0)
    pdy = _median_spec(sub[..., 1][valid[1]]) if ny else np.float32(0)
    flag = int(nx >= max(min_valid, 1))

    def mean_spec(v):                       # exact fixed-point sum, one double division (spec: mean)
        if v.size == 0:
            return np.float32(0)
        fixed = np.rint(v.astype(np.float32) * np.float32(1048576.0)).astype(np.int64)
        return np.float32(np.float64(int(fixed.sum())) / (np.float64(v.size) * np.float64(1048576.0)))
    return dict(mv=mv, sad=best.astype(np.uint32), sub=sub, flags=flags, pdx=pdx, pdy=pdy, flag=flag,
                nvalid_x=nx, nvalid_y=ny, n_textured=int(textured.sum()),
                mean_dx=mean_spec(sub[..., 0][valid[0]]), mean_dy=mean_spec(sub[..., 1][valid[1]]))
