"""The reference's per-pair displacement functions restated on cv2 (test infrastructure).

What the reference actually computes for a frame pair is not a block search but OpenCV's
iterative ECC alignment.  The arithmetic lives in the third-party wheel
``opencv_contrib_python==4.6.0.66`` (reference requirements.txt:6; this image has
cv2 4.13, so version skew is possible and cannot be measured offline).  This module restates
the two *drivers* around that library call so that tests and ``bench.py --impl reference``
can run the reference's CPU path without /root/reference being present:

* ``ecc_pair_video``  follows estimate_vel_vid.py:306-369  (pyramid ECC, half-res result)
* ``ecc_pair_raw``    follows estimate_vel.py:278-297      (single level, warm-started)
* ``dense_flow_mean`` follows estimate_vel_vid.py:375-391  (Farneback frame mean)

``tools/make_golden.py`` checks these against the reference's own function source (exec'd
from /root/reference in the build container) and stores the agreed outputs under
tests/golden/.
"""
from __future__ import annotations

import numpy as np

try:  # cv2 is part of the image; keep the import soft so oracle/sad.py stays usable without it
    import cv2 as cv
except Exception:  # pragma: no cover
    cv = None

# estimate_vel_vid.py:307-314 -- the literals of the video flavour
_LEVELS = 4           # halvings put in front of the full-resolution frame
_PASSES = 4
_CRITERIA_VID = (3, 200, 0.1)   # TERM_CRITERIA_EPS | TERM_CRITERIA_COUNT == 3; eps stays 0.1 (:308-309 build it once)


def _halvings(img, n):
    """[1/2^n, ..., 1/2, 1/1] area-averaged pyramid, coarsest first (estimate_vel_vid.py:318-322)."""
    out = [img]
    for _ in range(n):
        out.append(cv.resize(out[-1], None, fx=0.5, fy=0.5, interpolation=cv.INTER_AREA))
    return out[::-1]


def ecc_pair_video(prev, cur):
    """(pdx, pdy, flag) exactly as the video script's calc_ECC_transform returns them.

    Units are HALF-resolution pixels: the loop visits pyramid entries 0..3 of 5, so the finest
    level aligned is 1/2 scale and its translation is returned unscaled (:330, :342-346, :358).
    """
    pyr_p, pyr_c = _halvings(prev, _LEVELS), _halvings(cur, _LEVELS)
    col2 = np.array([[1, 1, 2], [1, 1, 2]], dtype=np.float32)
    carried = np.eye(2, 3, dtype=np.float32)
    warp = None
    flag = 0
    for _ in range(_PASSES):
        guess = carried * col2 ** (1 - _LEVELS)            # :329 -- translation scaled down to level 0
        for lvl in range(_LEVELS):
            try:
                _, warp = cv.findTransformECC(pyr_p[lvl], pyr_c[lvl], guess, motionType=cv.MOTION_TRANSLATION,
                                              criteria=_CRITERIA_VID)
                flag = 1
            except Exception:
                warp = guess                                # :337-341
                flag = 0
            if lvl != _LEVELS - 1:
                guess = warp * col2                         # :342-343
            else:
                carried = warp                              # :345-346
    return warp[0, 2], warp[1, 2], flag


class RawEcc:
    """estimate_vel.py:278-297: one ECC call per pair, <=1000 iterations, eps 1e-10, and one warp
    matrix object shared by every call (so each pair starts from the previous pair's answer)."""

    def __init__(self):
        self.warp = np.eye(2, 3, dtype=np.float32)

    def __call__(self, prev, cur):
        _, w = cv.findTransformECC(prev, cur, self.warp, motionType=cv.MOTION_TRANSLATION,
                                   criteria=(3, 1000, 1e-10))
        self.warp = w
        return w[0, 2], w[1, 2]


def dense_flow_mean(prev, cur):
    """Frame mean of Farneback flow with the reference's parameters (estimate_vel_vid.py:375-391)."""
    flow = cv.calcOpticalFlowFarneback(prev, cur, None, pyr_scale=0.5, levels=3, winsize=16, iterations=15,
                                       poly_n=5, poly_sigma=1.2, flags=1)
    return float(np.mean(flow[..., 0])), float(np.mean(flow[..., 1]))
