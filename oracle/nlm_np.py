"""NumPy restatement of cv.fastNlMeansDenoising / cv.fastNlMeansDenoisingMulti for 8-bit grey images
(test infrastructure; nothing in the product path imports this).

Reference call sites: estimate_vel_vid.py:226-232 (``denoising``: Multi, h=15, templateWindowSize=19,
searchWindowSize=41, temporal window 5 or 9) and :246-262 (``denoise_hf5``: which frames get the single-image
call, which the 5- and the 9-frame window).  The arithmetic lives in OpenCV's photo module (third party, not
vendored; the reference pins opencv_contrib_python==4.6.0.66, this image has 4.13), restated here from its
published algorithm:

* every frame of the temporal window is extended by ``search/2 + template/2`` pixels with BORDER_REFLECT_101;
* for every pixel and every (frame, dy, dx) of the search volume, ``dist`` = sum over the template window of
  squared differences between the patch around the pixel in the frame being denoised and the displaced patch;
* ``weight = table[dist >> shift]`` with ``1 << shift`` the power of two >= template area, and
  ``table[a] = cvRound(fixed * exp(-(a * 2^shift / area) / (h*h)))`` zeroed below ``0.001 * fixed``, where
  ``fixed = INT_MAX // (frames * search^2 * 255)`` (32-bit accumulators; ``h*h`` is a float32 product);
* result = ``(sum(weight * pixel) + sum(weight) // 2) // sum(weight)``.

tests/test_oracle.py pins this file against the installed cv2 bit for bit (single image, 5- and 9-frame
windows, frames smaller than the border); the CUDA kernel is compared with cv2 and with this file.
"""
from __future__ import annotations

import numpy as np

INT_MAX = 2 ** 31 - 1


def reflect101(i, n):
    """cv::borderInterpolate(i, n, BORDER_REFLECT_101), iterated like OpenCV does for borders wider than the image."""
    i = np.asarray(i).copy()
    if n == 1:
        return np.zeros_like(i)
    while True:
        bad = (i < 0) | (i >= n)
        if not bad.any():
            return i
        i = np.where(i < 0, -i, i)
        i = np.where(i >= n, 2 * n - 2 - i, i)


def weight_table(h, template_window, search_window, n_frames):
    """(table int64 truncated after the last non-zero weight, shift, fixed_point_mult)."""
    tws, sws = 2 * (template_window // 2) + 1, 2 * (search_window // 2) + 1
    fixed = min(INT_MAX // (n_frames * sws * sws * 255), INT_MAX)
    area = tws * tws
    shift = 0
    while (1 << shift) < area:
        shift += 1
    mult = float(1 << shift) / area
    n = int(65025 / mult + 1)
    dist = np.arange(n, dtype=np.float64) * mult
    hh = np.float32(h) * np.float32(h)                      # float product, as in calcWeight
    w = np.exp(-dist / np.float64(hh))
    table = np.rint(fixed * w).astype(np.int64)             # cvRound: round half to even
    table[table < 0.001 * fixed] = 0
    nz = np.nonzero(table)[0]
    return table[: (nz[-1] + 1 if len(nz) else 1)], shift, fixed


def denoise(frames, index, temporal_window=1, h=15.0, template_window=19, search_window=41):
    """frames (N,H,W) uint8 -> denoised frame ``index`` (H,W) uint8.  temporal_window 1 == fastNlMeansDenoising."""
    frames = np.asarray(frames, np.uint8)
    if frames.ndim == 2:
        frames = frames[None]
    N, H, W = frames.shape
    th, sh = template_window // 2, search_window // 2
    tws = 2 * th + 1
    half = temporal_window // 2
    if temporal_window % 2 == 0 or index - half < 0 or index + half >= N:
        raise ValueError("temporal window must be odd and lie inside the sequence")   # cv2 raises here as well
    b = th + sh
    rows, cols = reflect101(np.arange(-b, H + b), H), reflect101(np.arange(-b, W + b), W)
    ext = [frames[t][rows[:, None], cols[None, :]].astype(np.int64) for t in range(index - half, index + half + 1)]
    table, shift, _ = weight_table(h, template_window, search_window, temporal_window)
    table = np.concatenate([table, [0]])
    main = ext[half][sh:sh + H + 2 * th, sh:sh + W + 2 * th]
    est = np.zeros((H, W), np.int64)
    wsum = np.zeros((H, W), np.int64)
    ii = np.zeros((H + 2 * th + 1, W + 2 * th + 1), np.int64)
    for e in ext:
        for dy in range(-sh, sh + 1):
            for dx in range(-sh, sh + 1):
                sq = (main - e[sh + dy:sh + dy + H + 2 * th, sh + dx:sh + dx + W + 2 * th]) ** 2
                ii[1:, 1:] = sq.cumsum(0).cumsum(1)
                d = ii[tws:, tws:] - ii[:-tws, tws:] - ii[tws:, :-tws] + ii[:-tws, :-tws]
                wgt = table[np.minimum(d >> shift, len(table) - 1)]
                est += wgt * e[b + dy:b + dy + H, b + dx:b + dx + W]
                wsum += wgt
    return ((est + wsum // 2) // wsum).astype(np.uint8)


def reference_windows(n_frames, n_list=None):
    """Temporal window per frame as denoise_hf5 picks it (estimate_vel_vid.py:250-262); n_list = len(numberList)."""
    L = n_frames if n_list is None else n_list
    out = []
    for k in range(n_frames):
        if k <= 1 or k >= L - 3:
            out.append(1)
        elif k <= 4 or k >= L - 5:
            out.append(5)
        else:
            out.append(9)
    return out
