"""Drop-in functions with the reference's names, arguments and return values.

=====================================  ==================================================
reference                              here
=====================================  ==================================================
estimate_vel_vid.py:306 calc_ECC_transform(prev, cur) -> (pdx, pdy, flag)   ``calc_ECC_transform``
estimate_vel.py:286     calc_ECC_transform(prev, cur) -> (pdx, pdy)         ``calc_ECC_transform_raw``
estimate_vel_vid.py:493 end_process(clean_arrs) -> 5 lists                  ``end_process``
estimate_vel.py:333-378 module-level pair loop                              ``end_process_raw``
estimate_vel_vid.py:385 calculate_dense_flow(prev, cur) -> (mean u, mean v) ``calculate_dense_flow``
=====================================  ==================================================

The displacement comes from the CUDA block search (libme_b200.so), never from a CPU path: if the
library is missing these functions raise ImportError, and CUDA errors propagate as MEError.

Units.  The video script's ECC result is in HALF-resolution pixels (its pyramid loop stops one
level short, estimate_vel_vid.py:330,342-346) and everything downstream (gate ``pdx < 2`` :531,
``/ 611`` :556) is applied to that number.  ``units="reference"`` (default) therefore returns
full-resolution SAD displacements divided by 2; ``units="fullres"`` returns them as measured.
"""
from __future__ import annotations

import glob
import os
from typing import Optional, Sequence

import numpy as np

from . import velocity as _vel
from .engine import FLAG_TEXTURED, FLAG_X_PEAKED, FLAG_Y_PEAKED, MotionEstimator, SearchConfig

# Module-level state mirrors the reference's globals (estimate_vel_vid.py:73, :396-412).
pixels_per_mm = _vel.PIXELS_PER_MM
max_filament_speed = _vel.MAX_FILAMENT_SPEED_MM_MIN
default_config = SearchConfig(block=16, range=32)
default_device = 0

_engines = {}


def set_defaults(config: Optional[SearchConfig] = None, device: Optional[int] = None):
    """Choose the block size / search range used by the drop-in functions (the reference has no such knobs)."""
    global default_config, default_device
    if config is not None:
        default_config = config
    if device is not None:
        default_device = device


def _engine(height, width, config, device, max_frames, max_pairs) -> MotionEstimator:
    key = (height, width, config, device)
    eng = _engines.get(key)
    if eng is not None and (eng.max_frames < max_frames or eng.max_pairs < max_pairs):
        eng.close()
        eng = None
    if eng is None:
        eng = MotionEstimator(height, width, config, max_frames=max(max_frames, 2), max_pairs=max(max_pairs, 1),
                              device=device)
        _engines[key] = eng
    return eng


def close_engines():
    for e in _engines.values():
        e.close()
    _engines.clear()


def _scale(units):
    if units == "reference":
        return 0.5
    if units == "fullres":
        return 1.0
    raise ValueError("units must be 'reference' or 'fullres'")


def calc_ECC_transform(prevFrame, curFrame, *, config: Optional[SearchConfig] = None, device: Optional[int] = None,
                       units: str = "reference"):
    """(pdx, pdy, measurement_flag) for one frame pair -- estimate_vel_vid.py:306-369.

    ``measurement_flag`` is 1 when at least ``config.min_valid`` blocks entered the median (the
    reference sets 0 when ECC throws, :337-341).  Inputs are not modified.
    """
    prev = np.ascontiguousarray(prevFrame, dtype=np.uint8)
    cur = np.ascontiguousarray(curFrame, dtype=np.uint8)
    if prev.ndim != 2 or prev.shape != cur.shape:
        raise ValueError("calc_ECC_transform expects two (H, W) uint8 frames of equal shape")
    cfg = config or default_config
    eng = _engine(prev.shape[0], prev.shape[1], cfg, default_device if device is None else device, 2, 1)
    r = eng.estimate_pair(prev, cur)
    s = np.float32(_scale(units))
    return np.float32(r["pdx"] * s), np.float32(r["pdy"] * s), int(r["flag"])


def calc_ECC_transform_raw(prevFrame, curFrame, **kw):
    """(pdx, pdy) in full-resolution pixels -- the raw-image script's flavour, estimate_vel.py:286-297."""
    kw.setdefault("units", "fullres")
    pdx, pdy, _ = calc_ECC_transform(prevFrame, curFrame, **kw)
    return pdx, pdy


def calculate_dense_flow(prevFrame, curFrame, *, config: Optional[SearchConfig] = None, device: Optional[int] = None):
    """Frame mean of the per-block displacement field -- the role of estimate_vel_vid.py:385-391.

    The reference averages a Farneback flow field; here the dense field is the block search's own
    refined vectors (valid blocks only), so this is a second opinion that costs no extra search.
    """
    prev = np.ascontiguousarray(prevFrame, dtype=np.uint8)
    cur = np.ascontiguousarray(curFrame, dtype=np.uint8)
    cfg = config or default_config
    eng = _engine(prev.shape[0], prev.shape[1], cfg, default_device if device is None else device, 2, 1)
    eng.estimate_pair(prev, cur)
    b = eng.fetch_blocks(0)
    fx = (b["flags"] & (FLAG_TEXTURED | FLAG_X_PEAKED)) == (FLAG_TEXTURED | FLAG_X_PEAKED)
    fy = (b["flags"] & (FLAG_TEXTURED | FLAG_Y_PEAKED)) == (FLAG_TEXTURED | FLAG_Y_PEAKED)
    mx = float(np.mean(b["sub"][..., 0][fx])) if fx.any() else 0.0
    my = float(np.mean(b["sub"][..., 1][fy])) if fy.any() else 0.0
    return mx, my


def adaptive_histogram_equalization(noisy_arrs, *, clip_limit: float = 2.0, tiles=(8, 8), device: Optional[int] = None,
                                    chunk: int = 256):
    """(N,H,W) uint8 -> CLAHE-equalised (N,H,W) uint8 -- estimate_vel_vid.py:201-219, on the GPU.

    Same result as ``cv.createCLAHE(clipLimit=2.0, tileGridSize=(8,8)).apply`` per frame (bit-exact), written
    into one preallocated array instead of the reference's per-frame ``np.append`` (:215)."""
    frames = np.ascontiguousarray(noisy_arrs, dtype=np.uint8)
    if frames.ndim != 3:
        raise ValueError("noisy_arrs must be (N, H, W) uint8")
    n = frames.shape[0]
    chunk = max(1, min(chunk, n))
    eng = _engine(frames.shape[1], frames.shape[2], default_config, default_device if device is None else device,
                  max(chunk, 2), 1)
    out = np.empty_like(frames)
    for k in range(0, n, chunk):
        m = min(chunk, n - k)
        eng.upload(frames[k:k + m], first_slot=0)
        eng.clahe(0, m, clip_limit, tiles)
        out[k:k + m] = eng.download(0, m)
    return out


def reference_windows(n_frames, n_list=None):
    """Temporal window denoise_hf5 picks for frame k (estimate_vel_vid.py:250-262): the single-image call for the two
    first and three last frames, 5 frames for k <= 4 or k >= len(numberList) - 5, otherwise 9."""
    L = n_frames if n_list is None else int(n_list)
    return [1 if (k <= 1 or k >= L - 3) else 5 if (k <= 4 or k >= L - 5) else 9 for k in range(n_frames)]


def _denoise_frames(frames, windows, h, template_window, search_window, device, chunk):
    """Denoise frames[k] with temporal window windows[k]; device chunks carry the halo frames they need."""
    n = frames.shape[0]
    halo = max(w // 2 for w in windows)
    chunk = max(1, min(chunk, n))
    eng = _engine(frames.shape[1], frames.shape[2], default_config, default_device if device is None else device,
                  max(chunk + 2 * halo, 2), 1)
    out = np.empty_like(frames)
    for k in range(0, n, chunk):
        m = min(chunk, n - k)
        lo, hi = max(0, k - halo), min(n, k + m + halo)
        eng.upload(frames[lo:hi], first_slot=0)
        tw = [windows[i] if k <= i < k + m else 0 for i in range(lo, hi)]       # halo frames are only read (window 0)
        eng.denoise(0, tw, h, template_window, search_window)
        out[k:k + m] = eng.download(k - lo, m)
    return out


def denoising(arrays, numberIndex, num_frames_window, *, h: float = 15.0, template_window: int = 19, search_window: int = 41,
              device: Optional[int] = None):
    """One frame of ``arrays`` denoised from its temporal neighbours -- estimate_vel_vid.py:226-232
    (cv.fastNlMeansDenoisingMulti(h=15, templateWindowSize=19, searchWindowSize=41)), on the GPU, bit-exact.
    Raises ValueError where cv2 raises (window leaving the sequence, even window)."""
    frames = np.ascontiguousarray(arrays, dtype=np.uint8)
    if frames.ndim != 3:
        raise ValueError("arrays must be (N, H, W) uint8")
    half = int(num_frames_window) // 2
    if num_frames_window % 2 == 0 or numberIndex - half < 0 or numberIndex + half >= len(frames):
        raise ValueError("temporal window must be odd and lie inside the sequence")
    sub = frames[numberIndex - half:numberIndex + half + 1]
    tw = [0] * len(sub)
    tw[half] = int(num_frames_window)
    return _denoise_frames(sub, tw, h, template_window, search_window, device, len(sub))[half]


def denoise_hf5(eq_arrs, *, n_list: Optional[int] = None, h: float = 15.0, template_window: int = 19, search_window: int = 41,
                device: Optional[int] = None, chunk: int = 256, h5_path: Optional[str] = None):
    """(N,H,W) uint8 -> denoised (N,H,W) uint8 -- the role of estimate_vel_vid.py:246-299 on the GPU.

    Frame k gets the reference's choice of call (:250-262, ``n_list`` = len(numberList), default N): the single-image
    filter at the ends, the 5-frame and then the 9-frame temporal window inside.  Where the reference's call would
    raise (window leaving the array because the video was shorter than numberList) it prints and stops (:263-266);
    the same truncation happens here.  ``h5_path`` writes the ``clean_images`` dataset like :287-297 when h5py is
    importable (it is not in this image)."""
    frames = np.ascontiguousarray(eq_arrs, dtype=np.uint8)
    if frames.ndim != 3:
        raise ValueError("eq_arrs must be (N, H, W) uint8")
    n = frames.shape[0]
    windows = reference_windows(n, n_list)
    for k, w in enumerate(windows):                       # first frame whose call would throw ends the sequence
        if k - w // 2 < 0 or k + w // 2 >= n:
            print("something went wrong with denoising")
            n = k
            break
    if n == 0:
        return frames[:0].copy()
    out = _denoise_frames(frames, windows[:n] + [0] * (frames.shape[0] - n), h, template_window, search_window, device, chunk)[:n]
    if h5_path is not None:
        try:
            import h5py
        except ImportError:
            print(f"h5py is not installed: {h5_path} not written")
        else:
            with h5py.File(h5_path, "w") as f:
                ds = f.create_dataset("clean_images", data=out, maxshape=(len(out) + 10,) + out.shape[1:],
                                      chunks=(min(10, len(out)),) + out.shape[1:], dtype="uint8", compression="gzip",
                                      compression_opts=7, shuffle=True)
                ds.attrs["CLASS"] = "IMAGE"
                ds.attrs["IMAGE_VERSION"] = "1.2"
                ds.attrs["IMAGE_SUBCLASS"] = "IMAGE_GRAYSCALE"
                ds.attrs["IMAGE_MINMAXRANGE"] = np.array([0, 255], dtype=np.uint8)
                ds.attrs["IMAGE_WHITE_IS_ZERO"] = 0
    return out


class _RingSource:
    """d(anchor, k) for a host frame stack through a bounded device ring.

    Slot 0 always holds the current anchor; slots 1.. hold a sliding window of frames.  Nothing is
    assumed about how far k runs ahead of the anchor (a static scene never accepts a new anchor).
    """

    def __init__(self, frames, eng: MotionEstimator, scale):
        self.frames, self.eng, self.scale = frames, eng, np.float32(scale)
        self.cap = eng.max_frames - 1
        self.w0 = self.w1 = 0          # frames [w0, w1) are resident in slots 1..
        self.anchor = -1

    def _ensure_window(self, k_lo, k_hi):
        if k_lo >= self.w0 and k_hi <= self.w1:
            return
        self.w0 = k_lo
        self.w1 = min(k_lo + self.cap, len(self.frames))
        self.eng.upload(self.frames[self.w0:self.w1], first_slot=1)

    def __call__(self, anchor, ks):
        if anchor != self.anchor:
            self.eng.upload(self.frames[anchor:anchor + 1], first_slot=0)
            self.anchor = anchor
        out = []
        i = 0
        while i < len(ks):
            self._ensure_window(ks[i], ks[i] + 1)
            n = min(len(ks) - i, self.w1 - ks[i], self.eng.max_pairs)
            pairs = [(0, 1 + (k - self.w0)) for k in ks[i:i + n]]
            res = self.eng.search(pairs)
            out.extend((np.float32(r["pdx"] * self.scale), np.float32(r["pdy"] * self.scale), int(r["flag"])) for r in res)
            i += n
        return out


def _find_one(directory, pattern):
    hits = sorted(glob.glob(os.path.join(directory, pattern)))
    return hits[0] if hits else None


def end_process(clean_arrs, *, imagePath: Optional[str] = None, firstFrame: int = 0,
                numberList: Optional[Sequence[int]] = None, timestamps_ms: Optional[Sequence[float]] = None,
                config: Optional[SearchConfig] = None, device: Optional[int] = None, units: str = "reference",
                ring_frames: int = 512, max_pairs: int = 256, return_details: bool = False):
    """The video script's pair loop and velocity reduction -- estimate_vel_vid.py:493-583.

    Returns ``(out_information, velocity_list_x, velocity_list_y, tsList, motion_list_opt)`` exactly
    as the reference does (``return_details=True`` returns the full PairLoopResult instead).  The
    reference reads ``numberList``, ``imagePath`` and ``firstFrame`` from module globals; here they
    are keyword arguments.  Timestamps come from the first ``*tstamps.txt`` in ``imagePath`` (:516)
    unless ``timestamps_ms`` (one value per frame of the recording, file order) is given.
    """
    frames = np.ascontiguousarray(clean_arrs, dtype=np.uint8)
    if frames.ndim != 3:
        raise ValueError("clean_arrs must be (N, H, W) uint8")
    n = frames.shape[0] if numberList is None else min(frames.shape[0], len(numberList))   # zip() at :506
    if timestamps_ms is not None:
        lines = ["# timecode format v2"] + [repr(float(t)) for t in timestamps_ms]
    else:
        if imagePath is None:
            raise ValueError("end_process needs imagePath (directory with *tstamps.txt) or timestamps_ms")
        ts_file = _find_one(imagePath, "*tstamps.txt")
        if ts_file is None:
            raise FileNotFoundError(f"no *tstamps.txt in {imagePath} (estimate_vel_vid.py:516)")
        lines = _vel.read_timecodes(ts_file)
    cfg = config or default_config
    dev = default_device if device is None else device
    ring = max(2, min(ring_frames, n + 1))
    eng = _engine(frames.shape[1], frames.shape[2], cfg, dev, ring, max_pairs)
    src = _RingSource(frames[:n], eng, _scale(units))
    look = _vel.AnchoredPairs(src, n, max_lookahead=min(max_pairs, eng.max_frames - 1))
    res = _vel.anchored_pair_loop(n, look.get, _vel.timecode_lookup(lines), first_frame=firstFrame,
                                  pixels_per_mm=pixels_per_mm)
    res.searches, res.launches = look.searches, look.launches
    return res if return_details else res.as_tuple()


def end_process_raw(clean_slice, timestamps_us, *, config: Optional[SearchConfig] = None, device: Optional[int] = None,
                    max_pairs: int = 256, return_details: bool = False):
    """The raw-image script's loop -- estimate_vel.py:333-378: strictly consecutive pairs, so every
    pair of the recording is searched in a handful of batched launches.

    ``timestamps_us[k]`` is the microsecond value the reference reads for frame k from tstamps.csv.
    Returns ``(out_information, velocity_list_x, velocity_list_y, tsList)``.
    """
    frames = np.ascontiguousarray(clean_slice, dtype=np.uint8)
    n = frames.shape[0]
    cfg = config or default_config
    dev = default_device if device is None else device
    chunk = min(max_pairs, max(n - 1, 1))
    eng = _engine(frames.shape[1], frames.shape[2], cfg, dev, chunk + 1, chunk)
    disp = []
    for k0 in range(0, n - 1, chunk):
        k1 = min(k0 + chunk, n - 1)
        eng.upload(frames[k0:k1 + 1], first_slot=0)
        res = eng.search([(i, i + 1) for i in range(k1 - k0)])
        disp.extend((np.float32(r["pdx"]), np.float32(r["pdy"])) for r in res)
    out = _vel.consecutive_pair_loop(disp, timestamps_us, pixels_per_mm=pixels_per_mm)
    if return_details:
        return out
    return out.out_information, out.velocity_list_x, out.velocity_list_y, out.tsList
