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
from .engine import PAIR_DTYPE, MotionEstimator, SearchConfig

# Module-level state mirrors the reference's globals (estimate_vel_vid.py:73, :396-412).
pixels_per_mm = _vel.PIXELS_PER_MM
max_filament_speed = _vel.MAX_FILAMENT_SPEED_MM_MIN
# +-64 full-resolution pixels = +-32 of the reference's half-resolution pixels: the fastest example recording
# (116.4 mm/min) jumps up to 57 px between accepted frames, which a +-32 search cannot see.
# min_range = 6 grey levels (the library's own default is 2): on the example recordings after the reference's pre-processing
# most blocks are nearly flat, and their shallow minima pull the median towards the static pattern of the sensor; with the
# higher texture gate the per-pair agreement with the reference's ECC goes from 96.5 / 83.8 / 75.5 % to 99.3 / 92.9 / 80.5 %
# of pairs within 0.3 px and the travel deficit against its series from 7 / 18 / 27 % to 5 / 13 / 9 %, with no failed pair
# (higher gates leave some pairs without any valid block).
default_config = SearchConfig(block=16, range=64, min_range=6)
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
    s = np.float32(_scale(units))                 # bad arguments are refused before any device work
    cfg = config or default_config
    eng = _engine(prev.shape[0], prev.shape[1], cfg, default_device if device is None else device, 2, 1)
    r = eng.estimate_pair(prev, cur)
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
    r = eng.estimate_pair(prev, cur)          # the mean is part of the pair record (reduced on the GPU with the median)
    return float(r["mean_dx"]), float(r["mean_dy"])


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
        eng.download(0, m, out=out[k:k + m])
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
        eng.download(k - lo, m, out=out[k:k + m])
    return out


def _denoise_frames_sharded(frames, windows, h, template_window, search_window, device, chunk):
    """Frames [0, len(windows)) denoised across the ranks: contiguous ranges, halo frames read only, one gather."""
    import torch
    import torch.distributed as dist
    from . import shard
    n = len(windows)
    world, rank = dist.get_world_size(), dist.get_rank()
    lo, hi = shard.shard_range(n, rank, world)
    halo = max(w // 2 for w in windows)
    per = -(-n // world)
    dev = torch.cuda.current_device() if device is None else device
    eng = _engine(frames.shape[1], frames.shape[2], default_config, dev, max(per + 2 * halo, 2), 1)
    if hi > lo:
        f_lo, f_hi = max(0, lo - halo), min(frames.shape[0], hi + halo)
        eng.upload(frames[f_lo:f_hi], first_slot=0)
        tw = [windows[i] if lo <= i < hi else 0 for i in range(f_lo, f_hi)]
        eng.denoise(0, tw, h, template_window, search_window)
        first = lo - f_lo
    else:
        first = 0
    return shard.gather_frames_sharded(None, eng, first, hi - lo, n, frames.shape[1:])


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
                device: Optional[int] = None, chunk: int = 256, h5_path: Optional[str] = None,
                distributed: Optional[bool] = None):
    """(N,H,W) uint8 -> denoised (N,H,W) uint8 -- the role of estimate_vel_vid.py:246-299 on the GPU.

    Frame k gets the reference's choice of call (:250-262, ``n_list`` = len(numberList), default N): the single-image
    filter at the ends, the 5-frame and then the 9-frame temporal window inside.  Where the reference's call would
    raise (window leaving the array because the video was shorter than numberList) it prints and stops (:263-266);
    the same truncation happens here.  ``h5_path`` writes the ``clean_images`` dataset like :287-297 when h5py is
    importable (it is not in this image).

    ``distributed`` (default: whenever torch.distributed runs with more than one rank; call on every rank with the same
    frames): each rank denoises one contiguous frame range, uploading up to 4 halo frames on either side (the 9-frame
    window), and the cleaned ranges are all-gathered from device memory; every rank returns the whole array."""
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
    from . import shard
    if distributed is None:
        distributed = shard.is_distributed()
    if distributed and n >= 2:
        out = _denoise_frames_sharded(frames, windows[:n], h, template_window, search_window, device, chunk)
    else:
        out = _denoise_frames(frames, windows[:n] + [0] * (frames.shape[0] - n), h, template_window, search_window, device, chunk)[:n]
    if h5_path is not None and not write_clean_images_h5(out, h5_path):
        print(f"h5py is not installed: {h5_path} not written")
    return out


def write_clean_images_h5(frames, h5_path) -> bool:
    """The reference's cache file (estimate_vel_vid.py:287-297): dataset ``clean_images``, gzip level 7, chunks of 10
    frames, its five image attributes.  Returns False (nothing written) when h5py is not importable."""
    try:
        import h5py
    except ImportError:
        return False
    out = np.ascontiguousarray(frames, dtype=np.uint8)
    with h5py.File(h5_path, "w") as f:
        ds = f.create_dataset("clean_images", data=out, maxshape=(len(out) + 10,) + out.shape[1:],
                              chunks=(min(10, len(out)),) + out.shape[1:], dtype="uint8", compression="gzip",
                              compression_opts=7, shuffle=True)
        ds.attrs["CLASS"] = "IMAGE"
        ds.attrs["IMAGE_VERSION"] = "1.2"
        ds.attrs["IMAGE_SUBCLASS"] = "IMAGE_GRAYSCALE"
        ds.attrs["IMAGE_MINMAXRANGE"] = np.array([0, 255], dtype=np.uint8)
        ds.attrs["IMAGE_WHITE_IS_ZERO"] = 0
    return True


def _anchored_table(frames, eng: MotionEstimator, scale):
    """d(anchor_k, k) for every frame k >= 1 of a host frame stack: the anchor walk runs inside the library
    (me_anchored_scan), one call per window of frames that fits the device ring.

    Slot 0 is a spare for an anchor older than the window (a static scene never accepts a new anchor); slots 1.. hold
    the window.  Returns (anchor index per frame, list of (pdx, pdy, flag, dense mean x) per frame, searches, launches).
    """
    n = len(frames)
    cap = eng.max_frames - 1
    anchors = np.zeros(n, np.int64)
    pdx, pdy = np.zeros(n, np.float32), np.zeros(n, np.float32)
    flag, dense = np.zeros(n, np.int64), np.zeros(n, np.float64)
    anchor, k, searches, launches = 0, 1, 0, 0
    s = np.float32(scale)
    while k < n:
        w0 = min(anchor, k) if k - anchor < cap else k           # keep the anchor inside the window when it is recent enough
        w1 = min(w0 + cap, n)
        eng.upload(frames[w0:w1], first_slot=1)
        if anchor < w0:
            eng.upload(frames[anchor:anchor + 1], first_slot=0)
        a_slot = 1 + anchor - w0 if anchor >= w0 else 0
        rec, ns, nl = eng.anchored_scan(a_slot, 1 + k - w0, w1 - k, scale=float(s))
        searches += ns
        launches += nl
        slots = rec["anchor_slot"].astype(np.int64)
        anchors[k:w1] = np.where(slots == 0, anchor, w0 + slots - 1)     # slot 0 only ever holds the anchor older than the window
        pdx[k:w1] = rec["pair"]["pdx"] * s
        pdy[k:w1] = rec["pair"]["pdy"] * s
        flag[k:w1] = rec["pair"]["flag"]
        dense[k:w1] = rec["pair"]["mean_dx"]     # full-resolution pixels, the reference's Farneback mean (:524)
        acc = np.nonzero(rec["accepted"])[0]
        if len(acc):
            anchor = k + int(acc[-1])
        k = w1
    table = list(zip(pdx, pdy, flag.tolist(), dense.tolist()))
    return anchors, table, searches, launches


def _find_one(directory, pattern):
    hits = sorted(glob.glob(os.path.join(directory, pattern)))
    return hits[0] if hits else None


def end_process(clean_arrs, *, imagePath: Optional[str] = None, firstFrame: int = 0,
                numberList: Optional[Sequence[int]] = None, timestamps_ms: Optional[Sequence[float]] = None,
                config: Optional[SearchConfig] = None, device: Optional[int] = None, units: str = "reference",
                ring_frames: Optional[int] = None, max_pairs: int = 256, return_details: bool = False):
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
    if ring_frames is None:                           # whole recordings of the 640x64 shape stay resident; large frames in 256 MB windows
        ring_frames = max(512, (256 << 20) // (frames.shape[1] * frames.shape[2]))
    ring = max(3, min(ring_frames, n + 1))
    eng = _engine(frames.shape[1], frames.shape[2], cfg, dev, ring, max_pairs)
    # the sequential anchor walk (:531-534, :575) runs inside the library; the arithmetic on top of it stays the host mirror
    anchors, table, searches, launches = _anchored_table(frames[:n], eng, _scale(units))

    def pair(a, k):
        if anchors[k] != a:
            raise RuntimeError(f"anchor walk out of step at frame {k}: library compared with {anchors[k]}, host loop expects {a}")
        return table[k]

    res = _vel.anchored_pair_loop(n, pair, _vel.timecode_lookup(lines), first_frame=firstFrame, pixels_per_mm=pixels_per_mm)
    res.searches, res.launches = searches, launches
    return res if return_details else res.as_tuple()


def end_process_raw(clean_slice, timestamps_us, *, config: Optional[SearchConfig] = None, device: Optional[int] = None,
                    max_pairs: int = 256, return_details: bool = False, distributed: Optional[bool] = None):
    """The raw-image script's loop -- estimate_vel.py:333-378: strictly consecutive pairs, so every
    pair of the recording is searched in a handful of batched launches.

    ``timestamps_us[k]`` is the microsecond value the reference reads for frame k from tstamps.csv.
    Returns ``(out_information, velocity_list_x, velocity_list_y, tsList)``.

    ``distributed`` (default: whenever torch.distributed is initialised with more than one rank; call it on every
    rank with the same frames): the pair range is cut into one contiguous slice per rank, each rank uploads its frames
    plus the one-frame halo and searches them on its GPU (``device`` defaults to the rank's current CUDA device), and the
    32-byte pair records are all-gathered from device memory on a side stream (shard.RecordGather).  Every rank returns
    the whole series, byte-identical to the single-GPU result.
    """
    from . import shard
    frames = np.ascontiguousarray(clean_slice, dtype=np.uint8)
    n = frames.shape[0]
    cfg = config or default_config
    if distributed is None:
        distributed = shard.is_distributed()
    if distributed and n > 2:
        import torch
        dev = torch.cuda.current_device() if device is None else device
        rec = shard.search_consecutive_sharded(
            frames, lambda mf, mp: _engine(frames.shape[1], frames.shape[2], cfg, dev, mf, mp), PAIR_DTYPE, chunk=max_pairs)
        disp = list(zip(rec["pdx"], rec["pdy"]))
    else:
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


def end_process_recordings(recordings: Sequence[dict], **kw):
    """The video script's pair loop for SEVERAL recordings (BASELINE configs[4]: concurrent streams).  The anchored loop
    is sequential inside a recording (estimate_vel_vid.py:531-534, :575), so recordings -- not frame ranges -- are what
    the ranks share out: rank r runs ``end_process(**recordings[i], **kw)`` for its contiguous share of the list and the
    results (a few KB each) are exchanged with all_gather_object.  Without torch.distributed it is a plain loop.
    Every rank returns the list of PairLoopResult in input order."""
    from . import shard
    kw = dict(kw, return_details=True)
    if not shard.is_distributed():
        return [end_process(**r, **kw) for r in recordings]
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    lo, hi = shard.shard_range(len(recordings), rank, world)
    kw.setdefault("device", torch.cuda.current_device())
    mine = [end_process(**r, **kw) for r in recordings[lo:hi]]
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    return [r for p in parts for r in p]
