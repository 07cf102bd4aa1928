"""MotionEstimator: one GPU context of libme_b200 with NumPy in / NumPy out.

This is the host-side object behind the drop-in functions in ``api.py``; it mirrors the C ABI
one to one (include/me_b200.h) and adds nothing numerical of its own.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from ._lib import Geometry, PairResult, Params, check

FLAG_TEXTURED, FLAG_X_INTERIOR, FLAG_Y_INTERIOR, FLAG_X_PEAKED, FLAG_Y_PEAKED = 1, 2, 4, 8, 16

PAIR_DTYPE = np.dtype([("pdx", "<f4"), ("pdy", "<f4"), ("flag", "<i4"), ("nvalid_x", "<i4"),
                       ("nvalid_y", "<i4"), ("n_textured", "<i4"), ("mean_dx", "<f4"), ("mean_dy", "<f4")])


SCAN_DTYPE = np.dtype([("anchor_slot", "<i4"), ("accepted", "<i4"), ("pair", PAIR_DTYPE)])


@dataclass(frozen=True)
class SearchConfig:
    """Knobs of the block search (the reference's ECC call has none of them)."""
    block: int = 16
    range: int = 32
    min_range: int = 2
    subpixel: bool = True
    min_valid: int = 1
    min_curv: int = 1
    range_y: int = 0          # vertical range, 0 = the same as range; the recordings move along x only

    def as_struct(self) -> Params:
        return Params(self.block, self.range, self.min_range, int(self.subpixel), self.min_valid, self.min_curv, self.range_y)


class MotionEstimator:
    """Frame ring + search on one CUDA device.

    >>> me = MotionEstimator(1080, 1920, SearchConfig(16, 32), max_frames=33, max_pairs=32)
    >>> me.upload(frames)                       # (N,H,W) uint8 -> device slots 0..N-1
    >>> res = me.search([(k, k + 1) for k in range(32)])   # structured array, one record per pair
    """

    def __init__(self, height, width, config: SearchConfig = SearchConfig(), max_frames=64, max_pairs=64, device=0):
        self._L = _lib.lib()
        self._ctx = C.c_void_p()
        self.config = config
        prm = config.as_struct()
        check(self._L.me_create(int(device), int(height), int(width), C.byref(prm), int(max_frames), int(max_pairs),
                                C.byref(self._ctx)))
        g = Geometry()
        check(self._L.me_get_geometry(self._ctx, C.byref(g)))
        self.geometry = {name: getattr(g, name) for name, _ in Geometry._fields_}
        self.height, self.width = int(height), int(width)
        self.nbx, self.nby = g.nbx, g.nby
        self.max_frames, self.max_pairs = int(max_frames), int(max_pairs)
        self.device = int(device)

    # -- lifetime
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._L.me_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- ingest
    def _frames_arg(self, frames, channels):
        a = np.asarray(frames)
        if a.dtype != np.uint8:
            raise TypeError("frames must be uint8")
        want = 3 + (channels == 3)
        if a.ndim == want - 1:
            a = a[None]
        if a.ndim != want or a.shape[1] != self.height or a.shape[2] != self.width or (channels == 3 and a.shape[3] != 3):
            raise ValueError(f"frames of shape {a.shape} do not match (N,{self.height},{self.width}"
                             + (",3)" if channels == 3 else ")"))
        return np.ascontiguousarray(a)

    def upload(self, frames, first_slot=0):
        """(N,H,W) uint8 luma -> ring slots first_slot..first_slot+N-1 (async on the copy stream)."""
        a = self._frames_arg(frames, 1)
        check(self._L.me_upload(self._ctx, a.ctypes.data, a.strides[1], int(first_slot), a.shape[0]))
        return a.shape[0]

    def upload_bgr(self, frames, first_slot=0):
        """(N,H,W,3) uint8 BGR -> luma ring slots; conversion on the GPU (cv.cvtColor BGR2GRAY arithmetic)."""
        a = self._frames_arg(frames, 3)
        check(self._L.me_upload_bgr(self._ctx, a.ctypes.data, a.strides[1], int(first_slot), a.shape[0]))
        return a.shape[0]

    def clahe(self, first_slot=0, n_frames=1, clip_limit=2.0, tiles=(8, 8)):
        """CLAHE of ring slots in place on the GPU (cv.createCLAHE(clip_limit, tiles).apply, bit-exact)."""
        check(self._L.me_clahe(self._ctx, int(first_slot), int(n_frames), float(clip_limit), int(tiles[0]), int(tiles[1])))

    def denoise(self, first_slot, temporal_windows, h=15.0, template_window=19, search_window=41):
        """Non-local-means denoising of len(temporal_windows) ring slots in place on the GPU
        (cv.fastNlMeansDenoising / ...Multi with temporalWindowSize = temporal_windows[k], bit-exact)."""
        tw = np.ascontiguousarray(temporal_windows, np.int32)
        check(self._L.me_denoise(self._ctx, int(first_slot), int(tw.size), tw.ctypes.data_as(C.POINTER(C.c_int32)),
                                 float(h), int(template_window), int(search_window)))

    def download(self, first_slot=0, n_frames=1, out=None):
        """Ring slots -> (n_frames,H,W) uint8 host array (into `out` when given: a C-contiguous-row view of that shape)."""
        if out is None:
            out = np.empty((int(n_frames), self.height, self.width), np.uint8)
        elif (out.dtype != np.uint8 or out.shape != (int(n_frames), self.height, self.width) or out.strides[2] != 1
              or out.strides[0] != self.height * out.strides[1]):
            raise ValueError("out must be a (n_frames, H, W) uint8 array with contiguous pixels and evenly spaced rows")
        check(self._L.me_download(self._ctx, out.ctypes.data, out.strides[1], int(first_slot), int(n_frames)))
        return out

    # -- search
    @staticmethod
    def _pairs_arg(pairs):
        p = np.ascontiguousarray(np.asarray(pairs, dtype=np.int32).reshape(-1, 2))
        return p, p.ctypes.data_as(C.POINTER(C.c_int32))

    def launch(self, pairs):
        p, ptr = self._pairs_arg(pairs)
        check(self._L.me_search(self._ctx, ptr, p.shape[0]))
        return p.shape[0]

    def last_ticket(self) -> int:
        """Names the search just launched; valid until three more searches have been launched."""
        return int(self._L.me_last_ticket(self._ctx))

    def fetch_pairs(self, n, ticket=None):
        """Records of the last search, or of the search `ticket` names (several searches may be in flight)."""
        out = np.zeros(n, PAIR_DTYPE)
        if ticket is None:
            check(self._L.me_fetch_pairs(self._ctx, out.ctypes.data_as(C.POINTER(PairResult)), int(n)))
        else:
            check(self._L.me_fetch_pairs_ticket(self._ctx, int(ticket), out.ctypes.data_as(C.POINTER(PairResult)), int(n)))
        return out

    def search(self, pairs):
        """Launch + fetch: structured array (pdx, pdy, flag, nvalid_x, nvalid_y) per pair."""
        return self.fetch_pairs(self.launch(pairs))

    def search_device(self, d_ptr, pitch, n_frames, pairs, stream=None):
        """Search a device-resident (n_frames,H,pitch) uint8 stack (e.g. ``tensor.data_ptr()``)."""
        p, ptr = self._pairs_arg(pairs)
        check(self._L.me_search_device(self._ctx, C.c_void_p(int(d_ptr)), int(pitch), int(n_frames), ptr, p.shape[0],
                                       C.c_void_p(int(stream) if stream else None)))
        return p.shape[0]

    def anchored_scan(self, anchor_slot, first_slot, n_frames, scale=0.5, gate=2.0):
        """The video script's anchor walk (estimate_vel_vid.py:506-534, :575) over ring slots first_slot.. inside the library:
        (records, searches, launches); records[i] describes slot first_slot + i (anchor slot, accepted, pair record)."""
        out = np.zeros(int(n_frames), SCAN_DTYPE)
        ns, nl = C.c_int(), C.c_int()
        check(self._L.me_anchored_scan(self._ctx, int(anchor_slot), int(first_slot), int(n_frames), float(scale), float(gate),
                                       out.ctypes.data, C.byref(ns), C.byref(nl)))
        return out, ns.value, nl.value

    def copy_pair_records_device(self, d_dst, n_pairs, stream=None):
        """Records of the last search -> device memory at d_dst (e.g. a torch tensor's data_ptr()) on `stream`, no host bounce."""
        check(self._L.me_copy_pair_records_device(self._ctx, C.c_void_p(int(d_dst)), int(n_pairs),
                                                  C.c_void_p(int(stream) if stream else None)))

    def copy_frames_device(self, d_dst, dst_pitch, first_slot, n_frames, stream=None):
        """Ring slots -> device memory at d_dst (rows dst_pitch bytes apart), no host bounce."""
        check(self._L.me_copy_frames_device(self._ctx, C.c_void_p(int(d_dst)), int(dst_pitch), int(first_slot), int(n_frames),
                                            C.c_void_p(int(stream) if stream else None)))

    def copy_timer_start(self):
        check(self._L.me_copy_timer_start(self._ctx))

    def copy_timer_stop(self) -> float:
        ms = C.c_float()
        check(self._L.me_copy_timer_stop(self._ctx, C.byref(ms)))
        return float(ms.value)

    def fetch_blocks(self, pair_index=0):
        nb = (self.nby, self.nbx)
        mv = np.zeros(nb + (2,), np.int16)
        sad = np.zeros(nb, np.uint32)
        sub = np.zeros(nb + (2,), np.float32)
        flags = np.zeros(nb, np.uint8)
        check(self._L.me_fetch_blocks(self._ctx, int(pair_index), mv.ctypes.data, sad.ctypes.data, sub.ctypes.data,
                                      flags.ctypes.data))
        return dict(mv=mv, sad=sad, sub=sub, flags=flags)

    def estimate_pair(self, prev, cur):
        """One host frame pair -> PairResult fields as a dict (what calc_ECC_transform binds)."""
        prev = np.ascontiguousarray(prev, dtype=np.uint8)
        cur = np.ascontiguousarray(cur, dtype=np.uint8)
        if prev.shape != (self.height, self.width) or cur.shape != prev.shape:
            raise ValueError(f"frames must both be ({self.height},{self.width}) uint8")
        r = PairResult()
        check(self._L.me_estimate_pair(self._ctx, prev.ctypes.data, cur.ctypes.data, prev.strides[0], C.byref(r)))
        return dict(pdx=np.float32(r.pdx), pdy=np.float32(r.pdy), flag=int(r.flag), nvalid_x=int(r.nvalid_x),
                    nvalid_y=int(r.nvalid_y), n_textured=int(r.n_textured), mean_dx=np.float32(r.mean_dx), mean_dy=np.float32(r.mean_dy))

    # -- misc
    def sync(self):
        check(self._L.me_sync(self._ctx))

    def last_search_ms(self) -> float:
        ms = C.c_float()
        check(self._L.me_last_search_ms(self._ctx, C.byref(ms)))
        return float(ms.value)

    def timer_start(self):
        check(self._L.me_timer_start(self._ctx))

    def timer_stop(self) -> float:
        ms = C.c_float()
        check(self._L.me_timer_stop(self._ctx, C.byref(ms)))
        return float(ms.value)

    def launch_count(self) -> int:
        return int(self._L.me_launch_count(self._ctx))


def microbench_vabsdiff4(device=0, mode=0):
    """(VABSDIFF4 thread-instr / clk / SM, SM MHz seen, SM count) measured on `device`."""
    L = _lib.lib()
    ipc, mhz, sms = C.c_double(), C.c_double(), C.c_int()
    check(L.me_microbench_vabsdiff4(int(device), int(mode), C.byref(ipc), C.byref(mhz), C.byref(sms)))
    return float(ipc.value), float(mhz.value), int(sms.value)
