"""ctypes binding of oracle/libsad_oracle.so (test infrastructure; see oracle/__init__.py)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libsad_oracle.so")

FLAG_TEXTURED = 1
FLAG_X_INTERIOR = 2
FLAG_Y_INTERIOR = 4
FLAG_X_PEAKED = 8
FLAG_Y_PEAKED = 16


class Params(C.Structure):
    _fields_ = [("block", C.c_int32), ("range", C.c_int32), ("min_range", C.c_int32),
                ("subpixel", C.c_int32), ("min_valid", C.c_int32), ("min_curv", C.c_int32), ("range_y", C.c_int32)]


def build(force: bool = False) -> str:
    """Compile the C oracle with the committed Makefile (gcc, a second or two)."""
    src = os.path.join(_HERE, "sad_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-s", "-B", "libsad_oracle.so"], check=True)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, i16p, u32p, f32p, i32p = (C.POINTER(t) for t in (C.c_uint8, C.c_int16, C.c_uint32, C.c_float, C.c_int32))
        L.me_oracle_search.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_size_t, C.POINTER(Params),
                                       i16p, u32p, f32p, u8p, f32p, i32p, C.c_int]
        L.me_oracle_search.restype = C.c_int
        L.me_oracle_search_pairs.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_size_t, C.POINTER(Params),
                                             i32p, C.c_int, f32p, i32p, C.c_int]
        L.me_oracle_search_pairs.restype = C.c_int
        L.me_oracle_cost_surface.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int,
                                             C.c_int, C.c_int, u32p]
        L.me_oracle_cost_surface.restype = C.c_int
        L.me_oracle_bgr2gray.argtypes = [u8p, C.c_size_t, u8p]
        L.me_oracle_bgr2gray.restype = C.c_int
        L.me_oracle_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _frame(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    if a.ndim != 2:
        raise ValueError("frame must be (H, W) uint8")
    return a


def search(prev, cur, block=16, rng=32, min_range=2, subpixel=True, min_valid=1, min_curv=1, nthreads=0, range_y=0):
    """One frame pair through the C oracle.  Returns a dict of per-block and per-pair results."""
    prev, cur = _frame(prev), _frame(cur)
    if prev.shape != cur.shape:
        raise ValueError("frame shapes differ")
    H, W = prev.shape
    nbx, nby = W // block, H // block
    nb = nbx * nby
    mv = np.zeros((nby, nbx, 2), np.int16)
    sad = np.zeros((nby, nbx), np.uint32)
    sub = np.zeros((nby, nbx, 2), np.float32)
    flags = np.zeros((nby, nbx), np.uint8)
    pd = np.zeros(4, np.float32)
    pi = np.zeros(4, np.int32)
    prm = Params(block, rng, min_range, int(bool(subpixel)), min_valid, min_curv, range_y)
    rc = lib().me_oracle_search(_p(prev, C.c_uint8), _p(cur, C.c_uint8), H, W, W, C.byref(prm),
                                _p(mv, C.c_int16), _p(sad, C.c_uint32), _p(sub, C.c_float),
                                _p(flags, C.c_uint8), _p(pd, C.c_float), _p(pi, C.c_int32), nthreads)
    if rc:
        raise ValueError(f"me_oracle_search rc={rc} (H={H} W={W} B={block} R={rng})")
    assert nb == mv.shape[0] * mv.shape[1]
    return dict(mv=mv, sad=sad, sub=sub, flags=flags, pdx=pd[0], pdy=pd[1], flag=int(pi[0]),
                nvalid_x=int(pi[1]), nvalid_y=int(pi[2]), n_textured=int(pi[3]), mean_dx=pd[2], mean_dy=pd[3])


def search_pairs(frames, pairs, block=16, rng=32, min_range=2, subpixel=True, min_valid=1, min_curv=1, nthreads=0, range_y=0):
    """(N,H,W) frames and a (P,2) list of (prev,cur) indices -> (P,4) float32 (pdx,pdy,mean_dx,mean_dy), (P,4) int32 (flag,nx,ny,n_textured)."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
    N, H, W = frames.shape
    P = pairs.shape[0]
    pd = np.zeros((P, 4), np.float32)
    pi = np.zeros((P, 4), np.int32)
    prm = Params(block, rng, min_range, int(bool(subpixel)), min_valid, min_curv, range_y)
    rc = lib().me_oracle_search_pairs(_p(frames, C.c_uint8), N, H, W, W, C.byref(prm), _p(pairs, C.c_int32), P,
                                      _p(pd, C.c_float), _p(pi, C.c_int32), nthreads)
    if rc:
        raise ValueError(f"me_oracle_search_pairs rc={rc}")
    return pd, pi


def cost_surface(prev, cur, block, rng, bx, by):
    prev, cur = _frame(prev), _frame(cur)
    H, W = prev.shape
    U = 2 * rng + 1
    out = np.zeros((U, U), np.uint32)
    rc = lib().me_oracle_cost_surface(_p(prev, C.c_uint8), _p(cur, C.c_uint8), H, W, W, block, rng, bx, by,
                                      _p(out, C.c_uint32))
    if rc:
        raise ValueError(f"me_oracle_cost_surface rc={rc}")
    return out


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
    out = np.empty(bgr.shape[:-1], np.uint8)
    lib().me_oracle_bgr2gray(_p(bgr, C.c_uint8), out.size, _p(out, C.c_uint8))
    return out


def max_threads() -> int:
    return int(lib().me_oracle_max_threads())
