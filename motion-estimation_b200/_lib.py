"""ctypes binding of libme_b200.so -- the C ABI declared in include/me_b200.h.

There is no CPU fallback here by design: if the CUDA library has not been built, or cannot be
loaded, importing the binding raises with the build command to run.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.path.join(_HERE, "libme_b200.so")
HEADER = os.path.normpath(os.path.join(_HERE, "..", "include", "me_b200.h"))

NVCC_COMPILE_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
NVCC_LINK_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC"]


class MEError(RuntimeError):
    """A libme_b200 call returned a negative status."""

    def __init__(self, code, msg):
        super().__init__(f"libme_b200 error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("block", C.c_int32), ("range", C.c_int32), ("min_range", C.c_int32),
                ("subpixel", C.c_int32), ("min_valid", C.c_int32), ("min_curv", C.c_int32), ("range_y", C.c_int32)]


class PairResult(C.Structure):
    _fields_ = [("pdx", C.c_float), ("pdy", C.c_float), ("flag", C.c_int32),
                ("nvalid_x", C.c_int32), ("nvalid_y", C.c_int32), ("n_textured", C.c_int32),
                ("mean_dx", C.c_float), ("mean_dy", C.c_float)]


class ScanRecord(C.Structure):
    _fields_ = [("anchor_slot", C.c_int32), ("accepted", C.c_int32), ("pair", PairResult)]


class Geometry(C.Structure):
    _fields_ = [("height", C.c_int32), ("width", C.c_int32), ("nbx", C.c_int32), ("nby", C.c_int32),
                ("max_frames", C.c_int32), ("max_pairs", C.c_int32), ("tile_blocks_x", C.c_int32),
                ("threads_per_cta", C.c_int32), ("smem_bytes", C.c_int32), ("tiles_per_pair", C.c_int32),
                ("cand_px_per_pair", C.c_int64), ("cand_px_per_pair_valid", C.c_int64),
                ("sm_count", C.c_int32), ("uses_tma", C.c_int32), ("slabs_x", C.c_int32), ("slabs_y", C.c_int32),
                ("window_pitch", C.c_int32), ("resident_kernel", C.c_int32), ("tile_blocks_y", C.c_int32)]


INST_PITCHES = (144, 208, 240)
INST_BLOCKS = (8, 16, 32)
BUILD_DIR = os.path.join(_HERE, "build")


def sources():
    names = ("me_api.cu", "search_inst.cu", "sad_search.cuh", "sad_persistent.cuh", "sad_scan.cuh", "kernel_choice.h", "misc_kernels.cuh", "clahe.cuh", "nlm.cuh")
    return [os.path.join(_CSRC, n) for n in names] + [HEADER]


def _compile_jobs():
    """(object file, nvcc arguments) per translation unit: the host side + small kernels, and the search kernels of
    one (window pitch, block size) each, so that the build uses every core."""
    jobs = [(os.path.join(BUILD_DIR, "me_api.o"), [os.path.join(_CSRC, "me_api.cu")])]
    for p in INST_PITCHES:
        for b in INST_BLOCKS:
            jobs.append((os.path.join(BUILD_DIR, f"search_{p}_{b}.o"),
                         [f"-DME_INST_PITCH={p}", f"-DME_INST_B={b}", os.path.join(_CSRC, "search_inst.cu")]))
    return jobs


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/ for sm_100a with nvcc into libme_b200.so next to this file (in-tree): ten translation units
    in parallel, then one link."""
    from concurrent.futures import ThreadPoolExecutor
    srcs = sources()
    newest = max(os.path.getmtime(s) for s in srcs)
    if not force and os.path.exists(SO_PATH) and os.path.getmtime(SO_PATH) >= newest:
        return SO_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    extra = os.environ.get("ME_NVCC_EXTRA", "").split()             # experiments only, e.g. -DME_CW=16
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    os.makedirs(BUILD_DIR, exist_ok=True)
    stamp = os.path.join(BUILD_DIR, "flags.txt")
    flags_now = " ".join(NVCC_COMPILE_FLAGS + extra)
    if not os.path.exists(stamp) or open(stamp).read() != flags_now:
        force = True

    def compile_one(job):
        obj, args = job
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= newest:
            return
        cmd = [nvcc] + NVCC_COMPILE_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj] + args
        subprocess.run(cmd, check=True, env=env)

    jobs = _compile_jobs()
    with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as pool:
        list(pool.map(compile_one, jobs))
    open(stamp, "w").write(flags_now)
    subprocess.run([nvcc] + NVCC_LINK_FLAGS + ["-o", SO_PATH] + [j[0] for j in jobs], check=True, env=env)
    return SO_PATH


_lib = None


def lib():
    """The loaded library with argtypes set.  Raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: the CUDA extension has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc; no GPU needed to compile). "
            "There is deliberately no CPU fallback for the displacement search.")
    L = C.CDLL(SO_PATH)
    vp, i32p, u8p = C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
    L.me_last_error.restype = C.c_char_p
    L.me_version.restype = C.c_char_p
    L.me_create.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(Params), C.c_int, C.c_int, C.POINTER(vp)]
    L.me_destroy.argtypes = [vp]
    L.me_get_geometry.argtypes = [vp, C.POINTER(Geometry)]
    L.me_upload.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int]
    L.me_upload_bgr.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int]
    L.me_clahe.argtypes = [vp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int]
    L.me_denoise.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_int32), C.c_float, C.c_int, C.c_int]
    L.me_download.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int]
    L.me_search.argtypes = [vp, i32p, C.c_int]
    L.me_fetch_pairs.argtypes = [vp, C.POINTER(PairResult), C.c_int]
    L.me_fetch_pairs_ticket.argtypes = [vp, C.c_int, C.POINTER(PairResult), C.c_int]
    L.me_last_ticket.argtypes = [vp]
    L.me_last_ticket.restype = C.c_int
    L.me_fetch_blocks.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    L.me_estimate_pair.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(PairResult)]
    L.me_search_device.argtypes = [vp, vp, C.c_size_t, C.c_int, i32p, C.c_int, vp]
    L.me_sync.argtypes = [vp]
    L.me_last_search_ms.argtypes = [vp, C.POINTER(C.c_float)]
    L.me_timer_start.argtypes = [vp]
    L.me_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    L.me_launch_count.argtypes = [vp]
    L.me_launch_count.restype = C.c_int64
    L.me_microbench_vabsdiff4.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                          C.POINTER(C.c_int)]
    L.me_copy_timer_start.argtypes = [vp]
    L.me_copy_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
    L.me_anchored_scan.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.me_copy_pair_records_device.argtypes = [vp, vp, C.c_int, vp]
    L.me_copy_frames_device.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int, vp]
    L.me_search_static_scratch.argtypes = [vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(Params), i32p, C.c_int,
                                           vp, vp]
    L.me_search_static_scratch.restype = C.c_int
    for name in ("me_clahe", "me_denoise", "me_download", "me_create", "me_destroy", "me_get_geometry", "me_upload", "me_upload_bgr", "me_search",
                 "me_fetch_pairs", "me_fetch_blocks", "me_estimate_pair", "me_search_device", "me_sync",
                 "me_last_search_ms", "me_microbench_vabsdiff4", "me_timer_start", "me_timer_stop", "me_copy_timer_start",
                 "me_copy_timer_stop", "me_anchored_scan", "me_copy_pair_records_device", "me_copy_frames_device",
                 "me_fetch_pairs_ticket"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L


def check(rc):
    if rc < 0:
        raise MEError(rc, lib().me_last_error().decode("utf-8", "replace"))
    return rc


def exported_symbols():
    """Names declared in include/me_b200.h (parsed from the header, used by the ABI test)."""
    import re
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(me_[a-z0-9_]+)\s*\(", text)))
