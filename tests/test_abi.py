"""CPU tests of the C-ABI boundary: the shared library loads, exports every symbol the header
declares, reports errors instead of crashing when there is no GPU, and the Python binding refuses
to work without it (no CPU fallback)."""
import ctypes as C
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_are_exported(me):
    L = me._lib.lib()
    names = me._lib.exported_symbols()
    assert {"me_create", "me_destroy", "me_upload", "me_search", "me_fetch_pairs", "me_fetch_blocks",
            "me_estimate_pair", "me_search_device", "me_last_error"} <= set(names)
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/me_b200.h but not exported by libme_b200.so"
    out = subprocess.run(["nm", "-D", "--defined-only", me._lib.SO_PATH], capture_output=True, text=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    assert set(names) <= exported
    assert b"sm_100a" in L.me_version()


def test_struct_layouts_match_the_header(me):
    assert C.sizeof(me._lib.PairResult) == 32 and me.PAIR_DTYPE.itemsize == 32
    assert C.sizeof(me._lib.Params) == 28
    assert me._lib.Geometry.cand_px_per_pair.offset % 8 == 0


def test_argument_errors_do_not_need_a_gpu(me):
    L = me._lib.lib()
    ctx = C.c_void_p()
    prm = me._lib.Params(12, 8, 2, 1, 1, 1)                       # block 12 is not a built size
    assert L.me_create(0, 64, 64, C.byref(prm), 4, 4, C.byref(ctx)) == -2
    assert b"block size 12" in L.me_last_error()
    prm = me._lib.Params(16, 0, 2, 1, 1, 1)
    assert L.me_create(0, 64, 64, C.byref(prm), 4, 4, C.byref(ctx)) == -1
    prm = me._lib.Params(16, 8, 2, 1, 1, 1)
    assert L.me_create(0, 8, 64, C.byref(prm), 4, 4, C.byref(ctx)) == -1 and ctx.value is None
    assert L.me_search(None, None, 1) == -1
    assert L.me_destroy(None) == 0


def test_no_gpu_is_an_error_not_a_fallback(me):
    import torch
    if torch.cuda.is_available():
        pytest.skip("this check is for the CPU-only container")
    with pytest.raises(me.MEError) as e:
        me.MotionEstimator(64, 64, me.SearchConfig(16, 8), max_frames=2, max_pairs=1)
    assert e.value.code == -3                                     # ME_ECUDA: no device, and no CPU path to fall back to


def test_missing_library_fails_loudly():
    code = ("import motion_estimation_b200 as m, os\n"
            "m._lib.SO_PATH = os.path.join(os.path.dirname(m._lib.SO_PATH), 'absent.so')\n"
            "m._lib._lib = None\n"
            "try:\n    m._lib.lib()\nexcept ImportError as e:\n    print('IMPORTERROR', 'no CPU fallback' in str(e))\n")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, cwd=ROOT)
    assert "IMPORTERROR True" in out.stdout, out.stdout + out.stderr


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "motion-estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "sad_oracle" not in text.replace("oracle/sad_oracle.c", ""), f
    for f in ("estimate_vel_vid.py", "estimate_vel.py"):
        assert "oracle" not in open(os.path.join(ROOT, f)).read()
