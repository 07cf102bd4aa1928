"""motion-estimation_b200 -- B200-native displacement search for ruzko/motion-estimation's hot path.

The directory name carries a hyphen (layout contract); import it as ``motion_estimation_b200``
through the shim module of that name at the repository root.

Public surface (mirrors the reference's call sites):

* ``calc_ECC_transform(prev, cur)``            estimate_vel_vid.py:306 / estimate_vel.py:286
* ``end_process(clean_arrs, ...)``             estimate_vel_vid.py:493 / estimate_vel.py:333
* ``MotionEstimator`` / ``SearchConfig``       the batched engine underneath (C ABI: include/me_b200.h)
"""
from ._lib import MEError, build  # noqa: F401
from .engine import (FLAG_TEXTURED, FLAG_X_INTERIOR, FLAG_X_PEAKED, FLAG_Y_INTERIOR, FLAG_Y_PEAKED,  # noqa: F401
                     PAIR_DTYPE, MotionEstimator, SearchConfig, microbench_vabsdiff4)

__all__ = ["MotionEstimator", "SearchConfig", "MEError", "build", "microbench_vabsdiff4", "PAIR_DTYPE"]
