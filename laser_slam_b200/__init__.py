"""laser_slam_b200 -- B200-native batched 2D scan matching behind the reference's matcher entry points.

The product is libzhsm.so (CUDA, sm_100a) behind the C ABI in include/zhsm.h.  This package is only the
thin ctypes binding used by tests/ and bench.py; it never imports anything under oracle/ and it has no CPU
fallback: importing works anywhere, computing raises unless the library loads and a B200 is present.
"""
from .engine import (Engine, Laser, PlicpParams, RESULT_DTYPE, ZhsmError, lib_path, load_library,  # noqa: F401
                     LASER_TABLE, laser_from_table)

__all__ = ["Engine", "Laser", "PlicpParams", "RESULT_DTYPE", "ZhsmError", "lib_path", "load_library",
           "LASER_TABLE", "laser_from_table"]
