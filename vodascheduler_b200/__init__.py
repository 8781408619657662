"""vodascheduler_b200 — B200-native engine for Voda's allocation + placement hot path.

The compute lives in csrc/ (hand-written CUDA for sm_100a behind the C ABI of
include/voda_accel.h).  This package is the thin host side: batch layout, the ctypes
binding, and mirrors of the reference's SchedulerAlgorithm / munkres / PlacementManager
interfaces.  There is no CPU fallback: importing is cheap, but creating a Context without
the built library or without a CUDA device raises.
"""
from ._abi import ALGO_NAMES  # noqa: F401
from .batch import AllocBatch, PlaceBatch, PlaceOut  # noqa: F401


def __getattr__(name):
    if name in ("Context", "EngineError", "library_path", "load_library"):
        from . import _lib
        return getattr(_lib, name)
    if name in ("NewAlgorithmFactory", "TrainingJob", "JobInfo", "JobConfig", "SchedulerAlgorithm"):
        from . import algorithm
        return getattr(algorithm, name)
    raise AttributeError(name)
