"""entropy_coding_b200 - B200-native VVC CABAC engine (hot path of p-sawicki/entropy_coding).

The product is entropy_coding_b200/libcabac_b200.so (hand-written sm_100a CUDA
behind the C ABI of include/cabac_b200.h).  This package is the thin Python
face of that ABI plus the synthetic-workload generator used by tests and
bench.py.  There is no CPU implementation in here.
"""
from .engine import CabacEngine, CabacError  # noqa: F401
from . import sharding, types, workload  # noqa: F401

__all__ = ["CabacEngine", "CabacError", "sharding", "types", "workload"]
