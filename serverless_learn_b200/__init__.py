"""serverless_learn_b200 — B200-native hot path of sheaconlon/serverless_learn.

The product is the C ABI in include/sl_gpu.h (libsl_gpu.so: hand-written sm_100a
kernels) with the C++ role cores of include/sl_roles.h on top; this Python package is
a thin ctypes veneer over both, used by the tests and bench.py.  Nothing here computes
on the CPU: without the built library and a CUDA device every call raises.
"""
from .api import Replica, SlError  # noqa: F401

__all__ = ["Replica", "SlError"]
