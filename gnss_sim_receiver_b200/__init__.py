"""B200-native (sm_100a) PCPS acquisition and tracking multicorrelator behind the reference's block API.

Product code = include/gnss_b200.h (C ABI) + csrc/ (CUDA) + host/ (C++ adapters).  The Python modules here are
ctypes plumbing for tests and bench.py.  Nothing in this package imports or links oracle/.
"""
from . import _capi  # noqa: F401

__all__ = ["_capi"]
