"""osqp_b200 -- B200-native OSQP ADMM hot path (C host + hand-written CUDA, sm_100a only).

The product is two shared libraries built in-tree from osqp_b200/csrc:
  lib/libosqp_b200.so      plain-C OSQP API (include/osqp.h)
  lib/libosqp_cuda_pcg.so  CUDA engine + cuda_pcg linear-system plugin (include/osqp_cuda_pcg.h)
This Python package is a thin ctypes harness over that C ABI for tests and bench.py.
"""
from .capi import Solver, load_library, PRODUCT_LIB, PLUGIN_LIB  # noqa: F401
from . import problems  # noqa: F401
