"""spmv_b200 -- B200-native (sm_100a) hot path of LIBSPMV (Excalibur-SLE/spmv).

Distributed fp64 CSR SpMV (general and symmetric storage), its halo exchange and
the CG iteration, as hand-written CUDA behind a C ABI (``include/b200spmv.h``)
with a C++ host layer that mirrors the reference's API (``include/spmv/``).

This Python package is the thin harness over those libraries (ctypes): it is what
``tests/`` and ``bench.py`` drive.  There is no CPU compute path: loading fails
loudly when the CUDA library has not been built.
"""
from ._lib import lib, check, B200Error, LIB_PATH  # noqa: F401
