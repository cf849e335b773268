"""harfang3d_b200 — B200-native scene-update path for HARFANG scenes.

The product is the native library ``libhgb200.so`` (hand-written sm_100a CUDA kernels behind the C ABI declared in
``include/hgb200.h``) plus the C++ ``hg::Scene`` facade in ``csrc/facade``. This Python package is only the thin ctypes binding the
tests and ``bench.py`` drive it through; it contains no compute and there is NO CPU fallback: importing works anywhere (so the
symbol table can be checked without a GPU), creating a context without a CUDA device raises.
"""
from .capi import Context, HgbError, lib, library_path, Track  # noqa: F401

__all__ = ["Context", "HgbError", "lib", "library_path", "Track"]
