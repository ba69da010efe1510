"""irfinder_b200 — B200-native drop-in for IRFinder's BAM -> intron-retention counting core.

The product is the C-ABI shared library ``libirfinder_b200.so`` (CUDA, sm_100a; declared in
``include/irfinder_b200.h``).  This package is the thin Python host mirror of the reference's
``irfinder`` main (upstream src/irfinder/irfinder.cpp): load the four
BuildRef-format reference files, run a BAM, write the six output tables.  It binds the library with
ctypes and never computes anything itself: if the library is missing or no GPU is present, every
operation raises.
"""
from .api import IRFinderCore, IRFError, depth_scan_static, lib_path, load_library, TABLES  # noqa: F401

__all__ = ["IRFinderCore", "IRFError", "depth_scan_static", "lib_path", "load_library", "TABLES"]
