"""cuda_iblb_11_b200 — B200-native immersed-boundary lattice Boltzmann time-step path.

The product is `libiblb.so` (C ABI in include/iblb.h: hand-written sm_100a CUDA + C++ host code);
this package is its thin ctypes mirror plus the in-tree build recipe.  Nothing here imports the
oracle, and nothing falls back to the CPU.
"""
from .solver import (FLAG_DEV_KNOBS, FLAG_HALF_FORCE_BODY, FLAG_MUCUS, FLAG_NO_FUSE, FLAG_NO_GRAPH, FLAG_NO_PDL, FLAG_PICK_HIGHEST_K, LIB_PATH, SYMBOLS,
                     HaloHandle, IblbError, debug_violations, Params, Solver, lib, run)

__all__ = ["FLAG_DEV_KNOBS", "FLAG_HALF_FORCE_BODY", "FLAG_MUCUS", "FLAG_NO_FUSE", "FLAG_NO_GRAPH", "FLAG_NO_PDL", "FLAG_PICK_HIGHEST_K", "LIB_PATH", "SYMBOLS",
           "HaloHandle", "IblbError", "Params", "Solver", "debug_violations", "lib", "run"]
