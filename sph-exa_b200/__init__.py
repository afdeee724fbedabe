"""sph-exa_b200: B200-native (sm_100a) hot path of SPH-EXA behind a C ABI.

Layout: csrc/ (CUDA kernels + C ABI -> libsphexa_b200.so), binding.py (ctypes),
ops.py (operator-level mirror of the reference GPU interface), sim.py (Domain /
HydroVeProp mirror), ic.py (host-side initial conditions and constants).
The directory name contains a dash, so load it with importlib (see sphexa_b200_loader.py
at the repo root) under the module name `sph_exa_b200`.
"""
from . import binding, ic  # noqa: F401
from .binding import F64, MIXED, OPEN, PERIODIC, FIXED, SphError, Context, make_box  # noqa: F401
