#!/usr/bin/env python3
"""A few steps of one of the initial conditions on one GPU with the per-phase device times: the workload behind the
non-lattice captures in this directory.

  python profiles/cases.py <sedov|noh|turbulence|kh> <side> <stir 0|1> [warm-up steps]
  ATMIN=5 python profiles/cases.py turbulence 200 0      # Atwood ramp switched off
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sphexa_b200_loader  # noqa: E402

pkg = sphexa_b200_loader.load()
from sph_exa_b200 import sim as S  # noqa: E402
from sph_exa_b200.turbulence import Turbulence  # noqa: E402

ctx = pkg.Context(0)
case, side, stir = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
warm = int(sys.argv[4]) if len(sys.argv) > 4 else 4
f, box, s = pkg.ic.CASES[case](side, "f64")
if os.environ.get("ATMIN"):
    s["Atmin"] = float(os.environ["ATMIN"])
    s["Atmax"] = 2 * s["Atmin"]
n = f["x"].size
gs = S.HydroVeProp(ctx, box, s, capacity=n)
if stir:
    gs.attachTurbulence(Turbulence())
gs.upload(f)
for _ in range(warm):
    gs.step()
ph = {}
for _ in range(3):
    gs.step()
    for k, v in gs.phaseTimes().items():
        ph[k] = ph.get(k, 0) + v / 3
print(case, n, "stir", stir, {k: round(v, 2) for k, v in ph.items()}, flush=True)
