"""per-phase device times of one-GPU steps: python phase_times.py case side [steps]  (SPHB200_LIB selects the library)"""
import os, sys, json, numpy as np, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import sphexa_b200_loader
pkg = sphexa_b200_loader.load()
from sph_exa_b200 import sim as S
case, side, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 6
f, box, s = pkg.ic.CASES[case](side)
n = f["x"].size
ctx = pkg.Context(0)
gs = S.HydroVeProp(ctx, box, s, capacity=n)
gs.upload(f)
for _ in range(3): gs.step()
torch.cuda.synchronize()
acc = {}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    gs.step()
    for k, v in gs.phaseTimes().items():
        acc[k] = acc.get(k, 0.0) + v / steps
e1.record(); torch.cuda.synchronize()
print(json.dumps({"lib": os.environ.get("SPHB200_LIB", "default"), "case": case, "side": side, "n": n,
                  "ms_per_step": e0.elapsed_time(e1) / steps, "total_neighbors": int(gs.totalNeighbors()),
                  "phase_ms": {k: round(v, 3) for k, v in acc.items()}}), flush=True)
