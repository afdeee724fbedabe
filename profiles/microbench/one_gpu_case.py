"""one-GPU step time of a case at the size one rank of an N-rank run owns (weak-scaling reference)"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import sphexa_b200_loader
pkg = sphexa_b200_loader.load()
from sph_exa_b200 import sim as S
case, side, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 10
f, box, s = pkg.ic.CASES[case](side)
n = f["x"].size
ctx = pkg.Context(0)
gs = S.HydroVeProp(ctx, box, s, capacity=n)
gs.upload(f)
for _ in range(3): gs.step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps): gs.step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(f"ONE_GPU {case} side {side}: {n} particles, {ms:.3f} ms/step, {n / ms / 1e3:.1f} M particle-steps/s", flush=True)
