"""virtual ranks on one GPU with the sync stage profile (SPHB200_SYNC_PROFILE=1)"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
sys.path.insert(0, os.path.join(os.environ.get("GRAFT_REPO_ROOT", "/root/repo"), "tests"))
import sphexa_b200_loader
pkg = sphexa_b200_loader.load()
from test_gpu_domain_dist import make_group
case, side, P, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
ctx = pkg.Context(0)
group, comm, f, box, s = make_group(pkg, ctx, case, side, P, capacity_factor=1.6)
for _ in range(steps):
    group.step()
torch.cuda.synchronize()
infos = [r.info() for r in group.ranks]
print("own", [i.endIndex - i.startIndex for i in infos], "boxes", [i.numSurfaceBoxes for i in infos], "migrated in", [i.migratedIn for i in infos], flush=True)
