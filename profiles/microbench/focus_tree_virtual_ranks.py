"""The MAC-coarsened focus trees (sphb200_domain_update_focus) of P virtual ranks on one GPU: leaves per rank against
the guided bucket-64 tree the step path uses, peers, and the cost of the first call (converge) and of one update.

  python profiles/microbench/focus_tree_virtual_ranks.py sedov 200 8 1.0 [out.json]

side = GLOBAL lattice side (200 -> 8 M particles over P ranks)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.environ.get("GRAFT_REPO_ROOT", "/root/repo")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import sphexa_b200_loader  # noqa: E402

pkg = sphexa_b200_loader.load()
from test_gpu_domain_dist import make_group  # noqa: E402

case, side, P, theta = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), float(sys.argv[4])
ctx = pkg.Context(0)
group, comm, f, box, s = make_group(pkg, ctx, case, side, P, capacity_factor=1.6)
del f
for _ in range(2):
    group.step()
group.sync()
torch.cuda.synchronize()
launches0 = sum(r.ctx.launches for r in group.ranks)
t0 = time.perf_counter()
conv = group.updateFocus(theta)
torch.cuda.synchronize()
first_ms = (time.perf_counter() - t0) * 1e3
launches_first = sum(r.ctx.launches for r in group.ranks) - launches0
update_ms = []
for _ in range(3):
    group.step()
    group.sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    group.updateFocus(theta)
    torch.cuda.synchronize()
    update_ms.append((time.perf_counter() - t0) * 1e3)
infos = [r.info() for r in group.ranks]
trees = [r.focusTree() for r in group.ranks]
out = {
    "case": case, "global_side": side, "ranks": P, "theta": theta, "converged_after_first_call": bool(conv),
    "first_call_ms_all_ranks": first_ms, "launches_first_call_all_ranks": launches_first,
    "update_ms_all_ranks": update_ms, "update_ms_per_rank": float(np.median(update_ms)) / P,
    "owned": [int(i.endIndex - i.startIndex) for i in infos],
    "global_tree_leaves": int(infos[0].numGlobalLeaves),
    "guided_tree_leaves": [int(i.numFocusLeaves) for i in infos],
    "search_tree_leaves": [int(i.numSearchLeaves) for i in infos],
    "mac_focus_tree_leaves": [int(t[0].size - 1) for t in trees],
    "mac_focus_leaves_inside_own_range": [
        int(np.searchsorted(t[0], i.assignment[i.rank + 1]) - np.searchsorted(t[0], i.assignment[i.rank]))
        for t, i in zip(trees, infos)],
    "peers": [t[2] for t in trees],
    "max_leaf_count_outside": [int(np.delete(t[1], np.s_[np.searchsorted(t[0], i.assignment[i.rank]):
                                                         np.searchsorted(t[0], i.assignment[i.rank + 1])]).max())
                               for t, i in zip(trees, infos)],
}
print(json.dumps(out), flush=True)
if len(sys.argv) > 5:
    with open(sys.argv[5], "w") as fh:
        json.dump(out, fh, indent=1)
