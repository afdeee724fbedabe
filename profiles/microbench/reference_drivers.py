#!/usr/bin/env python3
"""Kernel-level comparison with the reference's own GPU code on the same B200.

Runs the reference's self-checking performance drivers (domain/test/performance/{neighbor_driver,hilbert,octree}.cu,
built for sm_100a by oracle/build_cstone_perf.sh into baseline/_ref/) and this library's entry points on the same
workloads: same sizes, same distributions (uniform / clipped Gaussian as in domain/test/coord_samples/random.hpp),
same bucket sizes, h and ngmax.  The reference drivers time ONE call each (CUDA events around it, first launch of the
process included), so both the first call and the best of five warm calls are given for this library.

  python profiles/microbench/reference_drivers.py > profiles/reference_drivers_r02.json
"""
import ctypes as C
import json
import os
import re
import subprocess
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import sphexa_b200_loader  # noqa: E402

pkg = sphexa_b200_loader.load()
from sph_exa_b200 import ops, binding as B  # noqa: E402

REF = os.path.join(ROOT, "baseline", "_ref")
ctx = pkg.Context(0)
dev = torch.device("cuda", 0)


def run_ref(name):
    p = os.path.join(REF, name)
    if not os.path.exists(p):
        return None
    res = subprocess.run([p], capture_output=True, text=True, timeout=1200)
    return res.stdout + res.stderr


def timed(fn, warm=5):
    """(first call, best warm call) in ms, CUDA events on the current stream"""
    out = []
    for _ in range(1 + warm):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b))
    return out[0], min(out[1:])


def sfc_sorted(x, y, z, box):
    keys = ops.computeSfcKeysGpu(ctx, x, y, z, box)
    perm = ops.sortByKeyGpu(ctx, keys)
    idx = perm.long()
    return x[idx].contiguous(), y[idx].contiguous(), z[idx].contiguous(), keys


result = {}

# ---- neighbor search: 2 M uniform particles in the periodic unit box, h = 0.012, ngmax = 200, bucket 64
out = run_ref("cstone_neighbors_perf")
ref = None
if out:
    m = re.search(r"GPU time ([0-9.eE+-]+) s", out)
    ref = {"findNeighborsBT_ms": float(m.group(1)) * 1e3 if m else None,
           "self_check": "PASS" if "Neighbor counts: PASS" in out else "see log"}
rng = np.random.Generator(np.random.MT19937(42))
n = 2_000_000
x, y, z = (torch.from_numpy(rng.uniform(0.0, 1.0, n)).to(dev) for _ in range(3))
box = B.make_box([0, 1, 0, 1, 0, 1], [1, 1, 1])
x, y, z, keys = sfc_sorted(x, y, z, box)
leaves, counts, _ = ops.computeOctreeGpu(ctx, keys, 64)
tree = ops.Octree(ctx, leaves, box, counts)
h = torch.full((n,), 0.012, dtype=torch.float64, device=dev)
state = {}


def search():
    state["nb"], state["nc"], state["fails"] = ops.findNeighbors(ctx, tree, x, y, z, h, 0, n, box, 100, 200,
                                                                 iterate_h=False)


first, best = timed(search)
nc = state["nc"].long()
result["neighbor_search"] = {"workload": "2 M uniform particles, periodic unit box, h = 0.012, ngmax 200, bucket 64, "
                                         "stored lists", "reference": ref,
                             "ours": {"first_call_ms": first, "best_warm_ms": best,
                                      "mean_neighbors": float((nc - 1).float().mean())}}
del tree, state, x, y, z, h, keys
torch.cuda.empty_cache()

# ---- SFC keys: 32 M Hilbert keys (the reference driver starts from integer coordinates, this library from doubles)
out = run_ref("cstone_hilbert_perf")
ref = None
if out:
    m = re.search(r"compute time for (\d+) hilbert keys: ([0-9.eE+-]+) s", out)
    ref = {"hilbert_keys_ms": float(m.group(2)) * 1e3 if m else None, "keys": int(m.group(1)) if m else None,
           "input": "3 x uint32 per key"}
n = 32_000_000
x, y, z = (torch.from_numpy(rng.uniform(-1.0, 1.0, n)).to(dev) for _ in range(3))
box = B.make_box([-1, 1, -1, 1, -1, 1], [0, 0, 0])
kbuf = torch.empty(n, dtype=torch.int64, device=dev)
first, best = timed(lambda: ops.computeSfcKeysGpu(ctx, x, y, z, box, kbuf))
result["sfc_keys"] = {"workload": "32 M Hilbert keys", "reference": ref,
                      "ours": {"first_call_ms": first, "best_warm_ms": best, "input": "3 x double per key"}}
del x, y, z, kbuf
torch.cuda.empty_cache()

# ---- octree: 2 M clipped-Gaussian particles in (-1, 1)^3, bucket 16
out = run_ref("cstone_octree_perf")
ref = None
if out:
    def grab(pat):
        m = re.search(pat, out)
        return float(m.group(1)) * 1e3 if m else None
    ref = {"build_from_scratch_ms": grab(r"build time from scratch ([0-9.eE+-]+)"),
           "update_with_guess_ms": grab(r"build time with guess ([0-9.eE+-]+)"),
           "internal_build_ms": grab(r"internal build time ([0-9.eE+-]+)"),
           "halo_discovery_ms": grab(r"\nhalo discovery ([0-9.eE+-]+)"),
           "leaves": int(re.search(r"build time from scratch [0-9.eE+-]+ nNodes\(tree\): (\d+)", out).group(1))}
n = 2_000_000
x, y, z = (torch.from_numpy(np.clip(rng.normal(0.0, 0.4, n), -1.0, 1.0)).to(dev) for _ in range(3))
x, y, z, keys = sfc_sorted(x, y, z, box)
st = {}


def build():
    st["leaves"], st["counts"], st["it"] = ops.computeOctreeGpu(ctx, keys, 16)


b_first, b_best = timed(build)
u_first, u_best = timed(lambda: ops.updateOctreeGpu(ctx, keys, 16, st["leaves"], st["counts"]))
i_first, i_best = timed(lambda: st.__setitem__("tree", ops.Octree(ctx, st["leaves"])))
tree = st["tree"]
nl = tree.numLeaves
radii = torch.full((nl,), 0.01, dtype=torch.float32, device=dev)
flags = torch.zeros(nl, dtype=torch.int32, device=dev)
p = lambda t: C.c_void_p(t.data_ptr())


def halos():
    flags.zero_()
    B.check(B.lib.sphb200_find_halos(ctx.h, p(tree.prefixes), p(tree.childOffsets), p(tree.internalToLeaf),
                                     p(st["leaves"]), p(radii), C.byref(box), 0, nl // 4, p(flags), None))


h_first, h_best = timed(halos)
result["octree"] = {"workload": "2 M clipped-Gaussian particles, bucket 16", "reference": ref,
                    "ours": {"leaves": int(nl), "build_from_scratch_ms": [b_first, b_best],
                             "update_with_guess_ms": [u_first, u_best], "internal_build_ms": [i_first, i_best],
                             "halo_discovery_ms": [h_first, h_best], "halo_leaves_flagged": int(flags.sum()),
                             "note": "[first call, best of 5 warm calls]; the python wrappers allocate their outputs "
                                     "inside the timed call"}}
print(json.dumps(result, indent=1))
