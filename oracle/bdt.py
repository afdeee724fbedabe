"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the block-time-step building blocks of the reference.

PARITY UNPINNED for the pieces marked (gpu-only): the reference implements them only as CUDA kernels
(`HydroVeBdtProp` throws on CPUs, ve_hydro_bdt.hpp:118), which cannot run where this oracle is built, and its tests
hold no golden vectors for them.  They are restated here from the kernel sources, function by function, with the
file:line each one follows.  Pieces built from host-callable reference functions (positionUpdate, energyUpdate,
findRungRanges) are pinned through oracle/_ref instead (oracle/ref.py).

Only tests/ may import this module.
"""
import numpy as np

GROUP_SIZE = 32  # TravConfig::targetSize = GpuConfig::warpSize (cstone/traversal/find_neighbors.cuh:62)


def leaf_levels(leaves):
    """treeLevel of every leaf: key range = 8^(21 - level)"""
    rng = np.diff(leaves.astype(np.uint64))
    log2 = np.array([int(r).bit_length() - 1 for r in rng])
    return 21 - log2 // 3


def compute_groups(first, last, x, y, z, leaves, layout, box_lim, tol_factor=2.0):
    """(gpu-only) cstone::computeGroupSplits (domain/include/cstone/traversal/groups.cuh:150-330) as used by
    sph::computeSpatialGroups (sph/include/sph/groups.cu:30-47): returns the group delimiters.

    groupSplitsKernel: per fixed run of 32 bodies, nodeVolume = warpMin of the volume of each body's leaf cell in the
    unit cube (:213-222), distCrit = cbrt(nodeVolume) * tolFactor (:223); bodies are compared in box-normalized
    coordinates pos = x * ilx (:229); findSplits sets a split after lane l if |pos[l+1] - pos[l]|^2 > distCrit^2 (:96-103).
    makeSplits + two scans turn the masks into delimiters (:108-135, :311-327)."""
    n = last - first
    if n == 0:
        return np.array([first], np.uint32)
    inv = [1.0 / (box_lim[1] - box_lim[0]), 1.0 / (box_lim[3] - box_lim[2]), 1.0 / (box_lim[5] - box_lim[4])]
    levels = leaf_levels(leaves)
    idx = np.arange(first, last)
    leaf_of = np.searchsorted(layout, idx, side="right") - 1
    lvl = levels[leaf_of]
    runs = (idx - first) // GROUP_SIZE
    max_level = np.zeros(runs.max() + 1, np.int64)
    np.maximum.at(max_level, runs, lvl)
    dist_crit = np.ldexp(1.0, -max_level) * np.float64(np.float32(tol_factor))
    px, py, pz = x[first:last] * inv[0], y[first:last] * inv[1], z[first:last] * inv[2]
    dx, dy, dz = np.diff(px), np.diff(py), np.diff(pz)
    d2 = dx * dx + dy * dy + dz * dz
    split = np.zeros(n, bool)  # split[t]: t starts a group
    split[1:] = d2 > dist_crit[runs[1:]] ** 2
    split[(idx - first) % GROUP_SIZE == 0] = True
    starts = idx[split]
    return np.concatenate([starts, [last]]).astype(np.uint32)


def group_divv_timestep(Krho, start, end, divv, group_dt):
    """(gpu-only) groupDivvKernel (sph/include/sph/ts_groups.cu:16-35): float running max from -inf over divv"""
    out = group_dt.copy()
    for g, (s, e) in enumerate(zip(start, end)):
        local_max = np.float32(-np.inf)
        if e > s:
            local_max = max(local_max, np.float32(divv[s:e].astype(np.float32).max()))
        with np.errstate(divide="ignore"):
            out[g] = min(out[g], np.float32(Krho) / np.abs(local_max))
    return out


def group_acc_timestep(eta_acc, start, end, ax, ay, az, group_dt):
    """(gpu-only) groupAccKernel (ts_groups.cu:49-69): max |a|^2 as float, dt = etaAcc / sqrt(sqrt(maxAcc))"""
    out = group_dt.copy()
    a2 = ax * ax + ay * ay + az * az
    for g, (s, e) in enumerate(zip(start, end)):
        m = np.float32(0)
        if e > s:
            m = max(m, np.float32(a2[s:e].astype(np.float32).max()))
        with np.errstate(divide="ignore"):
            out[g] = min(out[g], np.float32(eta_acc) / np.sqrt(np.sqrt(m, dtype=np.float32), dtype=np.float32))
    return out


def group_min_timestep(start, end, dt, group_dt):
    """(gpu-only) warpMin of the per-lane Courant steps folded into groupDt (momentum_energy_gpu.cu:98-104)"""
    out = group_dt.copy()
    for g, (s, e) in enumerate(zip(start, end)):
        if e > s:
            out[g] = min(out[g], np.float32(dt[s:e].min()))
    return out


def butterfly(i):
    """cstone::butterfly (domain/include/cstone/primitives/math.hpp:26-31): 0 for 0, else 1 + countTrailingZeros"""
    return 0 if i == 0 else (i & -i).bit_length()


def active_rung(substep, num_rungs):
    """HydroVeBdtProp::activeRung (ve_hydro_bdt.hpp:106-110)"""
    if substep == 0 or substep >= (1 << (num_rungs - 1)):
        return 0
    return butterfly(substep)
