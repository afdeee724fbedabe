"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the block-time-step building blocks of the reference.

The pieces marked (gpu-only) exist in the reference only as CUDA kernels (`HydroVeBdtProp` throws on CPUs,
ve_hydro_bdt.hpp:118), which cannot run where this oracle is built, and its tests hold no golden vectors for them.
They are restated here from the kernel sources, function by function, with the file:line each one follows, and are
pinned INDIRECTLY on the GPU box: tests/test_gpu_reference_kernels.py compares the library with the reference's own
kernels compiled for sm_100a (computeGroupSplits, groupDivvTimestepGpu, groupAccTimestepGpu) and
tests/test_gpu_groups.py compares the same library functions with this restatement.  PINNED through oracle/_ref
(tests/test_bdt_oracle.py): the per-particle updates built from the reference's positionUpdate / energyUpdate, and the time-step hierarchy -- rung_timestep, minimum_group_dt,
find_rung_ranges are checked against the reference's own sph/ts_rungs.hpp compiled for host memory
(oracle/ref_bdt.cpp).  What stays sequenced by this module alone: the order in which HydroVeBdtProp calls those pieces
(ve_hydro_bdt.hpp:280-420: which rungs drift, which advance, when groups are re-extracted).

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


# ---------------------------------------------------------------------------------------------------------------------
# HydroVeBdtProp (main/src/propagator/ve_hydro_bdt.hpp:51-420), one rank, g = 0
# ---------------------------------------------------------------------------------------------------------------------
MAX_RUNGS = 4  # Timestep::maxNumRungs (sph/include/sph/timestep.h:44)
FLT_MAX = np.finfo(np.float32).max


class Timestep:
    """sph::Timestep (sph/include/sph/timestep.h:42-66); float members are np.float32"""

    def __init__(self):
        self.nextDt = np.float32(0)
        self.elapsedDt = np.float32(0)
        self.totDt = np.float32(0)
        self.numRungs = 1
        self.substep = 0
        self.rungRanges = [0] * (MAX_RUNGS + 1)
        self.dt_m1 = np.zeros(MAX_RUNGS, np.float32)
        self.dt_drift = np.zeros(MAX_RUNGS, np.float32)

    def copy(self):
        t = Timestep()
        t.__dict__.update(self.__dict__)
        t.rungRanges = list(self.rungRanges)
        t.dt_m1, t.dt_drift = self.dt_m1.copy(), self.dt_drift.copy()
        return t


def sort_group_dt(group_dt, num):
    """sortGroupDt (ts_rungs.hpp:55-66): group_dt[:num] ascending in place (stable), returns the ordering"""
    order = np.argsort(group_dt[:num], kind="stable").astype(np.uint32)
    group_dt[:num] = group_dt[:num][order]
    return order


def rung_timestep(group_dt, num_groups, max_dt):
    """rungTimestep (ts_rungs.hpp:131-145) with computeMinTimestep (:78-101, fastFraction 0.4) on one rank; sorts
    group_dt[:num_groups] in place; returns (ordering, Timestep)"""
    indices = sort_group_dt(group_dt, num_groups)
    gd = group_dt
    m0, m1 = gd[0], gd[int(np.float32(0.4) * np.float32(num_groups))]
    num_rungs = min(int(np.log2(np.float32(m1 / m0))) + 1, MAX_RUNGS)
    rr = find_rung_ranges(m0, gd[:num_groups], num_rungs)
    m0 = min(np.float32(max_dt), m0)
    ts = Timestep()
    ts.nextDt, ts.totDt = np.float32(m0), np.float32(m0 * np.float32(1 << num_rungs))
    ts.numRungs, ts.rungRanges = num_rungs, rr
    return indices, ts


def minimum_group_dt(ts, group_dt, num_groups, num_groups_tot):
    """minimumGroupDt (ts_rungs.hpp:147-157): sorts group_dt[:num_groups] in place, the ordering is continued by a
    sequence up to num_groups_tot (computeMinTimestep :97); returns (ordering, dt, rungRanges)"""
    indices = np.concatenate([sort_group_dt(group_dt, num_groups),
                              np.arange(num_groups, num_groups_tot, dtype=np.uint32)])
    m0 = group_dt[0]
    rr = find_rung_ranges(m0, group_dt[:num_groups], MAX_RUNGS)
    time_left = np.float32(ts.totDt - ts.elapsedDt)
    substeps_left = (1 << ts.numRungs) - ts.substep
    dt = min(m0, np.float32(time_left / np.float32(substeps_left)))
    return indices, dt, rr


def find_rung_ranges(min_dt, group_dt, num_rungs):
    """findRungRanges<false> (sph/include/sph/ts_rungs.hpp:115-129); group_dt sorted float32"""
    n = len(group_dt)
    rr = [0] + [n] * MAX_RUNGS
    for rung in range(1, num_rungs):
        max_dt_rung = np.float32(1 << rung) * np.float32(min_dt)
        rr[rung] = int(np.searchsorted(group_dt, max_dt_rung, side="left"))
    return rr


def brute_force_neighbors(x, y, z, h, targets, box, ng0, ngmax, iterate_h):
    """all j != i with |r_ij|^2 < (2 h_i)^2 (minimum image on periodic axes, findneighbors.hpp:95-165), ascending j, with
    the coupled h / neighbor-count iteration of sph/find_neighbors.hpp:24-35: the result of the reference search when
    its tree covers every neighbor (searchExtFactor large enough).  Returns {i: list}, nc (count + 1) and updates h."""
    lim, bc = box
    n = x.size
    ngmin = ng0 // 4
    lists, nc = {}, {}
    todo = np.asarray(targets, np.int64)
    it = np.zeros(n, np.int64)
    while todo.size:
        nxt = []
        for c0 in range(0, todo.size, 256):
            t = todo[c0:c0 + 256]
            d2 = 0
            for a, dd in zip((x, y, z), range(3)):
                dx = a[t][:, None] - a[None, :]
                if bc[dd] == 1:
                    L = lim[2 * dd + 1] - lim[2 * dd]
                    dx = dx - L * np.rint(dx * (1.0 / L))
                d2 = d2 + dx * dx
            hi = h[t]
            inside = d2 < (hi.dtype.type(4.0) * hi * hi).astype(np.float64)[:, None]
            inside[np.arange(t.size), t] = False
            for k, i in enumerate(t):
                js = np.nonzero(inside[k])[0]
                nc_sph = js.size + 1
                if iterate_h and (ngmin > nc_sph or (nc_sph - 1) > ngmax) and it[i] < 10:
                    it[i] += 1
                    T = h.dtype.type
                    # updateH (sph/include/sph/kernels.hpp:26-32)
                    h[i] = h[i] * T(0.5) * np.power(T(1) + T(1023.0) * T(ng0) / T(nc_sph), T(1.0 / 10.0))
                    nxt.append(i)
                else:
                    lists[int(i)] = js[:ngmax].astype(np.uint32)
                    nc[int(i)] = nc_sph
        todo = np.asarray(nxt, np.int64)
    return lists, nc


class RefBdtSim:
    """HydroVeBdtProp sequenced over the reference's CPU templates (oracle/_ref): fullSync / partialSync /
    computeForces / computeRungs / integrate as in ve_hydro_bdt.hpp:172-376.  The reference runs this propagator on
    GPUs only; here its GPU operators are replaced by their CPU twins applied to the active groups' particles, the
    group kernels by the restatements above, the search during partial syncs by the exact all-pairs search."""

    def __init__(self, fields, box, settings, safety_steps=0, eta_acc=0.2, eps=0.005, **kw):
        from .ref_pipeline import RefSim
        self.rs = RefSim(fields, box, settings, **kw)
        rs = self.rs
        self.n = rs.n
        self.rung = np.zeros(self.n, np.uint8)
        self.ts = Timestep()
        self.ts.dt_m1[:] = np.float32(settings["minDt"])
        self.prev = self.ts.copy()
        self.safety_steps = safety_steps
        self.eta_acc_eps = np.float32(eta_acc * np.sqrt(eps))
        self.ext = np.float32(1)
        self.nb = np.zeros((self.n, rs.ngmax), np.uint32)
        self.nc = np.zeros(self.n, np.uint32)
        self.group_dt = None
        self.turb = None  # a ref.RefTurbulence makes this TurbVeBdtProp

    # ---- sync
    def full_sync(self):
        rs = self.rs
        rs.sync()
        self.rung = np.ascontiguousarray(self.rung[rs.perm])
        self.ext = np.float32(1)
        f = rs.f
        delim = compute_groups(0, self.n, f["x"], f["y"], f["z"], rs.ftree[0], rs.layout, rs.box[0])
        self.groups = (delim[:-1].copy(), delim[1:].copy())
        self.active = self.groups
        self.group_dt = np.full(self.groups[0].size, FLT_MAX, np.float32)
        self.all_active = True

    def partial_sync(self):
        self.ext = np.float32(self.ext * np.float32(1.012))
        highest = butterfly(self.ts.substep)
        lo, hi = self.ts.rungRanges[0], self.ts.rungRanges[highest]
        self.active = (self.ts_groups[0][lo:hi], self.ts_groups[1][lo:hi])
        self.all_active = False

    def _runs(self):
        """the active particles as maximal contiguous index ranges"""
        s, e = self.active
        if s.size == 0:
            return []
        order = np.argsort(s, kind="stable")
        s, e = s[order], e[order]
        runs, cs, ce = [], int(s[0]), int(e[0])
        for a, b in zip(s[1:], e[1:]):
            if int(a) == ce:
                ce = int(b)
            else:
                runs.append((cs, ce))
                cs, ce = int(a), int(b)
        runs.append((cs, ce))
        return runs

    def compute_forces(self):
        from . import ref
        rs = self.rs
        if active_rung(self.ts.substep, self.ts.numRungs) == 0:
            self.full_sync()
        else:
            self.partial_sync()
        f, n, K, box, ngmax, wh, whd = rs.f, rs.n, rs.K, rs.box, rs.ngmax, rs.wh, rs.whd
        nb, nc = self.nb, self.nc
        if self.all_active:
            nbl, ncl, f["h"] = ref.find_neighbors(f["x"], f["y"], f["z"], f["h"], 0, n, box, rs.tree, rs.layout,
                                                  rs.centers, rs.sizes, rs.ng0, ngmax, True)
            nb[:] = nbl.reshape(n, ngmax)
            nc[:] = ncl
        else:
            targets = np.concatenate([np.arange(a, b) for a, b in self._runs()] or [np.zeros(0, np.int64)])
            lists, ncd = brute_force_neighbors(f["x"], f["y"], f["z"], f["h"], targets, box, rs.ng0, ngmax, True)
            for i, js in lists.items():
                nb[i, :js.size] = js
                nc[i] = ncd[i]
        runs = self._runs()
        for a, b in runs:
            ref.xmass(a, b, ngmax, K, box, nb[a:b], nc[a:b], f["x"], f["y"], f["z"], f["h"], f["m"], wh, whd, f["xm"])
        for a, b in runs:
            ref.ve_def_gradh(a, b, ngmax, K, box, nb[a:b], nc[a:b], f["x"], f["y"], f["z"], f["h"], f["m"], wh, whd,
                             f["xm"], f["kx"], f["gradh"])
        ref.eos(0, n, f["temp"], f["m"], f["kx"], f["xm"], f["gradh"], rs.s["mui"], rs.s["gamma"], f["prho"], f["c"])
        dV = [f[k] for k in ("dV11", "dV12", "dV13", "dV22", "dV23", "dV33")] if rs.av_clean else None
        for a, b in runs:
            ref.iad_divv_curlv(a, b, ngmax, K, box, nb[a:b], nc[a:b], f["x"], f["y"], f["z"], f["vx"], f["vy"],
                               f["vz"], f["h"], wh, whd, f["xm"], f["kx"], f["c11"], f["c12"], f["c13"], f["c22"],
                               f["c23"], f["c33"], f["divv"], f["curlv"], dV)
        na = self.active[0].size
        gdt = self.group_dt
        gdt[:na] = group_divv_timestep(rs.s.get("Krho", 0.06), self.active[0], self.active[1], f["divv"], gdt[:na])
        for a, b in runs:
            ref.av_switches(a, b, ngmax, K, box, nb[a:b], nc[a:b], f["x"], f["y"], f["z"], f["vx"], f["vy"], f["vz"],
                            f["h"], f["c"], f["c11"], f["c12"], f["c13"], f["c22"], f["c23"], f["c33"], wh, whd,
                            f["kx"], f["xm"], f["divv"], rs.minDt, 0.05, 1.0, 0.2, f["alpha"])
        # one call per group: its return value is the group minimum of the Courant step (momentum_energy_gpu.cu:98-104)
        for g, (a, b) in enumerate(zip(self.active[0], self.active[1])):
            a, b = int(a), int(b)
            dtc = ref.momentum_energy(a, b, ngmax, K, box, nb[a:b], nc[a:b], f["x"], f["y"], f["z"], f["vx"], f["vy"],
                                      f["vz"], f["h"], f["m"], f["prho"], f["c"], f["c11"], f["c12"], f["c13"],
                                      f["c22"], f["c23"], f["c33"], 0.1, 0.2, 10.0, wh, f["kx"], f["xm"], f["alpha"],
                                      dV, f["ax"], f["ay"], f["az"], f["du"], rs.s.get("Kcour", 0.2), rs.av_clean)
            gdt[g] = min(gdt[g], np.float32(dtc))
        gdt[:na] = group_acc_timestep(self.eta_acc_eps, self.active[0], self.active[1], f["ax"], f["ay"], f["az"],
                                      gdt[:na])
        if self.turb is not None:
            # TurbVeBdtProp::computeForces (main/src/propagator/turb_ve.hpp:68-73): stir the active rungs afterwards
            self.turb.drive_groups(self.active[0], self.active[1], f["x"], f["y"], f["z"], f["ax"], f["ay"], f["az"],
                                   rs.minDt)

    # ---- time-step hierarchy
    def compute_rungs(self):
        ts = self.ts
        high = active_rung(ts.substep, ts.numRungs)
        num_tot = self.groups[0].size if high == 0 else ts.rungRanges[MAX_RUNGS]
        if high == 0:
            self.prev = ts.copy()
            max_dt = np.float32(ts.dt_m1[0] * np.float32(1.1))
            ng = self.groups[0].size
            indices, ts = rung_timestep(self.group_dt, ng, max_dt)
            if self.safety_steps > 0:
                ts.numRungs = 1
                ts.rungRanges[1:] = [ng] * MAX_RUNGS
                self.safety_steps -= 1
            self.ts = ts
        else:
            ng = ts.rungRanges[high]
            indices, ts.nextDt, rr = minimum_group_dt(ts, self.group_dt, ng, num_tot)
            ts.rungRanges[:high] = rr[:high]
        ts = self.ts
        if high == 0 or high > 1:
            src = self.groups if high == 0 else self.ts_groups
            if high > 1:
                self.groups = self.ts_groups
            nout = ts.rungRanges[MAX_RUNGS]
            self.ts_groups = (src[0][indices[:nout]], src[1][indices[:nout]])
        self.rungs = [(self.ts_groups[0][ts.rungRanges[r]:ts.rungRanges[r + 1]],
                       self.ts_groups[1][ts.rungRanges[r]:ts.rungRanges[r + 1]]) for r in range(ts.numRungs)]

    @staticmethod
    def _members(grp):
        if grp[0].size == 0:
            return np.zeros(0, np.uint32)
        return np.concatenate([np.arange(a, b, dtype=np.uint32) for a, b in zip(grp[0], grp[1])])

    def integrate(self):
        from . import ref
        self.compute_rungs()
        rs, ts, f = self.rs, self.ts, self.rs.f
        lowest_drift = butterfly(ts.substep + 1)
        is_last = active_rung(ts.substep + 1, ts.numRungs) == 0
        box = rs.box if is_last else ([0, 1, 0, 1, 0, 1], [0, 0, 0])
        args = (f["x"], f["y"], f["z"], f["vx"], f["vy"], f["vz"], f["x_m1"], f["y_m1"], f["z_m1"], f["ax"], f["ay"],
                f["az"], f["temp"], f["du"], f["du_m1"])
        mui, gamma = rs.s["mui"], rs.s["gamma"]
        for i in range(ts.numRungs):
            use_rung = ts.substep == ts.substep % (1 << i)
            advance = i < lowest_drift
            dt = ts.nextDt
            dt_m1 = (self.prev.dt_m1 if use_rung else ts.dt_m1).copy()
            idx = self._members(self.rungs[i])
            if advance:
                if ts.dt_drift[i] > 0:
                    ref.drift(idx, 0.0, ts.dt_drift[i], dt_m1, self.rung, *args, mui, gamma)
                ref.positions_rung(idx, np.float32(ts.dt_drift[i] + dt), dt_m1, self.rung, box, *args, f["h"], mui,
                                   gamma)
                ts.dt_m1[i] = np.float32(ts.dt_drift[i] + dt)
                ts.dt_drift[i] = 0
                self.rung[idx] = i
            else:
                ref.drift(idx, np.float32(ts.dt_drift[i] + dt), ts.dt_drift[i], dt_m1, self.rung, *args, mui, gamma)
                ts.dt_drift[i] = np.float32(ts.dt_drift[i] + dt)
        for a, b in self._runs():
            ref.update_h(a, b, rs.ng0, self.nc, f["h"])
        ts.substep += 1
        ts.elapsedDt = np.float32(ts.elapsedDt + ts.nextDt)
        rs.ttot += float(ts.nextDt)
        rs.minDt_m1 = rs.minDt
        rs.minDt = float(ts.nextDt)

    def step(self):
        self.compute_forces()
        self.integrate()
