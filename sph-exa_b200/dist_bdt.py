"""HydroVeBdtProp on P ranks (main/src/propagator/ve_hydro_bdt.hpp:51-420): block time steps over the SFC-partitioned
domain of dist.DistHydroVeProp.

What the ranks share (ts_rungs.hpp:77-93): the minimum group time step and the time step of the fastest 40 % of the
groups, both MIN-all-reduced; everything else (group tables, the sort of the group time steps, rung ranges, drift and
advance) is local to a rank.  Between two full syncs a rank only exchanges halos (partialSync: x, y, z, h, then the usual
four exchanges inside computeForces); the halo search radius is widened by setHaloFactor(1 + numRungs / 40) at the full
sync so that the halo set stays valid while particles drift.

The per-rank sequencing is the one of sphb200_sim_bdt_* (csrc/api.cu), written against the tier-1 operators.
"""
import ctypes as C

import numpy as np

from .binding import RestartBdt
from .dist import DistHydroVeProp

MAX_RUNGS = 4  # Timestep::maxNumRungs (sph/include/sph/timestep.h:44)
FLT_MAX = float(np.finfo(np.float32).max)


def butterfly(i):
    """cstone::butterfly (domain/include/cstone/primitives/math.hpp:26-31)"""
    return 0 if i == 0 else (i & -i).bit_length()


def active_rung(substep, num_rungs):
    """HydroVeBdtProp::activeRung (ve_hydro_bdt.hpp:106-110)"""
    if substep == 0 or substep >= (1 << (num_rungs - 1)):
        return 0
    return butterfly(substep)


class Timestep:
    """sph::Timestep (sph/include/sph/timestep.h:42-66); the float members are kept as np.float32"""

    def __init__(self):
        f32 = np.float32
        self.nextDt, self.elapsedDt, self.totDt = f32(0), f32(0), f32(0)
        self.numRungs, self.substep = 1, 0
        self.rungRanges = [0] * (MAX_RUNGS + 1)
        self.dt_m1 = np.zeros(MAX_RUNGS, np.float32)
        self.dt_drift = np.zeros(MAX_RUNGS, np.float32)

    def copy(self):
        t = Timestep()
        t.__dict__.update(self.__dict__)
        t.rungRanges = list(self.rungRanges)
        t.dt_m1, t.dt_drift = self.dt_m1.copy(), self.dt_drift.copy()
        return t


class DistHydroVeBdtProp(DistHydroVeProp):
    # "rung" is a conserved field of this propagator (ve_hydro_bdt.hpp:90): it follows the particles through the
    # reorder, the migration and the halo layout.  Kept as int32 here (the field gather moves 4- and 8-byte elements)
    # and handed to the group kernels as bytes.
    CONSERVED = DistHydroVeProp.CONSERVED + ["rung"]

    def __init__(self, *args, safety_steps=0, eta_acc=0.2, eps=0.005, **kw):
        super().__init__(*args, **kw)
        self.ts = Timestep()
        self.ts.dt_m1[:] = np.float32(self.minDt)
        self.prev = self.ts.copy()
        self.safety_steps = safety_steps
        self.eta_acc_eps = float(np.float32(eta_acc * np.sqrt(eps)))
        self.search_ext = 1.0
        self.all_active = True
        self.groups = self.ts_groups = None
        self.group_dt = self.group_indices = None
        self._halo_layout = False

    # ---- data in / out
    def upload(self, arrays):
        super_names = DistHydroVeProp.CONSERVED
        t = self.torch
        for k in super_names:
            want = np.float64 if k in self.B.ALWAYS_F64 else (np.float64 if self.dtype == self.B.F64 else np.float32)
            self.f[k] = t.from_numpy(np.ascontiguousarray(arrays[k], dtype=want)).to(self.dev)
        rung = arrays.get("rung")
        n = self.f["x"].numel()
        self.f["rung"] = (t.zeros(n, dtype=t.int32, device=self.dev) if rung is None
                          else t.from_numpy(np.ascontiguousarray(rung, np.int32)).to(self.dev))

    # ---- HydroVeBdtProp::save / load (ve_hydro_bdt.hpp:153-168): Timestep::loadOrStore with the prefix "ts::"
    def _restart_bdt(self, own):
        if not self.isSynced():
            # like sphb200_sim_save: only numRungs, dt_m1 and the rungs travel, so a dump from the middle of a
            # hierarchy would resume into a different run (the reference writes output only when isSynced())
            raise self.B.SphError("restart: the block-time-step hierarchy is not synced; dumps are only taken where "
                                  "isSynced() holds")
        rung = np.ascontiguousarray(own["rung"], np.uint8)
        self._keep_rung = rung  # the struct holds a raw pointer
        return RestartBdt(int(self.ts.numRungs), (C.c_float * MAX_RUNGS)(*self.ts.dt_m1), rung.ctypes.data)

    def _restart_extra(self, files, pieces, arrays):
        count = len(arrays["x"])
        if not all(f.has_field("rung") for f in files):
            return  # a dump of the global-time-step propagator: rung 0 everywhere, time-step record as constructed
        rung = np.zeros(count, np.uint8)
        for f, first, cnt, offset in pieces:
            f.read_bytes("rung", first, cnt, rung, offset=offset)
        arrays["rung"] = rung
        self.ts = Timestep()
        self.ts.numRungs = int(files[0].attribute("ts::numRungs", np.int64)[0])
        self.ts.dt_m1[:] = files[0].attribute("ts::dt_m1", np.float32)
        self.prev = self.ts.copy()
        self._halo_layout = False

    def isSynced(self):
        """Propagator::isSynced (ve_hydro_bdt.hpp:220)"""
        return active_rung(self.ts.substep, self.ts.numRungs) == 0

    def _zero_dependent(self):
        t, f, n = self.torch, self.f, self.n_total
        for k in self.DEPENDENT:
            f[k] = t.zeros(n, dtype=self.real_t, device=self.dev)
        f["du"] = t.zeros(n, dtype=t.float64, device=self.dev)
        f["nc"] = t.zeros(n, dtype=t.int32, device=self.dev)

    # ---- HydroVeBdtProp::fullSync / partialSync (ve_hydro_bdt.hpp:172-218)
    def _full_sync(self):
        t, ops, f = self.torch, self.ops, self.f
        self.halo_factor = 1.0 + float(self.ts.numRungs) / 40
        if self._halo_layout:
            # arrays still carry the halos of the finished hierarchy: the owned range is the state that is synced
            for k in self.CONSERVED:
                f[k] = f[k][self.start:self.end].contiguous()
        self.sync(lazy=False)
        self._halo_layout = True
        self.search_ext = 1.0
        a, b = self.start, self.end
        # the search tree: its leaves / layout index the [halos | owned | halos] arrays
        self.groups = ops.computeGroups(self.ctx, a, b, f["x"], f["y"], f["z"], self.tree, self.box)
        self.active_view = self.groups.view()
        self.all_active = True
        ng = self.groups.numGroups
        self.group_dt = t.full((max(ng, 1),), FLT_MAX, dtype=t.float32, device=self.dev)
        self._zero_dependent()
        self.dt_particle = t.empty(self.n_total, dtype=self.real_t, device=self.dev)

    def _partial_sync(self):
        self.exchange_halos(["x", "y", "z", "h"])
        self.search_ext *= 1.012
        highest = butterfly(self.ts.substep)
        lo, hi = self.ts.rungRanges[0], self.ts.rungRanges[highest]
        self.active_view = self.ts_groups.view(lo, hi)
        self.active_count = hi - lo
        self.all_active = False

    # ---- HydroVeBdtProp::computeForces (ve_hydro_bdt.hpp:222-283)
    def computeForces(self):
        t, ops, f, B = self.torch, self.ops, self.f, self.B
        if self.isSynced():
            self._full_sync()
            self.active_count = self.groups.numGroups
        else:
            self._partial_sync()
        n, a, b = self.n_total, self.start, self.end
        hc = self._hc()
        mask = gmask = None
        if not self.all_active:
            mask = ops.markGroups(self.ctx, self.active_view, n)
            gmask = t.zeros(n, dtype=t.uint8, device=self.dev)
        gdt = self.group_dt
        with ops.scope(self.ctx, active=mask, searchExtFactor=self.search_ext, dtParticle=self.dt_particle,
                       gatherMask=gmask):
            if self.all_active:
                self.nb, nc, self.fails = ops.findNeighbors(self.ctx, self.tree, f["x"], f["y"], f["z"], f["h"], a, b,
                                                            self.box, self.ng0, self.ngmax, True)
                f["nc"][a:b] = nc
            else:
                tv = self.tree.view()
                B.check(B.lib.sphb200_find_neighbors(self.ctx.h, C.byref(tv), B.ptr(f["x"]), B.ptr(f["y"]),
                                                     B.ptr(f["z"]), B.ptr(f["h"]), self.dtype, C.c_uint32(a),
                                                     C.c_uint32(b), C.byref(self.box), C.c_uint(self.ng0),
                                                     C.c_uint(self.ngmax), 1, B.ptr(self.nb), B.ptr(f["nc"][a:]), None))
            ops.computeXMass(self.ctx, a, b, f, self.box, hc, self.nb)
            self.exchange_halos(["xm"])
            ops.computeVeDefGradh(self.ctx, a, b, f, self.box, hc, self.nb)
            ops.computeEOS(self.ctx, a, b, f, hc)
            self.exchange_halos(["vx", "vy", "vz", "prho", "c", "kx"])
            ops.computeIadDivvCurlv(self.ctx, a, b, f, self.box, hc, self.nb)
            ops.groupDivvTimestep(self.ctx, float(self.s.get("Krho", 0.06)), self.active_view, f["divv"], gdt)
            self.exchange_halos(["c11", "c12", "c13", "c22", "c23", "c33", "divv"])
            ops.computeAVswitches(self.ctx, a, b, f, self.box, hc, self.nb)
            self.exchange_halos(["alpha"])
            f_struct = ops.fields_struct(f)
            B.check(B.lib.sphb200_momentum_energy(self.ctx.h, self.dtype, C.c_uint32(a), C.c_uint32(b),
                                                  C.byref(self.box), C.byref(hc), B.ptr(self.nb), C.byref(f_struct),
                                                  None))
            ops.groupMinTimestep(self.ctx, self.active_view, self.dt_particle, gdt)
            ops.groupAccTimestep(self.ctx, self.eta_acc_eps, self.active_view, f["ax"], f["ay"], f["az"], gdt)
            if self.turb is not None:
                # TurbVeBdtProp (turb_ve.hpp:68-73): the active rungs are stirred after the whole of computeForces, i.e.
                # after the acceleration time step was taken; the kernel honours the scope's active mask, every rank
                # advances the same OU sequence once per substep
                self.turb.drive(self.ctx, a, b, f, self.minDt)

    # ---- ts_rungs.hpp:77-157 with the two all-reduces
    def _global_min(self, values):
        t = self.torch
        v = t.tensor(values, dtype=t.float32, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(v, op=self.dist.ReduceOp.MIN, group=self.group)
        return [np.float32(x) for x in v.tolist()]

    def _compute_rungs(self):
        t, ops, ts = self.torch, self.ops, self.ts
        high = active_rung(ts.substep, ts.numRungs)
        if high == 0:
            self.prev = ts.copy()
            max_dt = np.float32(ts.dt_m1[0] * np.float32(1.1))
            ng = self.groups.numGroups
            # rungTimestep (ts_rungs.hpp:131-145)
            self.group_indices = ops.sortGroupDt(self.ctx, self.group_dt, ng)
            _, local = ops.rungRanges(self.ctx, -1.0, self.group_dt, ng, 1)
            m0, m1 = self._global_min([local[0], local[1]])
            num_rungs = min(int(np.log2(np.float32(m1 / m0))) + 1, MAX_RUNGS)
            rr, _ = ops.rungRanges(self.ctx, float(m0), self.group_dt, ng, num_rungs)
            m0 = min(max_dt, m0)
            ts = Timestep()
            ts.nextDt, ts.totDt = np.float32(m0), np.float32(m0 * np.float32(1 << num_rungs))
            ts.numRungs, ts.rungRanges = num_rungs, rr
            if self.safety_steps > 0:
                ts.numRungs = 1
                ts.rungRanges[1:] = [ng] * MAX_RUNGS
                self.safety_steps -= 1
            self.ts = ts
        else:
            # minimumGroupDt (ts_rungs.hpp:147-157)
            ng, tot = ts.rungRanges[high], ts.rungRanges[MAX_RUNGS]
            head = ops.sortGroupDt(self.ctx, self.group_dt, ng)
            tail = t.arange(ng, tot, dtype=t.int32, device=self.dev)
            self.group_indices = t.cat([head, tail])
            rr, local = ops.rungRanges(self.ctx, -1.0, self.group_dt, ng, 1)
            (m0,) = self._global_min([local[0]])
            rr, _ = ops.rungRanges(self.ctx, float(m0), self.group_dt, ng, MAX_RUNGS)
            time_left = np.float32(ts.totDt - ts.elapsedDt)
            substeps_left = (1 << ts.numRungs) - ts.substep
            ts.nextDt = min(m0, np.float32(time_left / np.float32(substeps_left)))
            ts.rungRanges[:high] = rr[:high]
        ts = self.ts
        if high == 0 or high > 1:
            src = self.groups if high == 0 else self.ts_groups
            if high > 1:
                self.groups = self.ts_groups
            self.ts_groups = ops.extractGroups(self.ctx, src, self.group_indices, 0, ts.rungRanges[MAX_RUNGS])
        self.rungs = [self.ts_groups.view(ts.rungRanges[r], ts.rungRanges[r + 1]) for r in range(ts.numRungs)]

    # ---- HydroVeBdtProp::integrate (ve_hydro_bdt.hpp:329-376)
    def integrate(self):
        t, ops, f, ts0 = self.torch, self.ops, self.f, self.ts
        # the active view of this substep must survive the regrouping (it points into the old tables)
        active_view, keep = self.active_view, (self.groups, self.ts_groups)
        self._compute_rungs()
        ts = self.ts
        hc = self._hc()
        fs = ops.fields_struct(f)
        rung8 = f["rung"].to(t.uint8)
        lowest_drift = butterfly(ts.substep + 1)
        is_last = active_rung(ts.substep + 1, ts.numRungs) == 0
        open_box = self.pkg.make_box([0, 1, 0, 1, 0, 1], [0, 0, 0])
        box = self.box if is_last else open_box
        for i in range(ts.numRungs):
            use_rung = ts.substep == ts.substep % (1 << i)
            advance = i < lowest_drift
            dt = ts.nextDt
            dt_m1 = (self.prev.dt_m1 if use_rung else ts.dt_m1).copy()
            if advance:
                if ts.dt_drift[i] > 0:
                    ops.driftPositions(self.ctx, self.dtype, self.rungs[i], 0.0, float(ts.dt_drift[i]), dt_m1, rung8, hc, fs)
                ops.computePositionsGroups(self.ctx, self.dtype, self.rungs[i], float(np.float32(ts.dt_drift[i] + dt)),
                                           dt_m1, rung8, box, hc, fs)
                ts.dt_m1[i] = np.float32(ts.dt_drift[i] + dt)
                ts.dt_drift[i] = 0
                ops.storeRung(self.ctx, self.rungs[i], i, rung8)
            else:
                ops.driftPositions(self.ctx, self.dtype, self.rungs[i], float(np.float32(ts.dt_drift[i] + dt)),
                                   float(ts.dt_drift[i]), dt_m1, rung8, hc, fs)
                ts.dt_drift[i] = np.float32(ts.dt_drift[i] + dt)
        f["rung"] = rung8.to(t.int32)
        ops.updateSmoothingLengthGroups(self.ctx, active_view, self.ng0, f["nc"], f["h"])
        del keep
        ts.substep += 1
        ts.elapsedDt = np.float32(ts.elapsedDt + ts.nextDt)
        self.ttot += float(ts.nextDt)
        self.minDt_m1 = self.minDt
        self.minDt = float(ts.nextDt)

    def step(self):
        self.computeForces()
        self.integrate()

    def download_owned(self, names=None):
        names = names or self.CONSERVED
        out = {}
        n_own = self.end - self.start
        for k in names:
            a = self.f[k]
            if a.numel() != n_own:
                a = a[self.start:self.end]
            out[k] = a.cpu().numpy()
        out["keys"] = self.own_keys.cpu().numpy().view(np.uint64)
        return out
