"""Python mirror of the Propagator / Domain boundary over the tier-2 C ABI.

`HydroVeProp` follows sphexa::HydroVeProp (main/src/propagator/ve_hydro.hpp:54-219): sync(),
computeForces(), integrate(); `Domain`-style getters startIndex/endIndex/nParticlesWithHalos
follow cstone::Domain (domain/include/cstone/domain/domain.hpp:382-429).  All device state lives in
the C library (sphb200_sim); this class only marshals host buffers.
"""
import ctypes as C
import numpy as np

from . import binding as B
from . import ic
from .binding import lib, check, vp

PHASES = ["domain::sync", "FindNeighbors", "XMass", "Normalization & Gradh", "EquationOfState",
          "IadVelocityDivCurl", "AVswitches", "MomentumAndEnergy", "Timestep+UpdateQuantities"]


def _host_fields(arrays, real):
    """dict name -> numpy array => sphb200_fields of host pointers (keeps references alive in `keep`)"""
    f = B.Fields()
    keep = []
    for n in B.FIELD_NAMES:
        a = arrays.get(n)
        if a is None:
            setattr(f, n, None)
            continue
        want = np.float64 if n in B.ALWAYS_F64 else (np.uint32 if n == "nc" else real)
        a = np.ascontiguousarray(a, dtype=want)
        keep.append(a)
        setattr(f, n, a.ctypes.data)
    return f, keep


class HydroVeProp:
    """one rank of an SPH-VE run; data stays resident on the GPU between steps"""

    def __init__(self, ctx, box, settings, capacity, dtype=B.F64, rank=0, num_ranks=1, bucket_size=None,
                 bucket_size_focus=64, av_clean=False, tables=None, creator=None):
        self.ctx = ctx
        self.dtype = dtype
        self.real = np.float64 if dtype == B.F64 else np.float32
        self.settings = dict(settings)
        n_glob = settings.get("numParticlesGlobal", capacity)
        if bucket_size is None:
            # sphexa.cpp:133-135
            bucket_size = max(bucket_size_focus, n_glob // (100 * num_ranks))
        wh, whd, K = tables if tables is not None else ic.kernel_tables(self.real)
        self._tables = (np.ascontiguousarray(wh, self.real), np.ascontiguousarray(whd, self.real))
        cfg = B.SimConfig()
        cfg.dtype, cfg.rank, cfg.numRanks = dtype, rank, num_ranks
        cfg.bucketSize, cfg.bucketSizeFocus, cfg.capacity = int(bucket_size), int(bucket_size_focus), int(capacity)
        cfg.box = B.make_box(*box)
        hc = cfg.hc
        hc.K, hc.ng0, hc.ngmax = K, int(settings.get("ng0", 100)), int(settings.get("ngmax", 150))
        hc.Kcour, hc.Krho = settings.get("Kcour", 0.2), settings.get("Krho", 0.06)
        hc.gamma, hc.muiConst = settings["gamma"], settings["mui"]
        hc.alphamin, hc.alphamax, hc.decay_constant = 0.05, 1.0, 0.2
        at0, at1 = settings.get("Atmin", 0.1), settings.get("Atmax", 0.2)  # ParticlesData::Atmin / Atmax
        hc.Atmin, hc.Atmax, hc.ramp = at0, at1, 1.0 / (at1 - at0)
        hc.minDt, hc.minDt_m1 = settings["minDt"], settings["minDt_m1"]
        hc.wh, hc.whd = self._tables[0].ctypes.data, self._tables[1].ctypes.data
        hc.avClean = int(av_clean)
        self.K = K
        self.capacity = capacity
        self.h = vp()
        if creator is None:
            check(lib.sphb200_sim_create(ctx.h, C.byref(cfg), C.byref(self.h)))
        else:
            check(creator(ctx.h, C.byref(cfg), C.byref(self.h)))

    def close(self):
        if self.h:
            lib.sphb200_sim_destroy(self.h)
            self.h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- data movement across the boundary
    def upload(self, arrays):
        n = len(arrays["x"])
        f, keep = _host_fields(arrays, self.real)
        check(lib.sphb200_sim_upload(self.h, C.c_size_t(n), C.byref(f)))
        self.ctx.sync()
        del keep

    def upload_async(self, f_struct, n):
        check(lib.sphb200_sim_upload(self.h, C.c_size_t(n), C.byref(f_struct)))

    def download(self, names=None, with_keys=False):
        names = names or B.CONSERVED
        cnt = self.endIndex - self.startIndex
        out = {}
        for n in names:
            want = np.float64 if n in B.ALWAYS_F64 else (np.uint32 if n == "nc" else self.real)
            out[n] = np.empty(cnt, want)
        f, keep = _host_fields(out, self.real)
        # _host_fields may copy; make sure we hand out the arrays that were written
        for n, a in zip([n for n in B.FIELD_NAMES if n in out], keep):
            out[n] = a
        keys = np.empty(cnt, np.uint64) if with_keys else None
        check(lib.sphb200_sim_download(self.h, C.byref(f), keys.ctypes.data_as(vp) if with_keys else None))
        if with_keys:
            out["keys"] = keys
        return out

    def step_host(self, fin, fout, n):
        """one time step on host-resident state (transfers overlapped with the kernels); fin/fout: binding.Fields of
        pinned host pointers for the 14 conserved fields"""
        check(lib.sphb200_sim_step_host(self.h, C.c_size_t(n), C.byref(fin), C.byref(fout)))

    def attachTurbulence(self, turb):
        """TurbVeProp (main/src/propagator/turb_ve.hpp:96-118): stir after the hydro forces of every step"""
        check(lib.sphb200_sim_attach_turbulence(self.h, turb.h if turb is not None else None))
        self._turb = turb

    def conserved(self):
        """computeConservedQuantities over the owned particles (observables/conserved_quantities.hpp:113-180)"""
        from . import ops
        out = (C.c_double * 10)()
        check(lib.sphb200_sim_conserved(self.h, out))
        return ops.conserved_dict(list(out))

    def machRMS(self, num_particles_global=0):
        """TurbulenceMachRMS (main/src/observables/turbulence_mach_rms.hpp:70-83): sqrt(sum |v|^2/c^2 / N)"""
        out = C.c_double(0)
        check(lib.sphb200_sim_mach_rms(self.h, C.c_uint64(num_particles_global), C.byref(out)))
        return out.value

    def download_into(self, f_struct):
        check(lib.sphb200_sim_download(self.h, C.byref(f_struct), None))

    # ---- Propagator interface
    def sync(self):
        check(lib.sphb200_sim_sync(self.h))

    def computeForces(self):
        check(lib.sphb200_sim_sync(self.h))
        check(lib.sphb200_sim_compute_forces(self.h))

    def forcesPhase(self, k):
        check(lib.sphb200_sim_forces_phase(self.h, k))

    def integrate(self, global_min_dt=-1.0):
        check(lib.sphb200_sim_integrate(self.h, C.c_double(global_min_dt)))

    def step(self):
        check(lib.sphb200_sim_step(self.h))

    def localTimestep(self):
        out = (C.c_double * 3)()
        check(lib.sphb200_sim_local_timestep(self.h, out))
        return tuple(out)

    # ---- Domain getters
    @property
    def startIndex(self):
        return lib.sphb200_sim_start_index(self.h)

    @property
    def endIndex(self):
        return lib.sphb200_sim_end_index(self.h)

    @property
    def nParticlesWithHalos(self):
        return lib.sphb200_sim_n_with_halos(self.h)

    @property
    def numLeaves(self):
        return lib.sphb200_sim_num_leaves(self.h)

    @property
    def numGlobalLeaves(self):
        return lib.sphb200_sim_num_global_leaves(self.h)

    @property
    def minDt(self):
        return lib.sphb200_sim_min_dt(self.h)

    @property
    def ttot(self):
        return lib.sphb200_sim_ttot(self.h)

    def boxLimits(self):
        """Domain::box(): [xmin, xmax, ymin, ymax, zmin, zmax] (open axes follow the particles)"""
        b = B.Box()
        lib.sphb200_sim_get_box(self.h, C.byref(b))
        return list(b.lim)

    def totalNeighbors(self):
        return lib.sphb200_sim_total_neighbors(self.h)

    def phaseTimes(self):
        ms = (C.c_float * 9)()
        check(lib.sphb200_sim_phase_times(self.h, ms))
        return dict(zip(PHASES, list(ms)))

    def deviceFields(self):
        f = B.Fields()
        check(lib.sphb200_sim_fields(self.h, C.byref(f)))
        return f

    def treeView(self):
        t = B.TreeView()
        check(lib.sphb200_sim_tree(self.h, C.byref(t)))
        return t

    def _d2h(self, dev_ptr, count, dtype):
        """copy `count` elements from a raw device pointer (tests / diagnostics)"""
        import torch
        out = np.empty(count, dtype)
        if count == 0:
            return out
        nbytes = out.nbytes
        rt = C.CDLL("libcudart.so.12") if not hasattr(HydroVeProp, "_rt") else HydroVeProp._rt
        HydroVeProp._rt = rt
        self.ctx.sync()
        rc = rt.cudaMemcpy(out.ctypes.data_as(vp), vp(dev_ptr), C.c_size_t(nbytes), 2)
        if rc != 0:
            raise B.SphError(f"cudaMemcpy D2H failed: {rc}")
        return out

    def leaves(self):
        t = self.treeView()
        return self._d2h(t.leaves, t.numLeafNodes + 1, np.uint64)

    def leafCounts(self):
        return self._d2h(lib.sphb200_sim_leaf_counts(self.h), self.numLeaves, np.uint32)

    def globalLeaves(self):
        return self._d2h(lib.sphb200_sim_global_leaves(self.h), self.numGlobalLeaves + 1, np.uint64)

    def globalCounts(self):
        return self._d2h(lib.sphb200_sim_global_counts(self.h), self.numGlobalLeaves, np.uint32)


class HydroVeBdtProp(HydroVeProp):
    """sphexa::HydroVeBdtProp (main/src/propagator/ve_hydro_bdt.hpp:51-420): SPH-VE with block time steps.

    Groups of <= 32 SFC-consecutive particles get their own time step (Courant, divv and acceleration criteria), the
    steps are binned into <= 4 power-of-two rungs, and between two full syncs only the groups of the rungs that are
    due are re-evaluated (sync() = fullSync or partialSync); the others are drifted."""

    def __init__(self, *args, safety_steps=0, eta_acc=0.0, eps=0.0, **kw):
        super().__init__(*args, **kw)
        check(lib.sphb200_sim_bdt_enable(self.h, int(safety_steps), C.c_double(eta_acc), C.c_double(eps)))

    def computeForces(self):
        check(lib.sphb200_sim_bdt_compute_forces(self.h))

    def integrate(self, global_min_dt=-1.0):
        check(lib.sphb200_sim_bdt_integrate(self.h))

    def step(self):
        check(lib.sphb200_sim_bdt_step(self.h))

    def timestep(self):
        """sph::Timestep of the running hierarchy + group counts (binding.BdtInfo)"""
        info = B.BdtInfo()
        check(lib.sphb200_sim_bdt_info(self.h, C.byref(info)))
        return info

    def isSynced(self):
        """Propagator::isSynced (ve_hydro_bdt.hpp:220): the next computeForces starts with a full sync"""
        return bool(self.timestep().isSynced)

    def rungs(self):
        """per-particle rung of the owned particles"""
        p = lib.sphb200_sim_bdt_rungs(self.h)
        return self._d2h(p + self.startIndex, self.endIndex - self.startIndex, np.uint8)

    def uploadRungs(self, rung):
        """rung of the particles of the last upload() (restart: the conserved field "rung", ve_hydro_bdt.hpp:94)"""
        rung = np.ascontiguousarray(rung, np.uint8)
        check(lib.sphb200_sim_bdt_upload_rungs(self.h, rung.ctypes.data_as(C.c_void_p), C.c_size_t(rung.size)))

    def groupDt(self):
        return self._d2h(lib.sphb200_sim_bdt_group_dt(self.h), self.timestep().numGroups, np.float32)

    def tsGroups(self):
        """groups of the last full sync in time-step order: (start, end) index arrays"""
        v = B.GroupView()
        check(lib.sphb200_sim_bdt_groups(self.h, C.byref(v)))
        return self._d2h(v.groupStart, v.numGroups, np.uint32), self._d2h(v.groupEnd, v.numGroups, np.uint32)
