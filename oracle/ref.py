"""TEST INFRASTRUCTURE ONLY: ctypes loader for oracle/_ref/libsphexa_ref.so.

That library is the reference's own CPU code (cstone + sph header templates from
/root/reference, compiled by oracle/Makefile behind oracle/ref_wrap.cpp).  It is
the checker, never the product: only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs import this module.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "_ref", "libsphexa_ref.so")
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(_PATH)
        for name in ("ref_momentum_energy_f64", "ref_momentum_energy_f32", "ref_rho_timestep_f64",
                     "ref_rho_timestep_f32", "ref_sphynx_3D_k"):
            getattr(_lib, name).restype = C.c_double
    return _lib


def P(a):
    """pointer to a contiguous numpy array (or NULL)"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)


def _box(box):
    lim = np.ascontiguousarray(box[0], dtype=np.float64)
    bc = np.ascontiguousarray(box[1], dtype=np.int32)
    return lim, bc


class TreeArgs(C.Structure):
    _fields_ = [("numLeaves", C.c_int), ("prefixes", C.c_void_p), ("childOffsets", C.c_void_p),
                ("internalToLeaf", C.c_void_p), ("levelRange", C.c_void_p), ("leaves", C.c_void_p),
                ("layout", C.c_void_p), ("centers", C.c_void_p), ("sizes", C.c_void_p)]


def num_threads():
    return lib().ref_num_threads()


def set_num_threads(n):
    lib().ref_set_num_threads(int(n))


def hilbert_encode(ix, iy, iz):
    ix, iy, iz = (np.ascontiguousarray(a, dtype=np.uint32) for a in (ix, iy, iz))
    keys = np.empty(ix.size, np.uint64)
    lib().ref_hilbert_encode(P(ix), P(iy), P(iz), C.c_size_t(ix.size), P(keys))
    return keys


def hilbert_decode(keys):
    keys = np.ascontiguousarray(keys, dtype=np.uint64)
    ix, iy, iz = (np.empty(keys.size, np.uint32) for _ in range(3))
    lib().ref_hilbert_decode(P(keys), C.c_size_t(keys.size), P(ix), P(iy), P(iz))
    return ix, iy, iz


def compute_sfc_keys(x, y, z, box):
    lim, bc = _box(box)
    keys = np.empty(x.size, np.uint64)
    lib().ref_compute_sfc_keys(P(x), P(y), P(z), C.c_size_t(x.size), P(lim), P(bc), P(keys))
    return keys


def assemble_cuboid(key_start, key_end, box, multiplicity, xb, yb, zb):
    """sphexa::assembleCuboid (main/src/init/grid.hpp:238-264): x, y, z in the reference's output order"""
    lim, bc = _box(box)
    m = np.ascontiguousarray(multiplicity, dtype=np.int32)
    xb, yb, zb = (np.ascontiguousarray(a, dtype=np.float64) for a in (xb, yb, zb))
    f = lib().ref_assemble_cuboid
    f.restype = C.c_size_t
    cap = int(xb.size) * int(m[0]) * int(m[1]) * int(m[2]) + 16
    x, y, z = (np.empty(cap, np.float64) for _ in range(3))
    n = f(C.c_uint64(int(key_start)), C.c_uint64(int(key_end)), P(lim), P(bc), P(m), P(xb), P(yb), P(zb),
          C.c_size_t(xb.size), P(x), P(y), P(z), C.c_size_t(cap))
    assert n <= cap
    return x[:n].copy(), y[:n].copy(), z[:n].copy()


def equi_distant_sfc_segment(rank, num_ranks, leaves_per_rank=100):
    """sphexa::equiDistantSfcSegments (main/src/init/grid.hpp:80-88)"""
    out = np.zeros(2, np.uint64)
    lib().ref_equidistant_sfc_segment(C.c_int(rank), C.c_int(num_ranks), C.c_int(leaves_per_rank), P(out))
    return int(out[0]), int(out[1])


def span_sfc_range(a, b):
    """cstone::spanSfcRange (domain/include/cstone/sfc/common.hpp:362-400): cells + the end key"""
    out = np.zeros(200, np.uint64)
    n = lib().ref_span_sfc_range(C.c_uint64(int(a)), C.c_uint64(int(b)), P(out))
    return out[:n + 1].copy()


def cut_sphere(radius, x, y, z):
    """sphexa::cutSphere (grid.hpp:267-300)"""
    x, y, z = (np.array(a, dtype=np.float64, copy=True) for a in (x, y, z))
    f = lib().ref_cut_sphere
    f.restype = C.c_size_t
    n = f(C.c_double(radius), P(x), P(y), P(z), C.c_size_t(x.size))
    return x[:n].copy(), y[:n].copy(), z[:n].copy()


def sort_by_key(keys):
    k = np.array(keys, dtype=np.uint64, copy=True)
    perm = np.empty(k.size, np.uint32)
    lib().ref_sort_by_key(P(k), C.c_size_t(k.size), P(perm))
    return k, perm


def compute_octree(keys, bucket, capacity=None):
    cap = capacity or max(64, 8 * (keys.size // max(1, bucket)) + 4096)
    leaves = np.empty(cap + 1, np.uint64)
    counts = np.empty(cap, np.uint32)
    it = C.c_int(0)
    nl = lib().ref_compute_octree(P(keys), C.c_size_t(keys.size), C.c_uint(bucket), P(leaves), P(counts), cap,
                                  C.byref(it))
    assert nl >= 0
    return leaves[:nl + 1].copy(), counts[:nl].copy(), it.value


def update_octree(keys, bucket, leaves, counts):
    nl = counts.size
    cap = 8 * nl + 64
    lv = np.empty(cap + 1, np.uint64)
    ct = np.empty(cap, np.uint32)
    lv[:nl + 1] = leaves
    ct[:nl] = counts
    conv = C.c_int(0)
    nn = lib().ref_update_octree(P(keys), C.c_size_t(keys.size), C.c_uint(bucket), P(lv), P(ct), nl, cap,
                                 C.byref(conv))
    assert nn >= 0
    return lv[:nn + 1].copy(), ct[:nn].copy(), bool(conv.value)


def compute_node_counts(leaves, keys):
    nl = leaves.size - 1
    counts = np.empty(nl, np.uint32)
    lib().ref_compute_node_counts(P(leaves), nl, P(keys), C.c_size_t(keys.size), P(counts))
    return counts


def rebalance_decision(leaves, counts, bucket):
    nl = counts.size
    ops = np.empty(nl + 1, np.int32)
    conv = C.c_int(0)
    lib().ref_rebalance_decision(P(leaves), P(counts), nl, C.c_uint(bucket), P(ops), C.byref(conv))
    return ops[:nl], bool(conv.value)


def build_octree(leaves):
    nl = leaves.size - 1
    ni = (nl - 1) // 7
    nn = nl + ni
    t = dict(prefixes=np.zeros(nn, np.uint64), childOffsets=np.zeros(nn + 1, np.int32),
             parents=np.zeros(max(1, (nn - 1) // 8), np.int32), levelRange=np.zeros(23, np.int32),
             internalToLeaf=np.zeros(nn, np.int32), leafToInternal=np.zeros(nn, np.int32))
    lib().ref_build_octree(P(leaves), nl, P(t["prefixes"]), P(t["childOffsets"]), P(t["parents"]),
                           P(t["levelRange"]), P(t["internalToLeaf"]), P(t["leafToInternal"]))
    t["numLeaves"], t["numInternal"], t["numNodes"] = nl, ni, nn
    t["leaves"] = leaves
    return t


def node_centers(prefixes, box):
    lim, bc = _box(box)
    n = prefixes.size
    centers = np.empty((n, 3), np.float64)
    sizes = np.empty((n, 3), np.float64)
    lib().ref_node_centers(P(prefixes), n, P(lim), P(bc), P(centers), P(sizes))
    return centers, sizes


def _sfx(T):
    return "f64" if np.dtype(T) == np.float64 else "f32"


def find_neighbors(x, y, z, h, first, last, box, tree, layout, centers, sizes, ng0, ngmax, iterate_h):
    """returns (neighbors[n,ngmax], nc[n] incl. self, h (modified copy))"""
    lim, bc = _box(box)
    h = h.copy()
    n = last - first
    nb = np.zeros((n, ngmax), np.uint32)
    nc = np.zeros(n, np.uint32)
    ta = TreeArgs(tree["numLeaves"], P(tree["prefixes"]), P(tree["childOffsets"]), P(tree["internalToLeaf"]),
                  P(tree["levelRange"]), P(tree["leaves"]), P(layout), P(centers), P(sizes))
    f = getattr(lib(), "ref_find_neighbors_" + _sfx(h.dtype))
    f(P(x), P(y), P(z), P(h), C.c_uint32(first), C.c_uint32(last), P(lim), P(bc), C.byref(ta), C.c_uint(ng0),
      C.c_uint(ngmax), P(nb), P(nc), int(iterate_h))
    return nb, nc, h


def tables(T, sinc_index=6.0):
    wh = np.empty(20000, T)
    whd = np.empty(20000, T)
    K = C.c_double(0)
    getattr(lib(), "ref_tables_" + _sfx(T))(C.c_double(sinc_index), P(wh), P(whd), C.byref(K))
    return wh, whd, K.value


def sphynx_3D_k(n):
    return lib().ref_sphynx_3D_k(C.c_double(n))


def _common(first, last, ngmax, K, box):
    lim, bc = _box(box)
    return (C.c_uint32(first), C.c_uint32(last), C.c_uint(ngmax), C.c_double(K), P(lim), P(bc)), (lim, bc)


def xmass(first, last, ngmax, K, box, nb, nc, x, y, z, h, m, wh, whd, xm):
    a, _k = _common(first, last, ngmax, K, box)
    getattr(lib(), "ref_xmass_" + _sfx(h.dtype))(*a, P(nb), P(nc), P(x), P(y), P(z), P(h), P(m), P(wh), P(whd), P(xm))


def ve_def_gradh(first, last, ngmax, K, box, nb, nc, x, y, z, h, m, wh, whd, xm, kx, gradh):
    a, _k = _common(first, last, ngmax, K, box)
    getattr(lib(), "ref_ve_def_gradh_" + _sfx(h.dtype))(*a, P(nb), P(nc), P(x), P(y), P(z), P(h), P(m), P(wh),
                                                        P(whd), P(xm), P(kx), P(gradh))


def eos(first, last, temp, m, kx, xm, gradh, mui, gamma, prho, c, rho=None, p=None):
    getattr(lib(), "ref_eos_" + _sfx(m.dtype))(C.c_uint32(first), C.c_uint32(last), P(temp), P(m), P(kx), P(xm),
                                               P(gradh), C.c_double(mui), C.c_double(gamma), P(prho), P(c), P(rho),
                                               P(p))


def iad_divv_curlv(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, wh, whd, xm, kx, c11, c12, c13, c22,
                   c23, c33, divv, curlv, dV=None):
    a, _k = _common(first, last, ngmax, K, box)
    dV = dV or [None] * 6
    getattr(lib(), "ref_iad_divv_curlv_" + _sfx(h.dtype))(*a, P(nb), P(nc), P(x), P(y), P(z), P(vx), P(vy), P(vz),
                                                          P(h), P(wh), P(whd), P(xm), P(kx), P(c11), P(c12), P(c13),
                                                          P(c22), P(c23), P(c33), P(divv), P(curlv),
                                                          *[P(d) for d in dV])


def av_switches(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, c, c11, c12, c13, c22, c23, c33, wh, whd,
                kx, xm, divv, min_dt, alphamin, alphamax, decay_constant, alpha):
    a, _k = _common(first, last, ngmax, K, box)
    getattr(lib(), "ref_av_switches_" + _sfx(h.dtype))(*a, P(nb), P(nc), P(x), P(y), P(z), P(vx), P(vy), P(vz), P(h),
                                                       P(c), P(c11), P(c12), P(c13), P(c22), P(c23), P(c33), P(wh),
                                                       P(whd), P(kx), P(xm), P(divv), C.c_double(min_dt),
                                                       C.c_double(alphamin), C.c_double(alphamax),
                                                       C.c_double(decay_constant), P(alpha))


def momentum_energy(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, m, prho, c, c11, c12, c13, c22, c23,
                    c33, atmin, atmax, ramp, wh, kx, xm, alpha, dV, ax, ay, az, du, kcour, av_clean):
    a, _k = _common(first, last, ngmax, K, box)
    dV = dV or [None] * 6
    return getattr(lib(), "ref_momentum_energy_" + _sfx(h.dtype))(
        *a, P(nb), P(nc), P(x), P(y), P(z), P(vx), P(vy), P(vz), P(h), P(m), P(prho), P(c), P(c11), P(c12), P(c13),
        P(c22), P(c23), P(c33), C.c_double(atmin), C.c_double(atmax), C.c_double(ramp), P(wh), P(kx), P(xm), P(alpha),
        *[P(d) for d in dV], P(ax), P(ay), P(az), P(du), C.c_float(kcour), int(av_clean))


def rho_timestep(first, last, divv, krho):
    return getattr(lib(), "ref_rho_timestep_" + _sfx(divv.dtype))(C.c_uint32(first), C.c_uint32(last), P(divv),
                                                                  C.c_double(krho))


def positions(first, last, dt, dt_m1, box, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az, temp, du, du_m1, h, mui,
              gamma):
    lim, bc = _box(box)
    getattr(lib(), "ref_positions_" + _sfx(h.dtype))(C.c_uint32(first), C.c_uint32(last), C.c_double(dt),
                                                     C.c_double(dt_m1), P(lim), P(bc), P(x), P(y), P(z), P(vx), P(vy),
                                                     P(vz), P(x_m1), P(y_m1), P(z_m1), P(ax), P(ay), P(az), P(temp),
                                                     P(du), P(du_m1), P(h), C.c_double(mui), C.c_double(gamma))


def drift(idx, dt, dt_back, dt_m1, rung, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az, temp, du, du_m1, mui, gamma):
    """driftKernel (sph/include/sph/positions_gpu.cu:47-90) for the particles idx (uint32), built from the reference's
    positionUpdate / energyUpdate"""
    idx = np.ascontiguousarray(idx, np.uint32)
    dt_m1 = np.ascontiguousarray(dt_m1, np.float32)
    getattr(lib(), "ref_drift_" + _sfx(vx.dtype))(P(idx), C.c_uint32(idx.size), C.c_float(dt), C.c_float(dt_back),
                                                  P(dt_m1), P(rung), P(x), P(y), P(z), P(vx), P(vy), P(vz), P(x_m1),
                                                  P(y_m1), P(z_m1), P(ax), P(ay), P(az), P(temp), P(du), P(du_m1),
                                                  C.c_double(mui), C.c_double(gamma))


def find_rung_ranges(min_dt, group_dt, num_rungs):
    """sph::findRungRanges<false> (sph/include/sph/ts_rungs.hpp:115-129) on sorted float32 group_dt"""
    gd = np.ascontiguousarray(group_dt, np.float32)
    out = np.zeros(5, np.uint32)
    lib().ref_find_rung_ranges(C.c_float(min_dt), P(gd), C.c_uint32(gd.size), C.c_int(num_rungs), P(out))
    return [int(v) for v in out]


def rung_timestep(group_dt, max_dt):
    """sph::rungTimestep (ts_rungs.hpp:131-145), the reference's own code on one rank (oracle/ref_bdt.cpp supplies
    the five device primitives it calls for host memory).  Returns (sorted group_dt, ordering, nextDt, totDt,
    numRungs, rungRanges)"""
    gd = np.ascontiguousarray(group_dt, np.float32).copy()
    idx = np.zeros(gd.size, np.uint32)
    next_dt, tot_dt, num_rungs, rr = C.c_float(0), C.c_float(0), C.c_int(0), np.zeros(5, np.uint32)
    lib().ref_rung_timestep(P(gd), P(idx), C.c_uint32(gd.size), C.c_float(max_dt), C.byref(next_dt), C.byref(tot_dt),
                            C.byref(num_rungs), P(rr))
    return gd, idx, np.float32(next_dt.value), np.float32(tot_dt.value), num_rungs.value, [int(v) for v in rr]


def minimum_group_dt(tot_dt, elapsed_dt, num_rungs, substep, ts_rung_ranges, group_dt, num_groups):
    """sph::minimumGroupDt (ts_rungs.hpp:147-157) as HydroVeBdtProp calls it inside a hierarchy
    (ve_hydro_bdt.hpp:312-313): group_dt[:num_groups] is sorted, the ordering is continued by a sequence up to
    ts_rung_ranges[4].  Returns (group_dt with its head sorted, ordering, dt, rungRanges)"""
    gd = np.ascontiguousarray(group_dt, np.float32).copy()
    tr = np.ascontiguousarray(ts_rung_ranges, np.uint32)
    idx = np.zeros(max(int(tr[4]), num_groups), np.uint32)
    dt, rr = C.c_float(0), np.zeros(5, np.uint32)
    lib().ref_minimum_group_dt(C.c_float(tot_dt), C.c_float(elapsed_dt), C.c_int(num_rungs), C.c_int(substep), P(tr),
                               P(gd), P(idx), C.c_uint32(num_groups), C.byref(dt), P(rr))
    return gd, idx, np.float32(dt.value), [int(v) for v in rr]


def positions_rung(idx, dt, dt_m1, rung, box, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az, temp, du, du_m1, h,
                   mui, gamma):
    """computePositionsKernel (positions_gpu.cu:118-161) for the particles idx"""
    idx = np.ascontiguousarray(idx, np.uint32)
    dt_m1 = np.ascontiguousarray(dt_m1, np.float32)
    lim, bc = _box(box)
    getattr(lib(), "ref_positions_rung_" + _sfx(h.dtype))(P(idx), C.c_uint32(idx.size), C.c_float(dt), P(dt_m1),
                                                          P(rung), P(lim), P(bc), P(x), P(y), P(z), P(vx), P(vy),
                                                          P(vz), P(x_m1), P(y_m1), P(z_m1), P(ax), P(ay), P(az),
                                                          P(temp), P(du), P(du_m1), P(h), C.c_double(mui),
                                                          C.c_double(gamma))


def update_h(first, last, ng0, nc, h):
    getattr(lib(), "ref_update_h_" + _sfx(h.dtype))(C.c_uint32(first), C.c_uint32(last), C.c_uint(ng0), P(nc), P(h))


def conserved_quantities(x, y, z, vx, vy, vz, m, temp, mui, gamma):
    """sphexa::localConservedQuantities (main/src/observables/conserved_quantities.hpp:48-107) ->
    [eKin, eInt, 0, linmom(3), angmom(3)]"""
    out = np.zeros(9)
    real = m.dtype
    fn = lib().ref_conserved_f64 if real == np.float64 else lib().ref_conserved_f32
    a = [np.ascontiguousarray(v, np.float64) for v in (x, y, z)]
    b = [np.ascontiguousarray(v, real) for v in (vx, vy, vz, m)]
    t = np.ascontiguousarray(temp, np.float64)
    fn(P(a[0]), P(a[1]), P(a[2]), P(b[0]), P(b[1]), P(b[2]), P(b[3]), P(t), C.c_size_t(m.size), C.c_double(mui),
       C.c_double(gamma), P(out))
    return out


class RefTurbulence:
    """sph::TurbulenceData<double, CpuTag> + the CPU branch of driveTurbulence (sph/include/sph/hydro_turb/*)"""

    def __init__(self, solWeight=0.5, stMaxModes=100000, Lbox=1.0, stEnergyPrefac=5.0e-3, stMachVelocity=0.3,
                 epsilon=1e-15, rngSeed=251299, stSpectForm=1, powerLawExp=5.0 / 3.0, anglesExp=2.0):
        L = lib()
        L.ref_turb_create.restype = C.c_void_p
        self.h = C.c_void_p(L.ref_turb_create(*[C.c_double(float(v)) for v in (
            solWeight, stMaxModes, Lbox, stEnergyPrefac, stMachVelocity, epsilon, rngSeed, stSpectForm, powerLawExp,
            anglesExp)]))
        self.numModes = L.ref_turb_num_modes(self.h)

    def state(self):
        n = self.numModes
        sc = np.zeros(4)
        out = dict(modes=np.zeros(3 * n), amplitudes=np.zeros(n), phases=np.zeros(6 * n), phasesReal=np.zeros(3 * n),
                   phasesImag=np.zeros(3 * n))
        lib().ref_turb_get(self.h, P(sc), P(out["modes"]), P(out["amplitudes"]), P(out["phases"]), P(out["phasesReal"]),
                           P(out["phasesImag"]))
        out.update(variance=sc[0], decayTime=sc[1], solWeight=sc[2], solWeightNorm=sc[3], numModes=n)
        return out

    def drive(self, first, last, x, y, z, ax, ay, az, dt):
        """in place on ax, ay, az (float64 or float32 numpy arrays)"""
        fn = lib().ref_turb_drive_f64 if ax.dtype == np.float64 else lib().ref_turb_drive_f32
        fn(self.h, C.c_uint32(first), C.c_uint32(last), P(x), P(y), P(z), P(ax), P(ay), P(az), C.c_double(dt))

    def drive_groups(self, starts, ends, x, y, z, ax, ay, az, dt):
        """driveTurbulence with a GroupView of the given groups (TurbVeBdtProp: the active rungs); one noise update"""
        fn = lib().ref_turb_drive_groups_f64 if ax.dtype == np.float64 else lib().ref_turb_drive_groups_f32
        starts = np.ascontiguousarray(starts, np.uint32)
        ends = np.ascontiguousarray(ends, np.uint32)
        fn(self.h, C.c_uint32(starts.size), P(starts), P(ends), P(x), P(y), P(z), P(ax), P(ay), P(az), C.c_double(dt))

    def close(self):
        if self.h:
            lib().ref_turb_destroy(self.h)
            self.h = None


# ---- multi-rank domain pieces (host functions of cstone)
def sfc_assignment(num_ranks, counts, leaves, old_bounds=None):
    """makeSfcAssignment + limitBoundaryShifts (domaindecomp.hpp:115-181): (bounds[P+1], rank counts[P])"""
    counts = np.ascontiguousarray(counts, np.uint32)
    leaves = np.ascontiguousarray(leaves, np.uint64)
    bounds = np.zeros(num_ranks + 1, np.uint64)
    rc = np.zeros(num_ranks, np.uint32)
    old = np.ascontiguousarray(old_bounds, np.uint64) if old_bounds is not None else None
    lib().ref_sfc_assignment(num_ranks, P(counts), P(leaves), counts.size, P(old), P(bounds), P(rc))
    return bounds, rc


def initial_global_tree(num_ranks):
    buf = np.zeros(1 << 16, np.uint64)
    n = lib().ref_initial_global_tree(num_ranks, P(buf), buf.size)
    assert n > 0
    return buf[:n + 1].copy()


def halo_radii(h, layout, first, last, search_ext=1.0):
    h = np.ascontiguousarray(h, np.float64)
    layout = np.ascontiguousarray(layout, np.uint32)
    radii = np.zeros(layout.size - 1, np.float32)
    lib().ref_halo_radii(P(h), P(layout), first, last, C.c_float(search_ext), P(radii))
    return radii


def find_halos(tree, leaves, radii, box, first, last):
    """findHalos (traversal/collisions.hpp:78-105) on a linked octree from build_octree: int flags per leaf"""
    lim, bc = _box(box)
    leaves = np.ascontiguousarray(leaves, np.uint64)
    radii = np.ascontiguousarray(radii, np.float32)
    flags = np.zeros(leaves.size - 1, np.int32)
    lib().ref_find_halos(P(tree["prefixes"]), P(tree["childOffsets"]), P(tree["internalToLeaf"]), P(leaves), P(radii),
                         P(lim), P(bc), first, last, P(flags))
    return flags


def node_layout(counts, halo_flags, first, last):
    counts = np.ascontiguousarray(counts, np.uint32)
    halo_flags = np.ascontiguousarray(halo_flags, np.int32)
    layout = np.zeros(counts.size + 1, np.uint32)
    lib().ref_node_layout(P(counts), P(halo_flags), counts.size, first, last, P(layout))
    return layout


def mach(vx, vy, vz, c, num_particles_global):
    """(localMachSquareSum, calculateMachRMS) of main/src/observables/turbulence_mach_rms.hpp:38-83, the reference's own
    templates instantiated on a four-vector dataset"""
    real = c.dtype
    fn = lib().ref_mach_f64 if real == np.float64 else lib().ref_mach_f32
    a = [np.ascontiguousarray(v, real) for v in (vx, vy, vz, c)]
    out = np.zeros(2)
    fn(P(a[0]), P(a[1]), P(a[2]), P(a[3]), C.c_size_t(c.size), C.c_uint64(num_particles_global), P(out))
    return float(out[0]), float(out[1])


# ---- focus tree (oracle/ref_focus.cpp): the reference's own CPU functions -----------------------------------------
def min_mac_r2(prefixes, inv_theta_eff, box):
    """computeMinMacR2 (traversal/macs.hpp:58-71) per node: [n, 4] = center, squared acceptance radius"""
    lim, bc = _box(box)
    prefixes = np.ascontiguousarray(prefixes, np.uint64)
    out = np.zeros((prefixes.size, 4), np.float64)
    lib().ref_min_mac_r2(P(prefixes), prefixes.size, C.c_float(inv_theta_eff), P(lim), P(bc), P(out))
    return out


def mark_macs(tree, centers4, box, focus_start, focus_end):
    """markMacs (traversal/macs.hpp:240-264) on a linked octree from build_octree: int8 per node"""
    lim, bc = _box(box)
    centers4 = np.ascontiguousarray(centers4, np.float64)
    out = np.zeros(tree["numNodes"], np.int8)
    lib().ref_mark_macs(P(tree["prefixes"]), P(tree["childOffsets"]), P(tree["parents"]), P(tree["levelRange"]),
                        P(tree["internalToLeaf"]), P(tree["leafToInternal"]), tree["numLeaves"], tree["numInternal"],
                        P(centers4), P(lim), P(bc), C.c_uint64(int(focus_start)), C.c_uint64(int(focus_end)), P(out))
    return out


def rebalance_decision_essential(tree, counts, macs, focus_start, focus_end, bucket):
    counts = np.ascontiguousarray(counts, np.uint32)
    macs = np.ascontiguousarray(macs, np.int8)
    ops = np.zeros(tree["numNodes"], np.int32)
    lib().ref_rebalance_decision_essential(P(tree["prefixes"]), tree["numNodes"], P(tree["childOffsets"]),
                                           P(tree["parents"]), P(counts), P(macs), C.c_uint64(int(focus_start)),
                                           C.c_uint64(int(focus_end)), C.c_uint(bucket), P(ops))
    return ops


def enforce_keys(keys, tree, ops):
    """enforceKeys (focus/rebalance.hpp:245-258): modifies ops in place, returns the ResolutionStatus"""
    keys = np.ascontiguousarray(keys, np.uint64)
    assert ops.dtype == np.int32
    return lib().ref_enforce_keys(P(keys), keys.size, P(tree["prefixes"]), P(tree["childOffsets"]), P(tree["parents"]),
                                  P(ops))


def protect_ancestors(tree, ops):
    assert ops.dtype == np.int32
    return bool(lib().ref_protect_ancestors(P(tree["prefixes"]), tree["numNodes"], P(tree["parents"]), P(ops)))


class FocusedOctreeSingleNode:
    """cstone::FocusedOctreeSingleNode<uint64_t> (focus/octree_focus.hpp:219-288), the reference's own class"""

    def __init__(self, bucket, theta):
        lib().ref_focus_create.restype = C.c_void_p
        self.h = C.c_void_p(lib().ref_focus_create(C.c_uint(bucket), C.c_float(theta)))

    def update(self, box, keys, focus_start, focus_end, mandatory=()):
        lim, bc = _box(box)
        keys = np.ascontiguousarray(keys, np.uint64)
        mk = np.ascontiguousarray(np.asarray(mandatory, np.uint64))
        return bool(lib().ref_focus_update(self.h, P(lim), P(bc), P(keys), C.c_size_t(keys.size),
                                           C.c_uint64(int(focus_start)), C.c_uint64(int(focus_end)),
                                           P(mk) if mk.size else None, mk.size))

    def leaves_counts(self):
        n = lib().ref_focus_num_leaves(self.h)
        leaves = np.zeros(n + 1, np.uint64)
        counts = np.zeros(n, np.uint32)
        lib().ref_focus_get(self.h, P(leaves), P(counts))
        return leaves, counts

    def __del__(self):
        try:
            lib().ref_focus_destroy(self.h)
        except Exception:  # noqa: BLE001
            pass


def find_peers_mac(rank, bounds, leaves, box, inv_theta_eff):
    """findPeersMac (traversal/peers.hpp:63-120): list of peer ranks"""
    lim, bc = _box(box)
    bounds = np.ascontiguousarray(bounds, np.uint64)
    leaves = np.ascontiguousarray(leaves, np.uint64)
    peers = np.zeros(bounds.size - 1, np.int32)
    lib().ref_find_peers_mac(rank, bounds.size - 1, P(bounds), P(leaves), leaves.size - 1, P(lim), P(bc),
                             C.c_float(inv_theta_eff), P(peers))
    return np.nonzero(peers)[0].tolist()
