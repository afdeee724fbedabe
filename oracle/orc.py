"""TEST INFRASTRUCTURE ONLY: ctypes loader + single-rank step driver for oracle/liboracle.so, the plain-C
restatement of the reference algorithms (oracle/sphexa_oracle.c).  f64 mode only."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


class OBox(C.Structure):
    _fields_ = [("lim", C.c_double * 6), ("len", C.c_double * 3), ("inv", C.c_double * 3), ("bc", C.c_int * 3)]


class OTree(C.Structure):
    _fields_ = [("childOffsets", C.c_void_p), ("internalToLeaf", C.c_void_p), ("layout", C.c_void_p),
                ("centers", C.c_void_p), ("sizes", C.c_void_p)]


OF_NAMES = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha", "xm", "kx",
            "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay", "az", "du",
            "dV11", "dV12", "dV13", "dV22", "dV23", "dV33"]


class OF(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in OF_NAMES]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
        _lib = C.CDLL(_PATH)
        _lib.orc_hilbert.restype = C.c_uint64
        _lib.orc_momentum_energy.restype = C.c_double
        _lib.orc_rho_timestep.restype = C.c_double
    return _lib


def P(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_box(box):
    b = OBox()
    lib().orc_make_box((C.c_double * 6)(*box[0]), (C.c_int * 3)(*box[1]), C.byref(b))
    return b


def hilbert(ix, iy, iz):
    return np.array([lib().orc_hilbert(int(a), int(b), int(c)) for a, b, c in zip(ix, iy, iz)], np.uint64)


def hilbert_decode(keys):
    out = np.zeros((len(keys), 3), np.uint32)
    a, b, c = C.c_uint(), C.c_uint(), C.c_uint()
    for i, k in enumerate(keys):
        lib().orc_hilbert_decode(C.c_uint64(int(k)), C.byref(a), C.byref(b), C.byref(c))
        out[i] = (a.value, b.value, c.value)
    return out[:, 0], out[:, 1], out[:, 2]


def sfc_keys(x, y, z, box):
    keys = np.empty(x.size, np.uint64)
    b = make_box(box)
    lib().orc_sfc_keys(P(x), P(y), P(z), C.c_size_t(x.size), C.byref(b), P(keys))
    return keys


def sort_by_key(keys):
    k = np.array(keys, np.uint64, copy=True)
    perm = np.empty(k.size, np.uint32)
    lib().orc_sort_by_key(P(k), C.c_size_t(k.size), P(perm))
    return k, perm


def compute_octree(keys, bucket):
    cap = max(4096, 16 * (keys.size // max(1, bucket)) + 4096)
    leaves = np.empty(cap + 2, np.uint64)
    counts = np.empty(cap + 2, np.uint32)
    it = C.c_int(0)
    nl = lib().orc_compute_octree(P(keys), C.c_size_t(keys.size), C.c_uint(bucket), P(leaves), P(counts), cap,
                                  C.byref(it))
    assert nl > 0
    return leaves[:nl + 1].copy(), counts[:nl].copy(), it.value


def update_octree(keys, bucket, leaves, counts):
    nl = counts.size
    cap = 4096 * min(nl, 64) + 8 * nl + 64
    nlv = np.empty(cap + 2, np.uint64)
    nct = np.empty(cap + 2, np.uint32)
    conv = C.c_int(0)
    nn = lib().orc_update_octree(P(keys), C.c_size_t(keys.size), C.c_uint(bucket), P(leaves), P(counts), nl, P(nlv),
                                 P(nct), cap, C.byref(conv))
    assert nn > 0
    return nlv[:nn + 1].copy(), nct[:nn].copy(), bool(conv.value)


def build_octree(leaves):
    nl = leaves.size - 1
    ni = (nl - 1) // 7
    nn = nl + ni
    t = dict(prefixes=np.zeros(nn, np.uint64), childOffsets=np.zeros(nn + 1, np.int32),
             parents=np.zeros(max(1, (nn - 1) // 8), np.int32), levelRange=np.zeros(23, np.int32),
             internalToLeaf=np.zeros(nn, np.int32), leafToInternal=np.zeros(nn, np.int32), numLeaves=nl,
             numInternal=ni, numNodes=nn, leaves=leaves)
    lib().orc_build_octree(P(leaves), nl, P(t["prefixes"]), P(t["childOffsets"]), P(t["parents"]), P(t["levelRange"]),
                           P(t["internalToLeaf"]), P(t["leafToInternal"]))
    return t


def node_centers(prefixes, box):
    n = prefixes.size
    c = np.empty((n, 3))
    s = np.empty((n, 3))
    b = make_box(box)
    lib().orc_node_centers(P(prefixes), n, C.byref(b), P(c), P(s))
    return c, s


def find_neighbors(x, y, z, h, first, last, box, tree, layout, centers, sizes, ng0, ngmax, iterate_h):
    h = h.copy()
    n = last - first
    nb = np.zeros((n, ngmax), np.uint32)
    nc = np.zeros(n, np.uint32)
    t = OTree(P(tree["childOffsets"]), P(tree["internalToLeaf"]), P(layout), P(centers), P(sizes))
    b = make_box(box)
    lib().orc_find_neighbors(P(x), P(y), P(z), P(h), C.c_uint32(first), C.c_uint32(last), C.byref(b), C.byref(t),
                             C.c_uint(ng0), C.c_uint(ngmax), int(iterate_h), P(nb), P(nc))
    return nb, nc, h


def tables(sinc_index=6.0):
    wh, whd = np.empty(20000), np.empty(20000)
    K = C.c_double(0)
    lib().orc_tables(C.c_double(sinc_index), P(wh), P(whd), C.byref(K))
    return wh, whd, K.value


def of_struct(f):
    s = OF()
    for n in OF_NAMES:
        a = f.get(n)
        setattr(s, n, a.ctypes.data if a is not None else None)
    return s


class OracleSim:
    """single-rank HydroVeProp sequence on the C restatement (same schedule as oracle/ref_pipeline.RefSim)"""

    CONSERVED = OF_NAMES[:14]

    def __init__(self, fields, box, settings, av_clean=False):
        self.f = {k: np.ascontiguousarray(v, np.float64).copy() for k, v in fields.items()}
        self.n = self.f["x"].size
        self.box = (list(box[0]), list(box[1]))
        self.s = dict(settings)
        self.ng0, self.ngmax = int(settings.get("ng0", 100)), int(settings.get("ngmax", 150))
        self.bucket = max(64, settings.get("numParticlesGlobal", self.n) // 100)
        self.wh, self.whd, self.K = tables()
        self.minDt, self.minDt_m1 = settings["minDt"], settings["minDt_m1"]
        self.gtree = self.ftree = None
        for k in OF_NAMES[14:31]:
            self.f[k] = np.zeros(self.n)
        if av_clean:
            for k in OF_NAMES[31:]:
                self.f[k] = np.zeros(self.n)
        self.av_clean = av_clean

    def _tree(self, tree, bucket):
        if tree is None:
            lv, ct, _ = compute_octree(self.keys, bucket)
        else:
            lv, ct, _ = update_octree(self.keys, bucket, tree[0], tree[1])
        return lv, ct

    def step(self):
        f, n, L = self.f, self.n, lib()
        lim, bc = self.box
        for d, k in enumerate("xyz"):
            if bc[d] != 1:
                lim[2 * d], lim[2 * d + 1] = float(f[k].min()), float(f[k].max())
        self.keys, perm = sort_by_key(sfc_keys(f["x"], f["y"], f["z"], self.box))
        self.gtree = self._tree(self.gtree, self.bucket)
        for k in self.CONSERVED:
            f[k] = np.ascontiguousarray(f[k][perm])
        self.ftree = self._tree(self.ftree, 64)
        tree = build_octree(self.ftree[0])
        centers, sizes = node_centers(tree["prefixes"], self.box)
        layout = np.zeros(self.ftree[1].size + 1, np.uint32)
        np.cumsum(self.ftree[1], out=layout[1:])
        self.nb, self.nc, f["h"] = find_neighbors(f["x"], f["y"], f["z"], f["h"], 0, n, self.box, tree, layout,
                                                  centers, sizes, self.ng0, self.ngmax, True)
        b, fs = make_box(self.box), of_struct(f)
        a = (C.c_uint32(0), C.c_uint32(n), C.c_uint(self.ngmax), C.c_double(self.K), C.byref(b), P(self.nb),
             P(self.nc), C.byref(fs))
        L.orc_xmass(*a, P(self.wh))
        L.orc_ve_def_gradh(*a, P(self.wh), P(self.whd))
        L.orc_eos(C.c_uint32(0), C.c_uint32(n), C.byref(fs), C.c_double(self.s["mui"]), C.c_double(self.s["gamma"]))
        L.orc_iad_divv_curlv(*a, P(self.wh))
        self.minDtRho = L.orc_rho_timestep(C.c_uint32(0), C.c_uint32(n), P(f["divv"]),
                                           C.c_double(self.s.get("Krho", 0.06)))
        L.orc_av_switches(*a, P(self.wh), C.c_double(self.minDt), C.c_double(0.05), C.c_double(1.0), C.c_double(0.2))
        self.minDtCourant = L.orc_momentum_energy(*a, P(self.wh), C.c_double(0.1), C.c_double(0.2), C.c_double(10.0),
                                                  C.c_float(self.s.get("Kcour", 0.2)), int(self.av_clean))
        dt = min(self.minDtCourant, self.minDtRho, 1.1 * self.minDt)
        self.minDt_m1, self.minDt = self.minDt, dt
        L.orc_positions_update_h(C.c_uint32(0), C.c_uint32(n), C.c_double(self.minDt), C.c_double(self.minDt_m1),
                                 C.byref(b), C.byref(fs), C.c_double(self.s["mui"]), C.c_double(self.s["gamma"]),
                                 C.c_uint(self.ng0), P(self.nc))


def conserved_quantities(f, first, last, mui, gamma, nc=None):
    """[eKin, eInt, egrav, linmom(3), angmom(3), ncsum] over [first,last) (conserved_quantities.hpp:48-159)"""
    out = np.zeros(10)
    a = {k: np.ascontiguousarray(f[k], np.float64) for k in ("x", "y", "z", "vx", "vy", "vz", "m", "temp")}
    ncp = P(np.ascontiguousarray(nc, np.uint32)) if nc is not None else None
    lib().orc_conserved_quantities(C.c_uint32(first), C.c_uint32(last), P(a["x"]), P(a["y"]), P(a["z"]), P(a["vx"]),
                                   P(a["vy"]), P(a["vz"]), P(a["m"]), P(a["temp"]), ncp, C.c_double(mui),
                                   C.c_double(gamma), P(out))
    return out


def turb_phases(ou, sol_weight, modes):
    """computePhases (hydro_turb/phases.hpp:46-72) -> (phasesReal, phasesImag)"""
    n = modes.size // 3
    pr, pi = np.zeros(3 * n), np.zeros(3 * n)
    lib().orc_turb_phases(n, 3, P(np.ascontiguousarray(ou)), C.c_double(sol_weight), P(np.ascontiguousarray(modes)),
                          P(pr), P(pi))
    return pr, pi


def stirring(first, last, x, y, z, ax, ay, az, modes, pr, pi, amplitudes, sol_weight_norm):
    """computeStirring (hydro_turb/stirring.hpp:99-122), in place on ax, ay, az (float64)"""
    lib().orc_stirring(C.c_uint32(first), C.c_uint32(last), P(x), P(y), P(z), P(ax), P(ay), P(az), int(amplitudes.size),
                       P(modes), P(pr), P(pi), P(amplitudes), C.c_double(sol_weight_norm))


def mach_square_sum(vx, vy, vz, c):
    """localMachSquareSum (main/src/observables/turbulence_mach_rms.hpp:38-56): sum of norm2(V) / (c*c) with V in
    double and c*c in the storage type of c.  Restated (5 lines); pinned against the reference's own template compiled
    into oracle/_ref (ref.mach, tests/test_oracle.py::test_mach_restatement_equals_reference)."""
    v2 = vx.astype(np.float64) ** 2 + vy.astype(np.float64) ** 2 + vz.astype(np.float64) ** 2
    return float(np.sum(v2 / (c * c).astype(np.float64)))


def mach_rms(vx, vy, vz, c, num_particles_global):
    """calculateMachRMS (turbulence_mach_rms.hpp:70-83), one rank"""
    return float(np.sqrt(mach_square_sum(vx, vy, vz, c) / num_particles_global))
