"""TEST INFRASTRUCTURE ONLY: the reference's CPU hot path sequenced like HydroVeProp.

Every numerical operation is executed by the reference's own templates through
oracle/_ref/libsphexa_ref.so (see oracle/ref_wrap.cpp); this file only sequences them the way
main/src/propagator/ve_hydro.hpp:116-217 and domain/include/cstone/domain/domain.hpp:196-244 do
for a single rank.  Used by tests/ (parity checker) and by bench.py as the timed CPU baseline
("kind": "reference").
"""
import time
import numpy as np

from . import ref

CONSERVED = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha"]
DEPENDENT = ["xm", "kx", "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay",
             "az"]


class RefSim:
    def __init__(self, fields, box, settings, real=np.float64, bucket_size=None, bucket_size_focus=64, av_clean=False):
        self.real = real
        self.f = {k: np.ascontiguousarray(v, dtype=(np.float64 if k in ("x", "y", "z", "temp") else real)).copy()
                  for k, v in fields.items()}
        self.n = self.f["x"].size
        self.box = (list(box[0]), list(box[1]))
        self.s = dict(settings)
        self.ng0, self.ngmax = int(settings.get("ng0", 100)), int(settings.get("ngmax", 150))
        self.bucket = bucket_size or max(bucket_size_focus, settings.get("numParticlesGlobal", self.n) // 100)
        self.bucket_focus = bucket_size_focus
        self.wh, self.whd, self.K = ref.tables(real)
        self.minDt, self.minDt_m1, self.ttot = settings["minDt"], settings["minDt_m1"], 0.0
        self.av_clean = av_clean
        self.gtree = None
        self.ftree = None
        for k in DEPENDENT:
            self.f[k] = np.zeros(self.n, real)
        self.f["du"] = np.zeros(self.n, np.float64)
        if av_clean:
            for k in ("dV11", "dV12", "dV13", "dV22", "dV23", "dV33"):
                self.f[k] = np.zeros(self.n, real)
        self.times = {}

    def _update_tree(self, tree, bucket):
        if tree is None:
            lv, ct, _ = ref.compute_octree(self.keys, bucket)
        else:
            lv, ct, _ = ref.update_octree(self.keys, bucket, tree[0], tree[1])
        return (lv, ct)

    def sync(self):
        t0 = time.perf_counter()
        f = self.f
        lim, bc = self.box
        for d, k in enumerate("xyz"):
            if bc[d] != 1:
                lim[2 * d], lim[2 * d + 1] = float(f[k].min()), float(f[k].max())
        keys = ref.compute_sfc_keys(f["x"], f["y"], f["z"], self.box)
        self.keys, self.perm = ref.sort_by_key(keys)
        self.gtree = self._update_tree(self.gtree, self.bucket)
        for k in CONSERVED:
            f[k] = np.ascontiguousarray(f[k][self.perm])
        self.ftree = self._update_tree(self.ftree, self.bucket_focus)
        self.tree = ref.build_octree(self.ftree[0])
        self.centers, self.sizes = ref.node_centers(self.tree["prefixes"], self.box)
        self.layout = np.zeros(self.ftree[1].size + 1, np.uint32)
        np.cumsum(self.ftree[1], out=self.layout[1:])
        self.times["domain::sync"] = time.perf_counter() - t0

    def computeForces(self):
        self.sync()
        f, n, K, box = self.f, self.n, self.K, self.box
        T = time.perf_counter
        t = T()
        self.nb, self.nc, f["h"] = ref.find_neighbors(f["x"], f["y"], f["z"], f["h"], 0, n, box, self.tree,
                                                      self.layout, self.centers, self.sizes, self.ng0, self.ngmax,
                                                      True)
        self.times["FindNeighbors"] = T() - t
        t = T()
        nb, nc, ngmax, wh, whd = self.nb, self.nc, self.ngmax, self.wh, self.whd
        ref.xmass(0, n, ngmax, K, box, nb, nc, f["x"], f["y"], f["z"], f["h"], f["m"], wh, whd, f["xm"])
        self.times["XMass"] = T() - t
        t = T()
        ref.ve_def_gradh(0, n, ngmax, K, box, nb, nc, f["x"], f["y"], f["z"], f["h"], f["m"], wh, whd, f["xm"],
                         f["kx"], f["gradh"])
        self.times["Normalization & Gradh"] = T() - t
        t = T()
        ref.eos(0, n, f["temp"], f["m"], f["kx"], f["xm"], f["gradh"], self.s["mui"], self.s["gamma"], f["prho"],
                f["c"])
        self.times["EquationOfState"] = T() - t
        t = T()
        dV = [f[k] for k in ("dV11", "dV12", "dV13", "dV22", "dV23", "dV33")] if self.av_clean else None
        ref.iad_divv_curlv(0, n, ngmax, K, box, nb, nc, f["x"], f["y"], f["z"], f["vx"], f["vy"], f["vz"], f["h"], wh,
                           whd, f["xm"], f["kx"], f["c11"], f["c12"], f["c13"], f["c22"], f["c23"], f["c33"],
                           f["divv"], f["curlv"], dV)
        self.minDtRho = ref.rho_timestep(0, n, f["divv"], self.s.get("Krho", 0.06))
        self.times["IadVelocityDivCurl"] = T() - t
        t = T()
        self.alpha_pre = f["alpha"].copy()  # alpha entering AVswitches, in this step's SFC order (for tests)
        ref.av_switches(0, n, ngmax, K, box, nb, nc, f["x"], f["y"], f["z"], f["vx"], f["vy"], f["vz"], f["h"],
                        f["c"], f["c11"], f["c12"], f["c13"], f["c22"], f["c23"], f["c33"], wh, whd, f["kx"], f["xm"],
                        f["divv"], self.minDt, 0.05, 1.0, 0.2, f["alpha"])
        self.times["AVswitches"] = T() - t
        t = T()
        self.minDtCourant = ref.momentum_energy(0, n, ngmax, K, box, nb, nc, f["x"], f["y"], f["z"], f["vx"], f["vy"],
                                                f["vz"], f["h"], f["m"], f["prho"], f["c"], f["c11"], f["c12"],
                                                f["c13"], f["c22"], f["c23"], f["c33"], 0.1, 0.2, 10.0, wh, f["kx"],
                                                f["xm"], f["alpha"], dV, f["ax"], f["ay"], f["az"], f["du"],
                                                self.s.get("Kcour", 0.2), self.av_clean)
        self.times["MomentumAndEnergy"] = T() - t

    def integrate(self):
        t = time.perf_counter()
        f, n = self.f, self.n
        # computeTimestep (ts_global.hpp:96-112), single rank
        dt = min(self.minDtCourant, self.minDtRho, 1.1 * self.minDt)
        self.ttot += dt
        self.minDt_m1, self.minDt = self.minDt, dt
        ref.positions(0, n, self.minDt, self.minDt_m1, self.box, f["x"], f["y"], f["z"], f["vx"], f["vy"], f["vz"],
                      f["x_m1"], f["y_m1"], f["z_m1"], f["ax"], f["ay"], f["az"], f["temp"], f["du"], f["du_m1"],
                      f["h"], self.s["mui"], self.s["gamma"])
        ref.update_h(0, n, self.ng0, self.nc, f["h"])
        self.times["Timestep+UpdateQuantities"] = time.perf_counter() - t

    def step(self):
        self.computeForces()
        self.integrate()
