"""Regenerates the committed golden fixtures.  Runs ONLY in the build container (needs /root/reference and
oracle/_ref/libsphexa_ref.so); the tests read the .npz files, never /root/reference.

  ve_fixture.npz     the 99-particle fixture of sph/test/example_data.txt (31 columns, order as in
                     sph/test/ve.cpp:74-78) + K = sphynx_3D_k(6) (ve.cpp:94)
  hilbert.npz        Hilbert keys of fixed integer coordinates computed by cstone::iHilbert, decodeHilbert
                     round trip, first-order curve (domain/test/unit/sfc/hilbert.cpp:101-201)
  sedov12_steps.npz  3 steps of Sedov 12^3 (f64) from the reference pipeline: keys, leaves, counts, nc,
                     all fields (used to pin the C restatement in oracle/ and as a box-independent fixture)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

from oracle import ref  # noqa: E402
from oracle.ref_pipeline import RefSim  # noqa: E402
import sphexa_b200_loader  # noqa: E402

pkg = sphexa_b200_loader.load()


def ve_fixture():
    d = np.loadtxt(os.path.join(REF, "sph/test/example_data.txt"))
    names = ("x y z vx vy vz h c c11 c12 c13 c22 c23 c33 p gradh rho0 sumwhrho0 sumwh dvxdx dvxdy dvxdz dvydx dvydy "
             "dvydz dvzdx dvzdy dvzdz alpha u divv").split()
    assert d.shape == (99, 31)
    out = {n: np.ascontiguousarray(d[:, i]) for i, n in enumerate(names)}
    out["K"] = np.float64(ref.sphynx_3D_k(6.0))
    np.savez_compressed(os.path.join(HERE, "ve_fixture.npz"), **out)


def hilbert():
    rng = np.random.Generator(np.random.MT19937(5489))
    n = 4096
    ix, iy, iz = (rng.integers(0, 1 << 21, n, dtype=np.uint32) for _ in range(3))
    # corners and first-order cells
    ix[:8] = [0, 0, 0, 0, 1, 1, 1, 1]
    iy[:8] = [0, 0, 1, 1, 0, 0, 1, 1]
    iz[:8] = [0, 1, 0, 1, 0, 1, 0, 1]
    ix[8:16] = ix[:8] << 20
    iy[8:16] = iy[:8] << 20
    iz[8:16] = iz[:8] << 20
    ix[16], iy[16], iz[16] = (1 << 21) - 1, (1 << 21) - 1, (1 << 21) - 1
    keys = ref.hilbert_encode(ix, iy, iz)
    dx, dy, dz = ref.hilbert_decode(keys)
    assert np.array_equal(dx, ix) and np.array_equal(dy, iy) and np.array_equal(dz, iz)
    np.savez_compressed(os.path.join(HERE, "hilbert.npz"), ix=ix, iy=iy, iz=iz, keys=keys)


def sedov_steps():
    f, box, s = pkg.ic.sedov(12)
    sim = RefSim(f, box, s)
    out = {}
    for k, v in f.items():
        out["init_" + k] = v
    for step in range(3):
        sim.step()
        pre = f"s{step}_"
        out[pre + "keys"] = sim.keys.copy()
        out[pre + "leaves"] = sim.ftree[0]
        out[pre + "counts"] = sim.ftree[1]
        out[pre + "nc"] = sim.nc.copy()
        out[pre + "nbsum"] = np.array([np.sort(sim.nb[i, :sim.nc[i] - 1]).astype(np.uint64).sum() for i in range(sim.n)],
                                      np.uint64)
        out[pre + "dt"] = np.float64(sim.minDt)
        for k in ("x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "xm", "kx", "gradh", "prho", "c", "c11",
                  "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay", "az", "du"):
            out[pre + k] = sim.f[k].copy()
    np.savez_compressed(os.path.join(HERE, "sedov12_steps.npz"), **out)


if __name__ == "__main__":
    ve_fixture()
    hilbert()
    sedov_steps()
    print("golden fixtures written to", HERE)
