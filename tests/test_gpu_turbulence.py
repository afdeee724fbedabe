"""GPU parity of the turbulence stirring row against sph::TurbulenceData / driveTurbulence compiled from the
reference headers (oracle/_ref).  Host-side state (mode table, OU phases, projected phases) must be bit-identical: same
arithmetic, same std::mt19937 / std::normal_distribution stream.  Accelerations: <= 1e-10 of max|a_stir| in f64 (device
vs host libm sin/cos differ in the last bits), float accuracy in mixed mode."""
import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_drive_matches_reference(pkg, ctx, ref, real):
    """three OU steps: phases bit-identical, accelerations within tolerance; [first,last) is a sub-range"""
    from sph_exa_b200.turbulence import Turbulence, computeStirring
    rt, gt = ref.RefTurbulence(), Turbulence()
    n = 20000
    rng = np.random.default_rng(2)
    x, y, z = [rng.uniform(-0.5, 0.5, n) for _ in range(3)]
    acc0 = [rng.normal(size=n).astype(real) for _ in range(3)]
    a_ref = [a.copy() for a in acc0]
    d = dict(x=to_dev(x), y=to_dev(y), z=to_dev(z), ax=to_dev(acc0[0]), ay=to_dev(acc0[1]), az=to_dev(acc0[2]))
    first, last = 11, n - 7
    tol = 1e-10 if real == np.float64 else 2e-5
    for step, dt in enumerate((1e-4, 3e-4, 2e-4)):
        rt.drive(first, last, x, y, z, *a_ref, dt)
        gt.drive(ctx, first, last, d, dt)
        sa, sb = gt.state(), rt.state()
        for k in ("phases", "phasesReal", "phasesImag"):
            assert np.array_equal(sa[k], sb[k]), (step, k)
        stir = max(np.abs(w.astype(np.float64) - a0).max() for w, a0 in zip(a_ref, acc0))
        for k, w in zip(("ax", "ay", "az"), a_ref):
            got = to_host(d[k])
            assert np.abs(got.astype(np.float64) - w).max() <= tol * stir, (step, k)
            assert np.array_equal(got[:first], acc0["xyz".index(k[1])][:first])     # outside the range: untouched
            assert np.array_equal(got[last:], acc0["xyz".index(k[1])][last:])
    # operator level (computeStirringGpu signature, per-mode sin/cos) on the same phases
    st = gt.state()
    d2 = dict(x=d["x"], y=d["y"], z=d["z"], ax=to_dev(np.zeros(n, real)), ay=to_dev(np.zeros(n, real)),
              az=to_dev(np.zeros(n, real)))
    computeStirring(ctx, 0, n, d2, to_dev(st["modes"]), to_dev(st["phasesReal"]), to_dev(st["phasesImag"]),
                    to_dev(st["amplitudes"]), st["solWeightNorm"])
    want = [np.zeros(n) for _ in range(3)]
    from oracle import orc
    orc.stirring(0, n, x, y, z, *want, st["modes"], st["phasesReal"], st["phasesImag"], st["amplitudes"],
                 st["solWeightNorm"])
    scale = max(np.abs(w).max() for w in want)
    for k, w in zip(("ax", "ay", "az"), want):
        assert np.abs(to_host(d2[k]).astype(np.float64) - w).max() <= tol * scale
    rt.close()
    gt.close()


def test_power_law_spectrum_uses_generic_kernel(pkg, ctx, ref):
    """spectForm 2 samples random angles: many distinct wave numbers -> per-mode sin/cos path"""
    from sph_exa_b200.turbulence import Turbulence
    kw = dict(stSpectForm=2, rngSeed=99)
    rt, gt = ref.RefTurbulence(**kw), Turbulence(**kw)
    n = 5000
    rng = np.random.default_rng(8)
    x, y, z = [rng.uniform(-0.5, 0.5, n) for _ in range(3)]
    a_ref = [np.zeros(n) for _ in range(3)]
    d = dict(x=to_dev(x), y=to_dev(y), z=to_dev(z), ax=to_dev(np.zeros(n)), ay=to_dev(np.zeros(n)), az=to_dev(np.zeros(n)))
    rt.drive(0, n, x, y, z, *a_ref, 1e-4)
    gt.drive(ctx, 0, n, d, 1e-4)
    scale = max(np.abs(w).max() for w in a_ref)
    for k, w in zip(("ax", "ay", "az"), a_ref):
        assert np.abs(to_host(d[k]) - w).max() <= 1e-10 * scale
    rt.close()
    gt.close()


def test_turb_ve_propagator(pkg, ctx, ref):
    """TurbVeProp (turb_ve.hpp:113-118): hydro forces, then stirring with the current minDt, then integrate -- tier-2
    sim with an attached turbulence object vs the reference pipeline + reference stirring, 3 steps"""
    from oracle.ref_pipeline import RefSim
    from sph_exa_b200 import sim as S
    from sph_exa_b200.turbulence import Turbulence
    f, box, s = pkg.ic.turbulence(24)
    rs = RefSim(f, box, s)
    rt, gt = ref.RefTurbulence(), Turbulence()
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, tables=(rs.wh, rs.whd, rs.K))
    gs.attachTurbulence(gt)
    gs.upload(f)
    for step in range(3):
        rs.sync()
        rs.computeForces()
        rt.drive(0, rs.n, rs.f["x"], rs.f["y"], rs.f["z"], rs.f["ax"], rs.f["ay"], rs.f["az"], rs.minDt)
        rs.integrate()
        gs.step()
        out = gs.download(names=["x", "y", "z", "vx", "vy", "vz", "ax", "ay", "az", "temp"])
        acc = max(np.abs(rs.f[k]).max() for k in ("ax", "ay", "az"))
        for k in ("ax", "ay", "az"):
            assert np.abs(out[k] - rs.f[k]).max() <= 1e-9 * acc, (step, k)
        for k in ("x", "y", "z", "vx", "vy", "vz", "temp"):
            assert np.abs(out[k] - rs.f[k]).max() <= 1e-9 * max(np.abs(rs.f[k]).max(), 1e-300), (step, k)
        assert abs(gs.minDt - rs.minDt) <= 1e-10 * rs.minDt
    gs.close()
    gt.close()
    rt.close()


def test_step_host_with_stirring_equals_step(pkg, ctx):
    """TurbVeProp through sphb200_sim_step_host: the last stage runs in 16 ranges, each stirred by its own kernel launch,
    but the Ornstein-Uhlenbeck sequence must advance once per time step — state and driver bit-identical to
    upload + step + download"""
    import torch
    from sph_exa_b200 import sim as S, binding as B
    from sph_exa_b200.turbulence import Turbulence
    f, box, s = pkg.ic.turbulence(24)
    n = f["x"].size
    assert n >= 16 * 128  # the ranged path
    a = S.HydroVeProp(ctx, box, s, capacity=n)
    b = S.HydroVeProp(ctx, box, s, capacity=n)
    ta, tb = Turbulence(), Turbulence()
    a.attachTurbulence(ta)
    b.attachTurbulence(tb)
    a.upload(f)
    bufs = []
    for _ in range(2):
        d = {k: torch.empty(n, dtype=torch.float64).pin_memory() for k in B.CONSERVED}
        bufs.append(d)
    for k in B.CONSERVED:
        bufs[0][k].numpy()[:] = f[k]

    def struct(d):
        st = B.Fields()
        for k in B.FIELD_NAMES:
            setattr(st, k, d[k].data_ptr() if k in d else None)
        return st

    for step in range(3):
        a.step()
        want = a.download()
        b.step_host(struct(bufs[0]), struct(bufs[1]), n)
        for k in B.CONSERVED:
            assert np.array_equal(bufs[1][k].numpy(), want[k]), (step, k)
        assert b.minDt == a.minDt
        assert np.array_equal(ta.state()["phases"], tb.state()["phases"]), step
        assert ta.rngState() == tb.rngState()
        bufs.reverse()
    a.close()
    b.close()
    ta.close()
    tb.close()
