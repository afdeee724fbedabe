"""GPU parity of the SPH-VE operators and the full time step against the reference CPU build.

Tolerance (north star): <= 1e-10 relative per step in fp64 mode, measured as max|gpu-ref| / max|ref|
per field (lattice forces cancel to ~0, a per-element relative error is undefined there).  Mixed mode
(reference default precision) is compared at float accuracy.
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import to_dev, to_host, slab_from_rowmajor, sorted_sets_equal, rel_err

pytestmark = pytest.mark.gpu

TOL = {np.float64: 1e-10, np.float32: 2e-4}
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def ref_state(pkg, case, side, real, steps, av_clean=False):
    """advance the reference pipeline `steps` steps, then run computeForces once more and keep everything"""
    from oracle.ref_pipeline import RefSim
    f, box, s = pkg.ic.CASES[case](side, "f64" if real == np.float64 else "mixed")
    sim = RefSim(f, box, s, real=real, av_clean=av_clean)
    for _ in range(steps):
        sim.step()
    pre = {k: v.copy() for k, v in sim.f.items()}
    pre_dt = (sim.minDt, sim.minDt_m1)
    sim.computeForces()
    return sim, pre, pre_dt, box, s


def device_fields(sim, names, real):
    d = {}
    for k in names:
        d[k] = to_dev(sim.f[k])
    d["nc"] = to_dev(sim.nc)
    return d


ALL = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha", "xm", "kx",
       "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay", "az", "du"]


@pytest.mark.parametrize("real", [np.float64, np.float32])
@pytest.mark.parametrize("case,side,steps", [("sedov", 20, 3), ("noh", 24, 2), ("kh", 40, 1)])
def test_operators_one_by_one(pkg, ctx, ref, real, case, side, steps):
    """each operator fed with the reference's inputs for that operator, outputs compared field by field"""
    import torch
    from sph_exa_b200 import ops
    av_clean = case == "kh"
    sim, pre, pre_dt, box, s = ref_state(pkg, case, side, real, steps, av_clean)
    n, ngmax = sim.n, sim.ngmax
    b = pkg.make_box(*sim.box)
    names = ALL + (["dV11", "dV12", "dV13", "dV22", "dV23", "dV33"] if av_clean else [])
    want = sim.f
    nb = slab_from_rowmajor(ctx, sim.nb, sim.nc, ngmax)
    wh, whd = to_dev(sim.wh), to_dev(sim.whd)
    hc = ops.hydro_const(sim.K, wh, whd, gamma=s["gamma"], muiConst=s["mui"], Kcour=s.get("Kcour", 0.2),
                         minDt=pre_dt[0], minDt_m1=pre_dt[1], avClean=av_clean)
    tol = TOL[real]

    def fresh(outputs):
        d = device_fields(sim, names, real)
        for k in outputs:
            d[k] = torch.full_like(d[k], -1.0)
        return d

    d = fresh(["xm"])
    ops.computeXMass(ctx, 0, n, d, b, hc, nb)
    assert rel_err(to_host(d["xm"]), want["xm"]) < tol

    d = fresh(["kx", "gradh"])
    ops.computeVeDefGradh(ctx, 0, n, d, b, hc, nb)
    assert rel_err(to_host(d["kx"]), want["kx"]) < tol
    assert rel_err(to_host(d["gradh"]), want["gradh"]) < tol

    d = fresh(["prho", "c"])
    ops.computeEOS(ctx, 0, n, d, hc)
    assert rel_err(to_host(d["prho"]), want["prho"]) < tol
    assert rel_err(to_host(d["c"]), want["c"]) < tol

    outs = ["c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv"] + (names[-6:] if av_clean else [])
    d = fresh(outs)
    ops.computeIadDivvCurlv(ctx, 0, n, d, b, hc, nb)
    # the off-diagonal c_ij (and dV_ij) of a near-lattice are cancellation residuals: scale by the whole tensor
    cscale = max(np.abs(want[k]).max() for k in ("c11", "c12", "c13", "c22", "c23", "c33"))
    for k in outs:
        scale = cscale if k.startswith("c") and k != "curlv" else max(np.abs(want[k]).max(), 1e-300)
        err = np.abs(to_host(d[k]).astype(np.float64) - want[k]).max() / scale
        # mixed mode: divv / curlv of the near-lattice are float cancellation residuals (measured 2.5e-4)
        assert err < (2 * tol if real == np.float32 and k in ("divv", "curlv") else tol), (k, err)

    # AV switches update alpha in place: start from the pre-force alpha
    d = fresh([])
    d["alpha"] = to_dev(sim.alpha_pre)
    ops.computeAVswitches(ctx, 0, n, d, b, hc, nb)
    assert rel_err(to_host(d["alpha"]), want["alpha"]) < tol

    d = fresh(["ax", "ay", "az", "du"])
    dt_c = ops.computeMomentumEnergy(ctx, 0, n, d, b, hc, nb)
    acc_scale = max(np.abs(want[k]).max() for k in ("ax", "ay", "az"))
    for k in ("ax", "ay", "az"):
        assert np.abs(to_host(d[k]).astype(np.float64) - want[k]).max() / acc_scale < tol, k
    assert rel_err(to_host(d["du"]), want["du"]) < tol
    assert abs(dt_c - sim.minDtCourant) <= tol * abs(sim.minDtCourant)

    dts = ops.computeTimestep(ctx, 0, n, d, hc)
    assert abs(dts[0] - sim.minDtCourant) <= tol * abs(sim.minDtCourant)
    assert abs(dts[1] - sim.minDtRho) <= tol * abs(sim.minDtRho)
    assert abs(dts[2] - min(sim.minDtCourant, sim.minDtRho, 1.1 * pre_dt[0])) <= tol * dts[2]

    # integrate
    sim.integrate()
    ops.computePositionsAndH(ctx, 0, n, d, b, hc, sim.minDt, sim.minDt_m1)
    for k in ("x", "y", "z", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "h"):
        a, w = to_host(d[k]).astype(np.float64), sim.f[k].astype(np.float64)
        scale = max(np.abs(w).max(), 1e-300)
        assert np.abs(a - w).max() / scale < tol, k


def test_golden_ve_fixture(pkg, ctx):
    """the reference's own known-answer test (sph/test/ve.cpp:112-233, example_data.txt) through the CUDA kernels:
    particle 0 with the other 98 as neighbors, open box, sinc^6, K = sphynx_3D_k(6)"""
    import torch
    from sph_exa_b200 import ops
    g = np.load(os.path.join(GOLDEN, "ve_fixture.npz"))
    n = 99
    f = {k: np.ascontiguousarray(g[k]) for k in g.files}
    K = float(f["K"].item())
    mpart = 3.781038064465603e26
    m = np.full(n, mpart)
    xm = mpart / f["rho0"]
    kx = K * xm / f["h"] ** 3
    prho = f["p"] / (kx * m * m * f["gradh"])
    nbrow = np.zeros((n, 98), np.uint32)
    nbrow[0] = np.arange(1, 99)
    nc = np.ones(n, np.uint32)
    nc[0] = 99
    nb = slab_from_rowmajor(ctx, nbrow, nc, 98)
    wh, whd, _ = pkg.ic.kernel_tables(np.float64)
    dwh, dwhd = to_dev(wh), to_dev(whd)
    box = pkg.make_box([-1e9, 1e9] * 3, [0, 0, 0])
    hc = ops.hydro_const(K, dwh, dwhd, ngmax=98, minDt=0.3, avClean=False)

    def dev(**over):
        d = dict(x=f["x"], y=f["y"], z=f["z"], h=f["h"], m=m, vx=f["vx"], vy=f["vy"], vz=f["vz"], c=f["c"],
                 c11=f["c11"], c12=f["c12"], c13=f["c13"], c22=f["c22"], c23=f["c23"], c33=f["c33"], xm=xm, kx=kx,
                 prho=prho, gradh=f["gradh"], alpha=f["alpha"], divv=f["divv"], curlv=np.zeros(n), ax=np.zeros(n),
                 ay=np.zeros(n), az=np.zeros(n), du=np.zeros(n), temp=np.zeros(n))
        d.update(over)
        out = {k: to_dev(v) for k, v in d.items()}
        out["nc"] = to_dev(nc)
        return out

    d = dev(xm=np.zeros(n))
    ops.computeXMass(ctx, 0, 1, d, box, hc, nb)
    xmass = to_host(d["xm"])[0]
    assert abs(mpart / xmass - 34.515038498081417) < 7.33e-7          # ve.cpp:229

    d = dev()
    ops.computeVeDefGradh(ctx, 0, 1, d, box, hc, nb)
    kxo, gh = to_host(d["kx"])[0], to_host(d["gradh"])[0]
    assert abs(kxo * mpart / xm[0] - 3.4662283566584293e1) < 8e-7      # ve.cpp:218-220
    assert abs(gh - 0.98699067585409861) < 5e-7
    assert abs(kxo - 1.0042661134076782) < 3e-7

    d = dev()
    ops.computeAVswitches(ctx, 0, 1, d, box, hc, nb)
    assert abs(to_host(d["alpha"])[0] - 0.93941905320351171) < 2e-9    # ve.cpp:121

    dV = {k: np.zeros(n) for k in ("dV11", "dV12", "dV13", "dV22", "dV23", "dV33")}
    d = dev(**dV)
    c_in = {k: f[k].copy() for k in ("c11", "c12", "c13", "c22", "c23", "c33")}
    ops.computeIadDivvCurlv(ctx, 0, 1, d, box, hc, nb)
    iad = [to_host(d[k])[0] for k in ("c11", "c12", "c13", "c22", "c23", "c33")]
    gold_iad = [1.9296619855715329e-18, -1.7838691836843698e-20, -1.2892885646884301e-20, 1.9482845913025683e-18,
                1.635410357476855e-20, 1.9246939006338132e-18]                 # ve.cpp:149-154
    for a, w in zip(iad, gold_iad):
        assert abs(a - w) < 1e-10 and abs(a - w) < 1e-7 * abs(w) + 1e-24
    # ve.cpp:124-141 evaluates divv/curlv with the FIXTURE's c_ij(0), which is not the IAD result above (different
    # normalisation); the fused kernel uses its own IAD result exactly like computeIadDivvCurlvImpl
    # (iad_divv_curlv.hpp:86-92), so that known answer is checked on the oracle side only (test_oracle.py).

    gv = {k: f[k] for k in ("dvxdx", "dvxdy", "dvxdz", "dvydx", "dvydy", "dvydz", "dvzdx", "dvzdy", "dvzdz")}
    sdV = dict(dV11=gv["dvxdx"], dV12=gv["dvxdy"] + gv["dvydx"], dV13=gv["dvxdz"] + gv["dvzdx"], dV22=gv["dvydy"],
               dV23=gv["dvydz"] + gv["dvzdy"], dV33=gv["dvzdz"])                # symmetrizeGradV ve.cpp:157-169
    for clean, gold in ((True, (-505548.68073726865, 303384.91384746187, -1767463.9739728321, 8.5525242525359648e12,
                                (0.023, 0.053, 0.043, 7.1e5))),
                        (False, (-521261.07791667967, -74471.016515749841, -1730426.827721074,
                                 7.1838438980436924e12, (0.022, 0.064, 0.042, 3.1e5)))):
        d = dev(**sdV)
        hc2 = ops.hydro_const(K, dwh, dwhd, ngmax=98, minDt=0.3, avClean=clean)
        ops.computeMomentumEnergy(ctx, 0, 1, d, box, hc2, nb)
        got = [to_host(d[k])[0] for k in ("ax", "ay", "az", "du")]
        for a, w, t in zip(got, gold[:4], gold[4]):
            assert abs(a - w) < t                                            # ve.cpp:186-209


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_full_steps_sedov(pkg, ctx, ref, real):
    """Config 1 scaled to 30^3: sync + forces + integrate for 4 steps through the tier-2 ABI;
    keys/leaves/counts/neighbor counts bit-exact, fields within tolerance, conserved totals"""
    from oracle.ref_pipeline import RefSim
    from sph_exa_b200 import sim as S
    dtype = pkg.F64 if real == np.float64 else pkg.MIXED
    f, box, s = pkg.ic.sedov(30, "f64" if real == np.float64 else "mixed")
    rs = RefSim(f, box, s, real=real)
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, dtype=dtype, tables=(rs.wh, rs.whd, rs.K))
    gs.upload(f)
    tol = TOL[real]
    for step in range(4):
        rs.step()
        gs.step()
        out = gs.download(names=["x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "du", "ax", "ay", "az", "nc",
                                 "xm", "kx", "c11", "divv"], with_keys=True)
        assert gs.numLeaves == rs.ftree[1].size
        assert gs.numGlobalLeaves == rs.gtree[1].size
        assert np.array_equal(gs.leaves(), rs.ftree[0])
        assert np.array_equal(gs.leafCounts(), rs.ftree[1])
        assert np.array_equal(gs.globalLeaves(), rs.gtree[0])
        assert np.array_equal(gs.globalCounts(), rs.gtree[1])
        if real == np.float64:
            assert np.array_equal(out["keys"], rs.keys), f"step {step}"
            assert np.array_equal(out["nc"], rs.nc), f"step {step}"
            assert gs.totalNeighbors() == int(rs.nc.sum())
        assert abs(gs.minDt - rs.minDt) <= tol * rs.minDt
        for k in ("x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "du", "xm", "kx", "c11", "divv"):
            assert rel_err(out[k], rs.f[k]) < tol * (step + 1), (step, k)
        acc = max(np.abs(rs.f[k]).max() for k in ("ax", "ay", "az"))
        for k in ("ax", "ay", "az"):
            assert np.abs(out[k].astype(np.float64) - rs.f[k]).max() / acc < tol * (step + 1), (step, k)
    # conserved totals (conserved_quantities.hpp:50-110)
    got = gs.download()
    tg, tr = pkg.ic.conserved_totals(got, s), pkg.ic.conserved_totals(rs.f, s)
    assert abs(tg["etot"] - tr["etot"]) <= tol * abs(tr["etot"])
    assert abs(tg["mass"] - tr["mass"]) <= 1e-12
    pscale = float(np.sum(np.abs(rs.f["m"].astype(np.float64) * rs.f["vx"].astype(np.float64)))) + 1e-300
    assert abs(tg["px"] - tr["px"]) / pscale < tol * 3
    gs.close()


@pytest.mark.parametrize("case,side", [("noh", 30), ("turbulence", 30), ("kh", 48)])
def test_full_steps_other_cases(pkg, ctx, ref, case, side):
    """synthetic Noh (open box, bbox reduction every step), turbulence (periodic, jittered) and KH (non-cubic box,
    density contrast) ICs: 3 steps, f64 mode"""
    from oracle.ref_pipeline import RefSim
    from sph_exa_b200 import sim as S
    f, box, s = pkg.ic.CASES[case](side)
    rs = RefSim(f, box, s)
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, tables=(rs.wh, rs.whd, rs.K))
    gs.upload(f)
    for step in range(3):
        rs.step()
        gs.step()
        out = gs.download(names=["x", "h", "vx", "temp", "alpha", "du", "nc"], with_keys=True)
        assert np.array_equal(out["keys"], rs.keys), (case, step)
        assert np.array_equal(gs.leaves(), rs.ftree[0])
        assert np.array_equal(out["nc"], rs.nc), (case, step)
        assert abs(gs.minDt - rs.minDt) <= 1e-10 * rs.minDt
        for k in ("x", "h", "vx", "temp", "alpha", "du"):
            assert rel_err(out[k], rs.f[k]) < 1e-10 * (step + 1), (case, step, k)
    gs.close()


def test_neighbor_lists_of_a_step(pkg, ctx, ref):
    """bit-exact neighbor sets of a full step (after sync + h iteration) read back from the simulation"""
    import ctypes as C
    from oracle.ref_pipeline import RefSim
    from sph_exa_b200 import sim as S, binding as B
    f, box, s = pkg.ic.turbulence(24)
    rs = RefSim(f, box, s)
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, tables=(rs.wh, rs.whd, rs.K))
    gs.upload(f)
    rs.step()
    gs.step()
    n, ngmax = rs.n, rs.ngmax
    nc = gs.download(names=["nc"])["nc"]
    import torch
    rm_d = torch.zeros((n, ngmax), dtype=torch.int32, device="cuda")
    assert B.lib.sphb200_neighbors_to_rowmajor(ctx.h, C.c_void_p(B.lib.sphb200_sim_neighbors(gs.h)),
                                               C.c_void_p(to_dev(nc).data_ptr()), C.c_uint32(n), C.c_uint(ngmax),
                                               C.c_void_p(rm_d.data_ptr())) == 0
    rm = to_host(rm_d, np.uint32)
    assert sorted_sets_equal(rm.copy(), nc, rs.nb.copy(), rs.nc, ngmax)
    gs.close()


def test_errors_are_reported(pkg, ctx):
    """precondition violations surface as exceptions carrying the library's message (reference: std::runtime_error)"""
    from sph_exa_b200 import sim as S
    f, box, s = pkg.ic.sedov(8)
    s2 = dict(s, ng0=200, ngmax=150)
    with pytest.raises(pkg.SphError, match="ng0 should be smaller than ngmax"):
        S.HydroVeProp(ctx, box, s2, capacity=512)
    gs = S.HydroVeProp(ctx, box, s, capacity=100)
    with pytest.raises(pkg.SphError, match="capacity"):
        gs.upload(f)
    gs.close()


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_step_host_equals_upload_step_download(pkg, ctx, real):
    """sphb200_sim_step_host stages the PCIe copies beside the kernels; the state it returns must be bit-identical to
    upload + step + download on the same inputs, step after step"""
    import torch
    from sph_exa_b200 import sim as S, binding as B
    dtype = pkg.F64 if real == np.float64 else pkg.MIXED
    f, box, s = pkg.ic.sedov(24, "f64" if real == np.float64 else "mixed")
    n = f["x"].size
    a = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    b = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    a.upload(f)
    bufs = []
    for _ in range(2):
        d = {}
        for k in B.CONSERVED:
            want = np.float64 if k in B.ALWAYS_F64 else real
            d[k] = torch.empty(n, dtype=torch.float64 if want == np.float64 else torch.float32).pin_memory()
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
        bufs.reverse()
    a.close()
    b.close()


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_conserved_quantities(pkg, ctx, ref, real):
    """device reduction of computeConservedQuantities vs the reference template (oracle/_ref), f64 and mixed types;
    tolerance: 1e-13 of the absolute-value sums (different but fixed summation order), ncsum exact"""
    from sph_exa_b200 import ops
    f, box, s = pkg.ic.turbulence(20, "f64" if real == np.float64 else "mixed")
    n = f["x"].size
    rng = np.random.default_rng(5)
    for k in ("vx", "vy", "vz"):
        f[k] = rng.normal(size=n).astype(real)
    nc = rng.integers(50, 150, n).astype(np.uint32)
    first, last = 17, n - 5
    want = ref.conserved_quantities(*[f[k][first:last] for k in ("x", "y", "z", "vx", "vy", "vz", "m", "temp")],
                                    s["mui"], s["gamma"])
    d = {k: to_dev(np.ascontiguousarray(f[k], np.float64 if k in ("x", "y", "z", "temp") else real))
         for k in ("x", "y", "z", "h", "m", "vx", "vy", "vz", "temp")}
    d["nc"] = to_dev(nc)
    hc = ops.hydro_const(1.0, 0, 0, gamma=s["gamma"], muiConst=s["mui"])
    got = ops.computeConservedQuantities(ctx, first, last, d, hc)
    mv = float(np.abs(f["m"].astype(np.float64) * f["vx"]).sum())
    tol = 1e-13 if real == np.float64 else 1e-6
    assert abs(got["ecin"] - want[0]) <= tol * want[0]
    assert abs(got["eint"] - want[1]) <= tol * want[1]
    for k, name in enumerate(("linmom_x", "linmom_y", "linmom_z")):
        assert abs(got[name] - want[3 + k]) <= tol * mv
    for k, name in enumerate(("angmom_x", "angmom_y", "angmom_z")):
        assert abs(got[name] - want[6 + k]) <= tol * mv * float(np.abs(f["x"]).max())
    assert got["ncsum"] == float(nc[first:last].sum())
    assert got["etot"] == got["ecin"] + got["eint"]


@pytest.mark.parametrize("real", [np.float64, np.float32])
@pytest.mark.parametrize("case,side", [("sedov", 20), ("noh", 24)])
def test_early_courant_and_ranged_momentum(pkg, ctx, ref, real, case, side):
    """the operator-level pieces of sphb200_sim_step_host's schedule: AVswitches + Courant == AVswitches followed by
    momentumEnergy's Courant step (bit for bit) and == the reference's value; momentumEnergy in ranges with the
    finished ranges integrated in between == one launch followed by one position update (bit for bit)"""
    import torch
    from sph_exa_b200 import ops
    sim, pre, pre_dt, box, s = ref_state(pkg, case, side, real, 2)
    n, ngmax = sim.n, sim.ngmax
    b = pkg.make_box(*sim.box)
    nb = slab_from_rowmajor(ctx, sim.nb, sim.nc, ngmax)
    wh, whd = to_dev(sim.wh), to_dev(sim.whd)
    hc = ops.hydro_const(sim.K, wh, whd, gamma=s["gamma"], muiConst=s["mui"], Kcour=s.get("Kcour", 0.2),
                         minDt=pre_dt[0], minDt_m1=pre_dt[1])

    def state():
        d = device_fields(sim, ALL, real)
        d["du"] = to_dev(sim.f["du"])
        d["alpha"] = to_dev(sim.alpha_pre)
        return d

    one = state()
    ops.computeAVswitches(ctx, 0, n, one, b, hc, nb)
    dt_mom = ops.computeMomentumEnergy(ctx, 0, n, one, b, hc, nb)
    dts_one = ops.computeTimestep(ctx, 0, n, one, hc)

    two = state()
    ops.computeAVswitchesCourant(ctx, 0, n, two, b, hc, nb)
    dts_two = ops.computeTimestep(ctx, 0, n, two, hc)
    assert dts_two == dts_one and dts_two[0] == dt_mom
    assert abs(dt_mom - sim.minDtCourant) <= TOL[real] * sim.minDtCourant
    assert np.array_equal(to_host(two["alpha"]), to_host(one["alpha"]))

    # ranged: uneven ranges (multiples of 128 from `first`), each integrated before the next one starts
    first, last = 128, n - 5
    ref_run = state()
    ops.computeAVswitches(ctx, first, last, ref_run, b, hc, nb[first // 32 * ngmax * 32:].contiguous())
    nb1 = nb[first // 32 * ngmax * 32:].contiguous()
    ref_run["nc"] = one["nc"]
    ops.computeMomentumEnergy(ctx, first, last, ref_run, b, hc, nb1)
    dt, dt_m1 = 0.9 * pre_dt[0], pre_dt[0]
    ops.computePositionsAndH(ctx, first, last, ref_run, b, hc, dt, dt_m1)

    rng_run = state()
    ops.computeAVswitches(ctx, first, last, rng_run, b, hc, nb1)
    cuts = [first, first + 128, first + 128 * 9, first + 128 * 10, last]
    for k, (rb, re) in enumerate(zip(cuts[:-1], cuts[1:])):
        ops.computeMomentumEnergyRange(ctx, first, last, rb, re, k == 0, rng_run, b, hc, nb1)
        ops.computePositionsAndH(ctx, rb, re, rng_run, b, hc, dt, dt_m1)
    for k in ("x", "y", "z", "vx", "vy", "vz", "x_m1", "temp", "du_m1", "h", "ax", "ay", "az", "du"):
        assert np.array_equal(to_host(rng_run[k]), to_host(ref_run[k])), k
    with pytest.raises(pkg.binding.SphError):
        ops.computeMomentumEnergyRange(ctx, first, last, first + 7, last, True, rng_run, b, hc, nb1)

    # next-step h into a separate array
    h_new = torch.full_like(one["h"], -1.0)
    ops.updateSmoothingLengthInto(ctx, first, last, sim.ng0, one["nc"], one["h"], h_new)
    got = to_host(h_new)
    assert np.array_equal(got[first:last], to_host(ref_run["h"])[first:last])
    assert np.all(got[:first] == -1.0) and np.all(got[last:] == -1.0)
    assert np.array_equal(to_host(one["h"]), sim.f["h"].astype(real))


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_mach_rms(pkg, ctx, ref, real):
    """machSquareSumGpu / TurbulenceMachRMS against the reference's own localMachSquareSum / calculateMachRMS
    (turbulence_mach_rms.hpp:38-83 compiled into oracle/_ref); run-to-run reproducible"""
    from sph_exa_b200 import ops, sim as S
    sim, pre, pre_dt, box, s = ref_state(pkg, "turbulence", 20, real, 2)
    n = sim.n
    d = device_fields(sim, ["x", "h", "vx", "vy", "vz", "c"], real)
    first, last = 37, n - 11
    want, _ = ref.mach(*(sim.f[k][first:last].astype(real) for k in ("vx", "vy", "vz", "c")), n)
    got = ops.machSquareSum(ctx, first, last, d)
    assert abs(got - want) <= (1e-13 if real == np.float64 else 1e-6) * want
    assert ops.machSquareSum(ctx, first, last, d) == got
    assert ops.machSquareSum(ctx, 5, 5, d) == 0.0
    # through the sim: after a step, c holds the EOS sound speed of that step
    f, box, st = pkg.ic.CASES["turbulence"](16, "f64" if real == np.float64 else "mixed")
    gs = S.HydroVeProp(ctx, box, st, capacity=f["x"].size, dtype=pkg.F64 if real == np.float64 else pkg.MIXED)
    gs.upload(f)
    gs.step()
    gs.step()
    out = gs.download(names=["vx", "vy", "vz", "c"])
    _, want = ref.mach(out["vx"], out["vy"], out["vz"], out["c"], f["x"].size)
    assert abs(gs.machRMS() - want) <= 1e-6 * want
    assert abs(gs.machRMS(8 * f["x"].size) - want / np.sqrt(8)) <= 1e-6 * want
    gs.close()
