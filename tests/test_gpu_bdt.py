"""Block time steps: the group operators of HydroVeBdtProp and the propagator itself on the GPU.

Reference: main/src/propagator/ve_hydro_bdt.hpp, sph/include/sph/{ts_rungs.hpp,ts_groups.cu,positions_gpu.cu}.  The
reference runs this propagator on GPUs only and its tests hold no vectors for it; the sequencing is pinned against the
reference's own GPU application in tests/test_gpu_reference_binary.py and the group kernels against the reference's own
kernels in tests/test_gpu_reference_kernels.py; here drift/advance are checked against the reference's positionUpdate/energyUpdate through oracle/_ref,
the rung search against findRungRanges<false>'s std::lower_bound restated with numpy, the whole propagator against
oracle/bdt.py::RefBdtSim, which sequences the reference's CPU operators over the active groups.
Tolerance: 1e-10 relative (f64) per substep like the global-time-step propagator; integers and group tables exact.
"""
import numpy as np
import pytest

from conftest import to_dev, to_host, slab_from_rowmajor, rel_err

pytestmark = pytest.mark.gpu


def test_sort_group_dt_and_rung_ranges_equal_reference_hierarchy(pkg, ctx, ref):
    """sphb200_sort_group_dt + sphb200_rung_ranges against the reference's own rungTimestep / minimumGroupDt
    (sph/ts_rungs.hpp:131-157 compiled for host memory, oracle/ref_bdt.cpp)"""
    from sph_exa_b200 import ops
    rng = np.random.Generator(np.random.MT19937(12))
    for n in (3, 640, 40_000):
        dt = np.exp(rng.uniform(np.log(1e-6), np.log(3e-5), n)).astype(np.float32)
        dt[rng.integers(0, n, n // 10)] = dt[0]
        want_dt, want_idx, next_dt, tot_dt, num_rungs, want_rr = ref.rung_timestep(dt, 1.0)
        d = to_dev(dt)
        idx = to_host(ops.sortGroupDt(ctx, d), np.uint32)
        assert np.array_equal(idx, want_idx) and np.array_equal(to_host(d), want_dt)
        _, mdt = ops.rungRanges(ctx, -1.0, d, n, 1)
        got_rungs = min(int(np.log2(np.float32(mdt[1]) / np.float32(mdt[0]))) + 1, 4)  # api.cu: bdtComputeRungs
        assert got_rungs == num_rungs and np.float32(mdt[0]) == next_dt
        rr, _ = ops.rungRanges(ctx, float(mdt[0]), d, n, got_rungs)
        assert rr == want_rr
        # inside the hierarchy: the head of the array is sorted again, boundaries for all rungs
        head = want_rr[2] if want_rr[2] > 2 else n
        dt2 = want_dt.copy()
        dt2[:head] = dt2[:head][rng.permutation(head)] * np.float32(0.9)
        d2 = to_dev(dt2)
        idx2 = to_host(ops.sortGroupDt(ctx, d2, head), np.uint32)
        w_dt, w_idx, w_min, w_rr = ref.minimum_group_dt(tot_dt, next_dt, num_rungs, 1, want_rr, dt2, head)
        assert np.array_equal(idx2, w_idx[:head]) and np.array_equal(to_host(d2), w_dt)
        rr2, mdt2 = ops.rungRanges(ctx, -1.0, d2, head, 4)
        assert rr2 == w_rr
        time_left, substeps_left = np.float32(tot_dt - next_dt), (1 << num_rungs) - 1
        assert min(np.float32(mdt2[0]), np.float32(time_left / np.float32(substeps_left))) == w_min


def test_sort_group_dt_and_rung_ranges(pkg, ctx):
    from oracle import bdt
    from sph_exa_b200 import ops
    rng = np.random.Generator(np.random.MT19937(11))
    for n in (1, 7, 5000, 70_001):
        dt = np.exp(rng.uniform(np.log(1e-6), np.log(1e-3), n)).astype(np.float32)
        dt[rng.integers(0, n, n // 10)] = dt[0]  # ties: the sort is stable
        if n > 100:
            dt[5] = np.finfo(np.float32).max
        d = to_dev(dt)
        idx = to_host(ops.sortGroupDt(ctx, d), np.uint32)
        want = np.argsort(dt, kind="stable")
        assert np.array_equal(idx, want)
        assert np.array_equal(to_host(d), dt[want])
        sdt = dt[want]
        for num_rungs in (1, 2, 3, 4):
            rr, mdt = ops.rungRanges(ctx, float(sdt[0]), d, n, num_rungs)
            assert rr == bdt.find_rung_ranges(sdt[0], sdt, num_rungs)
            assert mdt[0] == sdt[0] and mdt[1] == sdt[int(np.float32(0.4) * np.float32(n))]
        rr, _ = ops.rungRanges(ctx, -1.0, d, n, 4)
        assert rr == bdt.find_rung_ranges(sdt[0], sdt, 4)
        # partial sort of the leading groups only (minimumGroupDt)
        if n > 100:
            d2 = to_dev(dt)
            idx = to_host(ops.sortGroupDt(ctx, d2, n // 3), np.uint32)
            assert np.array_equal(idx, np.argsort(dt[:n // 3], kind="stable"))
            assert np.array_equal(to_host(d2)[n // 3:], dt[n // 3:])


@pytest.mark.parametrize("real", [np.float64, np.float32])
@pytest.mark.parametrize("case", ["sedov", "kh"])
def test_drift_and_advance_groups(pkg, ctx, ref, real, case):
    """driftPositionsGpu / computePositionsGpu on a subset of groups with per-particle rungs"""
    import torch
    from sph_exa_b200 import ops
    from test_gpu_hydro import ref_state, device_fields, ALL
    sim, pre, pre_dt, box, s = ref_state(pkg, case, 16 if case == "sedov" else 24, real, 2)
    n = sim.n
    names = ALL
    rng = np.random.Generator(np.random.MT19937(3))
    rung = rng.integers(0, 4, n).astype(np.uint8)
    dt_m1 = np.array([1.0e-6, 2.1e-6, 3.9e-6, 8.3e-6], np.float32) * np.float32(sim.minDt / 1e-6)
    # groups: runs of <= 32 particles, every third one selected
    delim = np.arange(0, n + 32, 32).clip(max=n).astype(np.uint32)
    sel = np.arange(0, delim.size - 1, 3)
    start, end = delim[:-1][sel].copy(), delim[1:][sel].copy()
    idx = np.concatenate([np.arange(a, b) for a, b in zip(start, end)]).astype(np.uint32)
    g = ops.Groups(0, n, to_dev(start), to_dev(end))
    b = pkg.make_box(*sim.box)
    wh, whd = to_dev(sim.wh), to_dev(sim.whd)
    hc = ops.hydro_const(sim.K, wh, whd, gamma=s["gamma"], muiConst=s["mui"])
    tol = 1e-13 if real == np.float64 else 2e-6
    moved = ("x", "y", "z", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1")

    def check(d, want):
        outside = np.ones(n, bool)
        outside[idx] = False
        for k in moved:
            got = to_host(d[k])
            assert np.array_equal(got[outside], sim.f[k][outside]), k  # the other groups are untouched
            assert rel_err(got[idx], want[k][idx]) < tol, k

    dt_step = np.float32(0.7 * sim.minDt)
    for dt, dt_back in ((dt_step, np.float32(0)), (np.float32(0), dt_step), (np.float32(1.5) * dt_step, dt_step)):
        d = device_fields(sim, names, real)
        d["du"] = to_dev(sim.f["du"])
        ops.driftPositions(ctx, pkg.binding.F64 if real == np.float64 else pkg.binding.MIXED, g.view(), dt, dt_back,
                           dt_m1, to_dev(rung).view(torch.uint8), hc, ops.fields_struct(d))
        w = {k: sim.f[k].copy() for k in moved + ("ax", "ay", "az", "du")}
        ref.drift(idx, dt, dt_back, dt_m1, rung, w["x"], w["y"], w["z"], w["vx"], w["vy"], w["vz"], w["x_m1"],
                  w["y_m1"], w["z_m1"], w["ax"], w["ay"], w["az"], w["temp"], w["du"], w["du_m1"], s["mui"],
                  s["gamma"])
        check(d, w)
        assert np.array_equal(to_host(d["x_m1"]), sim.f["x_m1"]) and np.array_equal(to_host(d["du_m1"]), sim.f["du_m1"])

    for use_box in (True, False):
        d = device_fields(sim, names, real)
        d["du"] = to_dev(sim.f["du"])
        bx = sim.box if use_box else ([0, 1, 0, 1, 0, 1], [0, 0, 0])
        ops.computePositionsGroups(ctx, pkg.binding.F64 if real == np.float64 else pkg.binding.MIXED, g.view(), dt_step,
                                   dt_m1, to_dev(rung).view(torch.uint8), pkg.make_box(*bx), hc, ops.fields_struct(d))
        w = {k: sim.f[k].copy() for k in moved + ("ax", "ay", "az", "du", "h")}
        ref.positions_rung(idx, dt_step, dt_m1, rung, bx, w["x"], w["y"], w["z"], w["vx"], w["vy"], w["vz"], w["x_m1"],
                           w["y_m1"], w["z_m1"], w["ax"], w["ay"], w["az"], w["temp"], w["du"], w["du_m1"], w["h"],
                           s["mui"], s["gamma"])
        check(d, w)

    # updateSmoothingLength on the groups
    d = device_fields(sim, ["h"], real)
    ops.updateSmoothingLengthGroups(ctx, g.view(), sim.ng0, d["nc"], d["h"])
    w = sim.f["h"].copy()
    for a, e in zip(start, end):
        ref.update_h(int(a), int(e), sim.ng0, sim.nc, w)
    assert rel_err(to_host(d["h"]), w) < (1e-14 if real == np.float64 else 1e-6)
    outside = np.ones(n, bool)
    outside[idx] = False
    assert np.array_equal(to_host(d["h"])[outside], sim.f["h"][outside])


def test_scoped_operators_touch_only_active_targets(pkg, ctx, ref):
    """with a scope set, search + pair operators process the marked targets only and give them the same results"""
    import torch
    from sph_exa_b200 import ops
    from test_gpu_hydro import ref_state, device_fields, ALL
    real = np.float64
    sim, pre, pre_dt, box, s = ref_state(pkg, "sedov", 20, real, 2)
    n, ngmax = sim.n, sim.ngmax
    b = pkg.make_box(*sim.box)
    wh, whd = to_dev(sim.wh), to_dev(sim.whd)
    hc = ops.hydro_const(sim.K, wh, whd, gamma=s["gamma"], muiConst=s["mui"], Kcour=s.get("Kcour", 0.2),
                         minDt=pre_dt[0], minDt_m1=pre_dt[1])
    delim = np.arange(0, n + 32, 32).clip(max=n).astype(np.uint32)
    # whole slabs, partial slabs and single particles
    start = np.concatenate([delim[:-1][::5], [delim[3] + 5, delim[7] + 1]]).astype(np.uint32)
    end = np.concatenate([delim[1:][::5], [delim[3] + 17, delim[7] + 2]]).astype(np.uint32)
    g = ops.Groups(0, n, to_dev(start), to_dev(end))
    mask = ops.markGroups(ctx, g.view(), n)
    m = to_host(mask).astype(bool)
    want_m = np.zeros(n, bool)
    for a, e in zip(start, end):
        want_m[a:e] = True
    assert np.array_equal(m, want_m)

    tree = ops.Octree(ctx, to_dev(sim.ftree[0]), b, to_dev(sim.ftree[1]))
    d = device_fields(sim, ALL, real)
    d["du"] = to_dev(sim.f["du"])
    # unscoped run = the expected values
    nb_all, nc_all, _ = ops.findNeighbors(ctx, tree, d["x"], d["y"], d["z"], d["h"], 0, n, b, sim.ng0, ngmax, True)
    full = device_fields(sim, ALL, real)
    full["du"] = to_dev(sim.f["du"])
    full["alpha"] = to_dev(sim.alpha_pre)
    full["nc"] = nc_all
    ops.computeXMass(ctx, 0, n, full, b, hc, nb_all)
    ops.computeVeDefGradh(ctx, 0, n, full, b, hc, nb_all)
    ops.computeEOS(ctx, 0, n, full, hc)
    ops.computeIadDivvCurlv(ctx, 0, n, full, b, hc, nb_all)
    ops.computeAVswitches(ctx, 0, n, full, b, hc, nb_all)
    ops.computeMomentumEnergy(ctx, 0, n, full, b, hc, nb_all)

    outs = ["xm", "kx", "gradh", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "alpha", "ax", "ay", "az",
            "du"]
    part = device_fields(sim, ALL, real)
    part["du"] = to_dev(sim.f["du"])
    part["alpha"] = to_dev(sim.alpha_pre)
    # inputs of later operators must be complete: start from the full results, then poison the outputs one by one
    nb2 = torch.full_like(nb_all, -1)
    nc2 = torch.full_like(nc_all, 77)
    dtp = torch.full((n,), -1.0, dtype=torch.float64, device="cuda")
    gmask = torch.zeros(n, dtype=torch.uint8, device="cuda")
    with ops.scope(ctx, active=mask, searchExtFactor=1.0, dtParticle=dtp, gatherMask=gmask):
        h2 = to_dev(sim.f["h"])
        lib_find(pkg, ctx, tree, part, h2, n, b, sim, nb2, nc2)
        assert np.array_equal(to_host(nc2)[m], to_host(nc_all)[m]) and np.all(to_host(nc2)[~m] == 77)
        # the gather mask = the active targets and everything on their lists (the reference lists give the same sets)
        want_g = np.zeros(n, bool)
        want_g[m] = True
        for i in np.nonzero(m)[0]:
            want_g[sim.nb[i, :min(sim.nc[i] - 1, ngmax)]] = True
        assert np.array_equal(to_host(gmask).astype(bool), want_g)
        assert want_g.sum() < n
        part["nc"] = torch.where(mask.bool(), nc2, nc_all)
        for k in ("xm",):
            part[k] = torch.full_like(full[k], -7.0)
        ops.computeXMass(ctx, 0, n, part, b, hc, nb2)
        got = to_host(part["xm"])
        assert np.array_equal(got[m], to_host(full["xm"])[m]) and np.all(got[~m] == -7.0)
        part["xm"] = full["xm"].clone()
        for k in ("kx", "gradh"):
            part[k] = torch.full_like(full[k], -7.0)
        ops.computeVeDefGradh(ctx, 0, n, part, b, hc, nb2)
        for k in ("kx", "gradh"):
            got = to_host(part[k])
            assert np.array_equal(got[m], to_host(full[k])[m]) and np.all(got[~m] == -7.0), k
            part[k] = full[k].clone()
        ops.computeEOS(ctx, 0, n, part, hc)
        iad = ["c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv"]
        for k in iad:
            part[k] = torch.full_like(full[k], -7.0)
        ops.computeIadDivvCurlv(ctx, 0, n, part, b, hc, nb2)
        for k in iad:
            got = to_host(part[k])
            assert np.array_equal(got[m], to_host(full[k])[m]) and np.all(got[~m] == -7.0), k
            part[k] = full[k].clone()
        ops.computeAVswitches(ctx, 0, n, part, b, hc, nb2)
        got = to_host(part["alpha"])
        assert np.array_equal(got[m], to_host(full["alpha"])[m]) and np.array_equal(got[~m], sim.alpha_pre[~m])
        part["alpha"] = full["alpha"].clone()
        for k in ("ax", "ay", "az", "du"):
            part[k] = torch.full_like(full[k], -7.0)
        ops.computeMomentumEnergy(ctx, 0, n, part, b, hc, nb2)
        for k in ("ax", "ay", "az", "du"):
            got = to_host(part[k])
            assert np.array_equal(got[m], to_host(full[k])[m]) and np.all(got[~m] == -7.0), k
        dth = to_host(dtp)
        assert np.all(dth[~m] == -1.0) and np.all(dth[m] > 0)
    # the per-target Courant steps reduce to the global one
    dt_c = ops.computeMomentumEnergy(ctx, 0, n, full, b, hc, nb_all)
    with ops.scope(ctx, dtParticle=dtp):
        ops.computeMomentumEnergy(ctx, 0, n, full, b, hc, nb_all)
    assert to_host(dtp).min() == dt_c


def lib_find(pkg, ctx, tree, d, h, n, b, sim, nb, nc):
    """sphb200_find_neighbors into caller-provided (pre-poisoned) buffers"""
    import ctypes as C
    from sph_exa_b200.binding import lib, check, ptr
    tv = tree.view()
    check(lib.sphb200_find_neighbors(ctx.h, C.byref(tv), ptr(d["x"]), ptr(d["y"]), ptr(d["z"]), ptr(h),
                                     pkg.binding.F64, C.c_uint32(0), C.c_uint32(n), C.byref(b), C.c_uint(sim.ng0),
                                     C.c_uint(sim.ngmax), 1, ptr(nb), ptr(nc), None))


def run_pair(pkg, ctx, case, side, steps, **kw):
    from oracle.bdt import RefBdtSim
    from sph_exa_b200 import sim as S
    f, box, s = pkg.ic.CASES[case](side)
    if case == "noh":
        # The Noh IC starts at u = 1e-20.  Drifting a particle back to the start of its hierarchy (driftKernel) then
        # recovers u = O(rounding noise of du * dt), and the sign of that residual selects the u < 0 branch of
        # energyUpdate (positions.hpp:58-59): the reference's result is discontinuous in the last bit of du there, on
        # any two implementations.  A floor of u = 1e-10 (still 7 orders below the post-shock energies) keeps the test
        # on the continuous side.
        cv = 8.317e7 / s["mui"] / (s["gamma"] - 1.0)
        f["temp"] = np.full_like(f["temp"], 1e-10 / cv)
    rs = RefBdtSim(f, box, s, **kw)
    gs = S.HydroVeBdtProp(ctx, box, s, capacity=f["x"].size, tables=(rs.rs.wh, rs.rs.whd, rs.rs.K), **kw)
    gs.upload(f)
    return rs, gs, steps


@pytest.mark.parametrize("case,side,steps", [("sedov", 16, 14), ("noh", 16, 10)])
def test_bdt_propagator_vs_reference_sequence(pkg, ctx, ref, case, side, steps):
    """HydroVeBdtProp over several hierarchies: rung tables exact, fields to 1e-10 per substep"""
    rs, gs, steps = run_pair(pkg, ctx, case, side, steps)
    seen_rungs, seen_partial = set(), 0
    names = ["x", "y", "z", "h", "vx", "vy", "vz", "x_m1", "temp", "du_m1", "alpha"]
    for step in range(steps):
        was_synced = gs.isSynced()
        rs.step()
        gs.step()
        ts, info = rs.ts, gs.timestep()
        assert info.numRungs == ts.numRungs and info.substep == ts.substep, step
        assert list(info.rungRanges) == list(ts.rungRanges), step
        assert abs(info.nextDt - ts.nextDt) <= 1e-6 * ts.nextDt
        assert np.allclose(np.array(info.dt_m1[:ts.numRungs]), ts.dt_m1[:ts.numRungs], rtol=1e-6)
        assert np.allclose(np.array(info.dt_drift[:ts.numRungs]), ts.dt_drift[:ts.numRungs], rtol=1e-6)
        assert info.numActiveGroups == rs.active[0].size
        gstart, gend = gs.tsGroups()
        assert np.array_equal(gstart, rs.ts_groups[0]) and np.array_equal(gend, rs.ts_groups[1]), step
        assert np.array_equal(gs.rungs(), rs.rung), step
        out = gs.download(names=names + ["nc"], with_keys=True)
        if was_synced:
            assert np.array_equal(out["keys"], rs.rs.keys)
        active = np.zeros(rs.n, bool)
        for a, e in zip(rs.active[0], rs.active[1]):
            active[a:e] = True
        assert np.array_equal(out["nc"][active], rs.nc[active]), step
        for k in names:
            assert rel_err(out[k], rs.rs.f[k]) < 1e-9, (step, k)
        seen_rungs.add(ts.numRungs)
        seen_partial += not was_synced
        assert abs(gs.minDt - rs.rs.minDt) <= 1e-6 * rs.rs.minDt and abs(gs.ttot - rs.rs.ttot) <= 1e-6 * rs.rs.ttot
    assert max(seen_rungs) >= 2 and seen_partial >= 3, (seen_rungs, seen_partial)
    gs.close()


@pytest.mark.parametrize("case,side,safety", [("sedov", 16, 100), ("kh", 24, 0)])
def test_bdt_single_rung(pkg, ctx, ref, case, side, safety):
    """numRungs = 1 (forced by safetySteps on Sedov, natural on the KH slab): every step is a full sync of all
    groups; conserved totals stay put"""
    rs, gs, steps = run_pair(pkg, ctx, case, side, 4, safety_steps=safety)
    e0 = None
    for step in range(steps):
        rs.step()
        gs.step()
        info = gs.timestep()
        assert info.numRungs == 1 and info.isSynced and info.numActiveGroups == info.numGroups
        out = gs.download(names=["x", "temp", "vx"])
        for k in out:
            assert rel_err(out[k], rs.rs.f[k]) < 1e-9, (step, k)
        c = gs.conserved()
        e0 = e0 or c["etot"]
        assert abs(c["etot"] - e0) < 1e-4 * abs(e0)
    gs.close()


def test_turb_ve_bdt_propagator(pkg, ctx, ref):
    """TurbVeBdtProp (main/src/propagator/turb_ve.hpp:52-96): block time steps with the stirring applied to the active
    rungs after the acceleration time step was taken; Sedov IC so that several rungs and partial syncs occur"""
    from oracle.bdt import RefBdtSim
    from sph_exa_b200 import sim as S
    from sph_exa_b200.turbulence import Turbulence
    f, box, s = pkg.ic.sedov(16)
    rs = RefBdtSim(f, box, s)
    rs.turb = ref.RefTurbulence()
    gt = Turbulence()
    gs = S.HydroVeBdtProp(ctx, box, s, capacity=f["x"].size, tables=(rs.rs.wh, rs.rs.whd, rs.rs.K))
    gs.attachTurbulence(gt)
    gs.upload(f)
    partial = 0
    for step in range(10):
        partial += not gs.isSynced()
        rs.step()
        gs.step()
        info = gs.timestep()
        assert info.numRungs == rs.ts.numRungs and list(info.rungRanges) == list(rs.ts.rungRanges), step
        out = gs.download(names=["x", "y", "z", "vx", "vy", "vz", "temp", "ax", "ay", "az"])
        acc = max(np.abs(rs.rs.f[k]).max() for k in ("ax", "ay", "az"))
        for k in ("ax", "ay", "az"):
            assert np.abs(out[k] - rs.rs.f[k]).max() <= 1e-9 * acc, (step, k)
        for k in ("x", "y", "z", "vx", "vy", "vz", "temp"):
            assert rel_err(out[k], rs.rs.f[k]) < 1e-9, (step, k)
    assert partial >= 5
    # the OU process advanced once per substep on both sides
    st_g, st_r = gt.state(), rs.turb.state()
    assert np.array_equal(st_g["phases"], st_r["phases"])
    gs.close()
    gt.close()
    rs.turb.close()
