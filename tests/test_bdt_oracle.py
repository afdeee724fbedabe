"""CPU checks of the block-time-step oracle (oracle/bdt.py, the drift/advance veneer in oracle/ref_wrap.cpp).

What pins it: ref.positions_rung == the reference's own positionUpdate/energyUpdate host path (ref.positions, itself
pinned by sph/test and the 3-step goldens) when all rungs are 0; drift obeys the time-reversibility the reference
documents for positionUpdate (sph/include/sph/positions.hpp:74-75); the all-pairs search equals cstone::findNeighbors
(findneighbors.hpp:95-165); butterfly / activeRung sequences follow cstone/primitives/math.hpp:26-31 and
ve_hydro_bdt.hpp:106-110; the time-step hierarchy (rungTimestep, minimumGroupDt, findRungRanges) equals the reference's
own sph/ts_rungs.hpp run on host memory (oracle/ref_bdt.cpp).
"""
import numpy as np
import pytest


def small_state(pkg, ref, case="sedov", side=10, steps=2):
    from oracle.ref_pipeline import RefSim
    f, box, s = pkg.ic.CASES[case](side)
    sim = RefSim(f, box, s)
    for _ in range(steps):
        sim.step()
    sim.computeForces()
    return sim, s


def test_butterfly_active_rung_sequences():
    from oracle import bdt
    assert [bdt.butterfly(i) for i in range(9)] == [0, 1, 2, 1, 3, 1, 2, 1, 4]
    # rung hierarchy of 4 rungs: 8 substeps, rung r is due every 2^r substeps
    assert [bdt.active_rung(k, 4) for k in range(9)] == [0, 1, 2, 1, 3, 1, 2, 1, 0]
    assert [bdt.active_rung(k, 1) for k in range(3)] == [0, 0, 0]
    assert [bdt.active_rung(k, 2) for k in range(4)] == [0, 1, 0, 0]


def test_find_rung_ranges_is_lower_bound():
    from oracle import bdt
    dt = np.array([1, 1, 1.5, 2, 2, 3.9, 4, 7.9, 8, 100], np.float32) * np.float32(1e-5)
    assert bdt.find_rung_ranges(dt[0], dt, 4) == [0, 3, 6, 8, 10]
    assert bdt.find_rung_ranges(dt[0], dt, 2) == [0, 3, 10, 10, 10]
    assert bdt.find_rung_ranges(dt[0], dt, 1) == [0, 10, 10, 10, 10]


def test_find_rung_ranges_equals_reference(ref):
    from oracle import bdt
    rng = np.random.Generator(np.random.MT19937(5))
    for n in (1, 2, 33, 1000):
        dt = np.sort(np.exp(rng.uniform(np.log(1e-6), np.log(1e-4), n)).astype(np.float32))
        for num_rungs in (1, 2, 3, 4):
            assert bdt.find_rung_ranges(dt[0], dt, num_rungs) == ref.find_rung_ranges(dt[0], dt, num_rungs)


def test_time_step_hierarchy_equals_reference_ts_rungs(ref):
    """rung_timestep / minimum_group_dt of the oracle against sph::rungTimestep / sph::minimumGroupDt themselves
    (ts_rungs.hpp:131-157, compiled with host stand-ins for the five device primitives they call)"""
    from oracle import bdt
    rng = np.random.Generator(np.random.MT19937(6))
    seen_rungs = set()
    for trial in range(60):
        n = int(rng.integers(1, 3000))
        spread = rng.choice([1.5, 3.0, 10.0, 100.0])
        dt = np.exp(rng.uniform(np.log(1e-6), np.log(1e-6 * spread), n)).astype(np.float32)
        if n > 10:
            dt[rng.integers(0, n, n // 8)] = dt[1]  # ties
        max_dt = np.float32(rng.choice([1.0, float(dt.min()) * 0.7]))
        w_dt, w_idx, w_next, w_tot, w_rungs, w_rr = ref.rung_timestep(dt, max_dt)
        gd = dt.copy()
        idx, ts = bdt.rung_timestep(gd, n, max_dt)
        assert np.array_equal(gd, w_dt) and np.array_equal(idx, w_idx)
        assert (ts.nextDt, ts.totDt, ts.numRungs, ts.rungRanges) == (w_next, w_tot, w_rungs, w_rr), trial
        seen_rungs.add(ts.numRungs)
        # a substep inside the hierarchy: the groups of the rungs that were due got new time steps
        for substep in range(1, 1 << (ts.numRungs - 1)):
            high = bdt.active_rung(substep, ts.numRungs)
            ng = ts.rungRanges[high]
            if ng == 0:
                continue
            gd2 = gd.copy()
            gd2[:ng] = (gd2[:ng] * rng.uniform(0.5, 1.5, ng).astype(np.float32))[rng.permutation(ng)]
            t2 = ts.copy()
            t2.substep, t2.elapsedDt = substep, np.float32(ts.nextDt * np.float32(substep))
            w_gd, w_idx2, w_dt2, w_rr2 = ref.minimum_group_dt(t2.totDt, t2.elapsedDt, t2.numRungs, substep,
                                                              ts.rungRanges, gd2, ng)
            idx2, dt2, rr2 = bdt.minimum_group_dt(t2, gd2, ng, ts.rungRanges[4])
            assert np.array_equal(gd2, w_gd) and np.array_equal(idx2, w_idx2[:idx2.size])
            assert dt2 == w_dt2 and rr2 == w_rr2, (trial, substep)
    assert seen_rungs == {1, 2, 3, 4}


def test_positions_rung_equals_reference_host_path(pkg, ref):
    sim, s = small_state(pkg, ref)
    n = sim.n
    names = ("x", "y", "z", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "ax", "ay", "az", "temp", "du", "du_m1")
    a = {k: sim.f[k].copy() for k in names + ("h",)}
    b = {k: sim.f[k].copy() for k in names + ("h",)}
    dt, dt_m1 = np.float32(0.8 * sim.minDt), np.float32(sim.minDt)
    ref.positions(0, n, float(dt), float(dt_m1), sim.box, *[a[k] for k in names], a["h"], s["mui"], s["gamma"])
    ref.positions_rung(np.arange(n, dtype=np.uint32), dt, [dt_m1] * 4, np.zeros(n, np.uint8), sim.box,
                       *[b[k] for k in names], b["h"], s["mui"], s["gamma"])
    for k in names:
        assert np.array_equal(a[k], b[k]), k
    # per-particle rung: particle i uses dt_m1[rung[i]]
    rung = (np.arange(n) % 4).astype(np.uint8)
    dts = np.array([1.0, 0.5, 2.0, 4.0], np.float32) * dt_m1
    c = {k: sim.f[k].copy() for k in names + ("h",)}
    ref.positions_rung(np.arange(n, dtype=np.uint32), dt, dts, rung, sim.box, *[c[k] for k in names], c["h"], s["mui"],
                       s["gamma"])
    for r in range(4):
        d = {k: sim.f[k].copy() for k in names + ("h",)}
        ref.positions(0, n, float(dt), float(dts[r]), sim.box, *[d[k] for k in names], d["h"], s["mui"], s["gamma"])
        sel = rung == r
        for k in ("x", "vx", "x_m1", "temp"):
            assert np.array_equal(c[k][sel], d[k][sel]), (r, k)


def test_drift_is_time_reversible_and_leaves_history(pkg, ref):
    sim, s = small_state(pkg, ref)
    n = sim.n
    names = ("x", "y", "z", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "ax", "ay", "az", "temp", "du", "du_m1")
    a = {k: sim.f[k].copy() for k in names}
    idx = np.arange(0, n, 3, dtype=np.uint32)
    rung = np.zeros(n, np.uint8)
    dt, dt_m1 = np.float32(0.9 * sim.minDt), [np.float32(sim.minDt)] * 4
    ref.drift(idx, dt, 0.0, dt_m1, rung, *[a[k] for k in names], s["mui"], s["gamma"])
    moved = np.abs(a["temp"][idx] - sim.f["temp"][idx]).max()
    assert moved > 0
    for k in ("x_m1", "y_m1", "z_m1", "du_m1", "ax", "du"):
        assert np.array_equal(a[k], sim.f[k]), k
    rest = np.ones(n, bool)
    rest[idx] = False
    for k in ("x", "vx", "temp"):
        assert np.array_equal(a[k][rest], sim.f[k][rest]), k
    # drifting back by the same dt recovers the state at which the accelerations were evaluated
    ref.drift(idx, 0.0, dt, dt_m1, rung, *[a[k] for k in names], s["mui"], s["gamma"])
    assert np.abs(a["x"] - sim.f["x"]).max() < 1e-15
    assert np.abs(a["temp"] - sim.f["temp"]).max() <= 1e-12 * np.abs(sim.f["temp"]).max()
    # drift(dt + e, dt) == drift(e, 0) applied to the recovered state: one call does both legs
    b = {k: sim.f[k].copy() for k in names}
    c = {k: sim.f[k].copy() for k in names}
    e = np.float32(0.25 * sim.minDt)
    ref.drift(idx, dt, 0.0, dt_m1, rung, *[b[k] for k in names], s["mui"], s["gamma"])
    ref.drift(idx, np.float32(dt + e), dt, dt_m1, rung, *[b[k] for k in names], s["mui"], s["gamma"])
    ref.drift(idx, np.float32(dt + e), 0.0, dt_m1, rung, *[c[k] for k in names], s["mui"], s["gamma"])
    assert np.abs(b["x"] - c["x"]).max() < 1e-15
    assert np.abs(b["temp"] - c["temp"]).max() <= 1e-12 * np.abs(c["temp"]).max()


def test_all_pairs_search_equals_reference_search(pkg, ref):
    from oracle import bdt
    sim, s = small_state(pkg, ref, "kh", 16, 1)
    f, n = sim.f, sim.n
    h0 = f["h"].copy() * 1.3  # force h iterations for some particles
    h1, h2 = h0.copy(), h0.copy()
    nb, nc, h1 = ref.find_neighbors(f["x"], f["y"], f["z"], h1, 0, n, sim.box, sim.tree, sim.layout, sim.centers,
                                    sim.sizes, sim.ng0, sim.ngmax, True)
    targets = np.arange(0, n, 7)
    lists, ncd = bdt.brute_force_neighbors(f["x"], f["y"], f["z"], h2, targets, sim.box, sim.ng0, sim.ngmax, True)
    nb = nb.reshape(n, sim.ngmax)
    for i in targets:
        assert ncd[int(i)] == nc[i]
        assert np.array_equal(lists[int(i)], np.sort(nb[i, :min(nc[i] - 1, sim.ngmax)]))
    assert np.array_equal(h2[targets], h1[targets])


def test_reference_sequence_hierarchy_invariants(pkg, ref):
    """RefBdtSim on Sedov 12^3: every hierarchy has 2^(numRungs-1) substeps, all rungs are advanced (nothing left
    drifted) when it ends, a rung-r group is advanced every 2^r substeps with dt_m1[r] = the time since its last
    advance, and the energy stays conserved."""
    from oracle import bdt
    f, box, s = pkg.ic.sedov(12)
    rs = bdt.RefBdtSim(f, box, s)
    tot0 = pkg.ic.conserved_totals(rs.rs.f, s)["etot"]
    hierarchies, t_sum = 0, 0.0
    for step in range(20):
        substep = rs.ts.substep if step else 0
        rs.step()
        ts = rs.ts
        t_sum += float(ts.nextDt)
        assert ts.rungRanges[0] == 0 and ts.rungRanges[bdt.MAX_RUNGS] == rs.groups[0].size
        assert all(ts.rungRanges[r] <= ts.rungRanges[r + 1] for r in range(bdt.MAX_RUNGS))
        if bdt.active_rung(ts.substep, ts.numRungs) == 0:  # hierarchy complete
            hierarchies += 1
            assert ts.substep == 1 << (ts.numRungs - 1)
            assert np.all(ts.dt_drift[:ts.numRungs] == 0)
            occupied = [r for r in range(ts.numRungs) if ts.rungRanges[r + 1] > ts.rungRanges[r]]
            for r in occupied:
                assert abs(ts.dt_m1[r] - (1 << r) * ts.nextDt) <= 1e-5 * ts.dt_m1[r] * (1 << r)
            # every particle carries the rung of the group it was advanced with
            for r in range(ts.numRungs):
                for a, e in zip(rs.rungs[r][0], rs.rungs[r][1]):
                    assert np.all(rs.rung[a:e] == r)
    assert hierarchies >= 2
    assert abs(rs.rs.ttot - t_sum) < 1e-12
    tot = pkg.ic.conserved_totals(rs.rs.f, s)["etot"]
    assert abs(tot - tot0) < 2e-3 * abs(tot0)


def test_product_host_logic_matches_oracle(pkg):
    """the host-side rung arithmetic of the distributed propagator (sph-exa_b200/dist_bdt.py) against the oracle's"""
    from oracle import bdt
    from sph_exa_b200 import dist_bdt
    for i in range(64):
        assert dist_bdt.butterfly(i) == bdt.butterfly(i)
        for nr in range(1, 5):
            assert dist_bdt.active_rung(i, nr) == bdt.active_rung(i, nr)
    a, b = dist_bdt.Timestep(), bdt.Timestep()
    assert a.numRungs == b.numRungs and a.substep == b.substep and list(a.rungRanges) == list(b.rungRanges)
    c = a.copy()
    c.dt_m1[0] = 1
    c.rungRanges[1] = 5
    assert a.dt_m1[0] == 0 and a.rungRanges[1] == 0
