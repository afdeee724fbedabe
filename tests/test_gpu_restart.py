"""Restart parity on the GPU: a run that is dumped and reloaded continues bit-identically.

Reference behaviour: `sphexa --init dump.h5:k` resumes from step k of a dump written by the main loop
(main/src/sphexa/sphexa.cpp:232-246, main/src/init/file_init.hpp:93-140); conserved fields, time-step history, box and
the turbulence driver state are what travels (ve_hydro.hpp:75, particles_data.hpp:142-193, turbulence_data.hpp:90-125).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run_and_compare(pkg, ctx, tmp_path, f, box, s, dtype, before, after, turb_factory=None, tag="dump"):
    from sph_exa_b200 import sim as S, restart, binding as B
    n = f["x"].size
    a = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    ta = turb_factory() if turb_factory else None
    if ta:
        a.attachTurbulence(ta)
    a.upload(f)
    for _ in range(before):
        a.step()
    path = str(tmp_path / tag)
    restart.save(a, path, iteration=before, num_particles_global=n)
    for _ in range(after):
        a.step()
    want = a.download(with_keys=True)

    b = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    tb = turb_factory() if turb_factory else None
    if tb:
        b.attachTurbulence(tb)
    it, n_glob = restart.load(b, path)
    assert (it, n_glob) == (before, n)
    assert b.ttot > 0
    for _ in range(after):
        b.step()
    got = b.download(with_keys=True)
    for k in list(B.CONSERVED) + ["keys"]:
        assert np.array_equal(got[k], want[k]), k
    assert (b.minDt, b.ttot) == (a.minDt, a.ttot)
    if ta:
        assert np.array_equal(ta.state()["phases"], tb.state()["phases"])
        assert ta.rngState() == tb.rngState()
        ta.close()
        tb.close()
    a.close()
    b.close()


@pytest.mark.parametrize("real", [np.float64, np.float32])
def test_sedov_restart_continues_bit_identically(pkg, ctx, tmp_path, real):
    dtype = pkg.F64 if real == np.float64 else pkg.MIXED
    f, box, s = pkg.ic.sedov(24, "f64" if real == np.float64 else "mixed")
    run_and_compare(pkg, ctx, tmp_path, f, box, s, dtype, before=3, after=3)


def test_open_box_restart(pkg, ctx, tmp_path):
    """Noh: open boundaries, the global box follows the particles and is part of the dump (box.hpp:168-175)"""
    f, box, s = pkg.ic.noh(20)
    run_and_compare(pkg, ctx, tmp_path, f, box, s, pkg.F64, before=2, after=2)


def test_turbulence_restart_carries_the_driver_state(pkg, ctx, tmp_path):
    """stirred run: the OU phases and the random engine resume where the dump left them"""
    from sph_exa_b200.turbulence import Turbulence
    f, box, s = pkg.ic.turbulence(20)
    run_and_compare(pkg, ctx, tmp_path, f, box, s, pkg.F64, before=2, after=3, turb_factory=Turbulence)


def test_turbulence_restart_refuses_a_driver_with_other_constants(pkg, ctx, tmp_path):
    """same number of modes, another Mach velocity: the forcing would resume silently wrong
    (TurbulenceData::loadOrStore reloads the constants from the file, turbulence_data.hpp:92-115)"""
    from sph_exa_b200 import sim as S, restart
    from sph_exa_b200.turbulence import Turbulence
    f, box, s = pkg.ic.turbulence(16)
    n = f["x"].size
    a = S.HydroVeProp(ctx, box, s, capacity=n)
    ta = Turbulence()
    a.attachTurbulence(ta)
    a.upload(f)
    a.step()
    path = str(tmp_path / "turb")
    restart.save(a, path, iteration=1, num_particles_global=n)
    b = S.HydroVeProp(ctx, box, s, capacity=n)
    tb = Turbulence(stMachVelocity=0.123)
    assert tb.numModes == ta.numModes
    b.attachTurbulence(tb)
    with pytest.raises(pkg.SphError, match="differs from the file"):
        restart.load(b, path)
    for o in (a, b, ta, tb):
        o.close()


def test_restart_refuses_another_avclean_setting(pkg, ctx, tmp_path):
    """avClean selects other kernels and gradient fields; a dump of one setting must not resume under the other"""
    from sph_exa_b200 import sim as S, restart
    f, box, s = pkg.ic.sedov(12)
    n = f["x"].size
    a = S.HydroVeProp(ctx, box, s, capacity=n, av_clean=False)
    a.upload(f)
    a.step()
    path = str(tmp_path / "noclean")
    restart.save(a, path, iteration=1, num_particles_global=n)
    b = S.HydroVeProp(ctx, box, s, capacity=n, av_clean=True)
    with pytest.raises(pkg.SphError, match="avClean"):
        restart.load(b, path)
    c = S.HydroVeProp(ctx, box, s, capacity=n, av_clean=False)
    assert restart.load(c, path) == (1, n)
    for o in (a, b, c):
        o.close()


def test_restart_into_other_precision(pkg, ctx, tmp_path):
    """an f64 dump starts a mixed-precision run: the reader converts, and the result equals a mixed run started from
    the rounded state"""
    from sph_exa_b200 import sim as S, restart, binding as B
    f, box, s = pkg.ic.sedov(20, "f64")
    n = f["x"].size
    a = S.HydroVeProp(ctx, box, s, capacity=n, dtype=pkg.F64)
    a.upload(f)
    a.step()
    path = str(tmp_path / "d64")
    restart.save(a, path, iteration=1, num_particles_global=n)
    state = a.download()
    a.close()

    b = S.HydroVeProp(ctx, box, s, capacity=n, dtype=pkg.MIXED)
    restart.load(b, path)
    b.step()
    got = b.download()
    b.close()

    with restart.RestartFile(path) as rf:
        info = rf.info
        s2 = dict(s, minDt=info.minDt, minDt_m1=info.minDt_m1)
    c = S.HydroVeProp(ctx, box, s2, capacity=n, dtype=pkg.MIXED)
    c.upload({k: (v if k in B.ALWAYS_F64 else v.astype(np.float32)) for k, v in state.items()})
    c.step()
    want = c.download()
    c.close()
    for k in B.CONSERVED:
        assert np.array_equal(got[k], want[k]), k


def test_block_time_step_restart(pkg, ctx, tmp_path):
    """HydroVeBdtProp::save / load (ve_hydro_bdt.hpp:153-168): a dump written where the hierarchy is synced carries
    ts::numRungs, ts::dt_m1 and the conserved field "rung"; the reloaded run starts a new hierarchy from them."""
    from sph_exa_b200 import sim as S, restart, binding as B
    f, box, s = pkg.ic.sedov(16, "f64")
    n = f["x"].size
    a = S.HydroVeBdtProp(ctx, box, s, capacity=n, dtype=pkg.F64)
    a.upload(f)
    # run until at least two hierarchies with more than one rung are complete
    hierarchies, steps = 0, 0
    while hierarchies < 3 or not a.isSynced():
        a.step()
        steps += 1
        if a.isSynced():
            hierarchies += 1
        assert steps < 200
    ts = a.timestep()
    assert ts.numRungs > 1
    path = str(tmp_path / "bdt")
    restart.save(a, path, iteration=steps, num_particles_global=n)
    rung_a = a.rungs()
    with restart.RestartFile(path) as rf:
        assert rf.attribute("ts::numRungs", np.int64)[0] == ts.numRungs
        assert list(rf.attribute("ts::dt_m1", np.float32)) == list(ts.dt_m1)
        got = np.empty(n, np.uint8)
        rf.read_bytes("rung", 0, n, got)
        assert np.array_equal(got, rung_a)

    def advance(p):
        # one full hierarchy
        p.step()
        k = 1
        while not p.isSynced():
            p.step()
            k += 1
        return k

    # a dump from the middle of a hierarchy would lose substep / drift state: refused
    a.step()
    ka = 1
    assert not a.isSynced()
    with pytest.raises(B.SphError, match="synced"):
        restart.save(a, str(tmp_path / "mid"), iteration=steps + 1, num_particles_global=n)
    while not a.isSynced():
        a.step()
        ka += 1
    want = a.download(with_keys=True)

    b = S.HydroVeBdtProp(ctx, box, s, capacity=n, dtype=pkg.F64)
    it, n_glob = restart.load(b, path)
    assert (it, n_glob) == (steps, n)
    tb = b.timestep()
    assert (tb.numRungs, list(tb.dt_m1), tb.substep) == (ts.numRungs, list(ts.dt_m1), 0)
    kb = advance(b)
    assert ka == kb
    got = b.download(with_keys=True)
    for k in list(B.CONSERVED) + ["keys"]:
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(b.rungs(), a.rungs())
    ta, tb = a.timestep(), b.timestep()
    assert (ta.numRungs, list(ta.dt_m1), ta.totDt) == (tb.numRungs, list(tb.dt_m1), tb.totDt)
    assert (b.minDt, b.ttot) == (a.minDt, a.ttot)
    a.close()
    b.close()
