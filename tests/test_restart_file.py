"""CPU tests of the restart container (csrc/checkpoint.cu is host code: no device needed).

The attribute names are the reference's (sph/include/sph/particles_data.hpp:169-189, cstone/sfc/box.hpp:170-171,
hydro_turb/turbulence_data.hpp:92-115); the per-rank slicing follows partitionRange
(main/src/io/ifile_io_hdf5.cpp:131-148)."""
import ctypes as C
import os

import numpy as np
import pytest


def make_fields(B, n, real, seed=0):
    rng = np.random.Generator(np.random.MT19937(seed))
    out = {}
    for name in B.CONSERVED:
        dt = np.float64 if name in B.ALWAYS_F64 else real
        out[name] = rng.normal(size=n).astype(dt)
    return out


def write_file(pkg, path, arrays, dtype, iteration=7, n_global=None, rank=0, num_ranks=1, turb=None,
               box=([-0.5, 0.5, 0.0, 2.0, -1.0, 1.0], [1, 0, 1])):
    B = pkg.binding
    n = len(arrays["x"])
    info = B.RestartInfo()
    info.iteration, info.numParticlesGlobal, info.numParticles = iteration, n_global or n, n
    info.dtype = dtype
    info.time, info.minDt, info.minDt_m1 = 0.125, 1e-4, 0.5e-4
    info.box = pkg.make_box(*box)
    hc = B.HydroConst()
    hc.K, hc.ng0, hc.ngmax = 0.79, 100, 150
    hc.Kcour, hc.Krho, hc.gamma, hc.muiConst = 0.2, 0.06, 5.0 / 3.0, 10.0
    hc.alphamin, hc.alphamax, hc.decay_constant = 0.05, 1.0, 0.2
    hc.Atmin, hc.Atmax, hc.ramp = 0.1, 0.2, 10.0
    f = B.Fields()
    for name in B.FIELD_NAMES:
        setattr(f, name, None)
    for name, a in arrays.items():
        setattr(f, name, a.ctypes.data)
    B.check(B.lib.sphb200_restart_write(path.encode(), C.byref(info), C.byref(hc), turb, rank, num_ranks, C.byref(f)))


@pytest.mark.parametrize("real", [np.float64, np.float32])
@pytest.mark.parametrize("n", [0, 1, 1001])
def test_round_trip(pkg, tmp_path, real, n):
    B = pkg.binding
    from sph_exa_b200 import restart
    dtype = B.F64 if real == np.float64 else B.MIXED
    arrays = make_fields(B, n, real)
    p = str(tmp_path / "dump.bin")
    write_file(pkg, p, arrays, dtype)
    assert not os.path.exists(p + ".part")
    with restart.RestartFile(p) as f:
        assert (f.info.iteration, f.info.numParticlesGlobal, f.info.numParticles) == (7, n, n)
        assert f.info.dtype == dtype and f.info.hasTurbulence == 0
        assert (f.info.time, f.info.minDt, f.info.minDt_m1) == (0.125, 1e-4, 0.5e-4)
        assert list(f.info.box.lim) == [-0.5, 0.5, 0.0, 2.0, -1.0, 1.0]
        assert list(f.info.box.bc) == [1, 0, 1] and list(f.info.box.len) == [1.0, 2.0, 2.0]
        # the reference's attribute names
        assert f.attribute("gamma")[0] == 5.0 / 3.0
        assert f.attribute("ngmax", np.int64)[0] == 150
        assert f.attribute("gravConstant")[0] == 0.0
        assert list(f.attribute("boundaryType", np.uint8)) == [1, 0, 1]
        assert f.attribute("no such attribute") is None
        got = {k: np.full(n, np.nan, v.dtype) for k, v in arrays.items()}
        f.read_fields(0, n, dtype, got)
        for k in arrays:
            assert np.array_equal(got[k], arrays[k]), k


def test_slices_and_precision_conversion(pkg, tmp_path):
    """any [first, first+count) can be read, at either precision (fileutils readField converts on the fly)"""
    B = pkg.binding
    from sph_exa_b200 import restart
    n = 517
    arrays = make_fields(B, n, np.float64, seed=3)
    p = str(tmp_path / "d.bin")
    write_file(pkg, p, arrays, B.F64)
    with restart.RestartFile(p) as f:
        got = {k: np.zeros(100, np.float64 if k in B.ALWAYS_F64 else np.float32) for k in arrays}
        f.read_fields(400, 100, B.MIXED, got)
        for k in arrays:
            want = arrays[k][400:500].astype(got[k].dtype)
            assert np.array_equal(got[k], want), k
        with pytest.raises(pkg.SphError, match="beyond the end"):
            f.read_fields(500, 100, B.MIXED, got)
    # f32 file read as f64
    arrays32 = make_fields(B, n, np.float32, seed=4)
    write_file(pkg, p, arrays32, B.MIXED)
    with restart.RestartFile(p) as f:
        got = {k: np.zeros(n, np.float64) for k in arrays32}
        f.read_fields(0, n, B.F64, got)
        for k in arrays32:
            assert np.array_equal(got[k], arrays32[k].astype(np.float64)), k


def test_partition_range():
    """same split as the reference reader: contiguous, exhaustive, first R % N ranks one longer"""
    import sphexa_b200_loader
    sphexa_b200_loader.load()
    from sph_exa_b200 import restart
    for total in (0, 1, 7, 100, 1003):
        for P in (1, 2, 3, 8):
            pos = 0
            for r in range(P):
                first, count = restart.slice_for_rank(total, r, P)
                assert first == pos
                assert count == total // P + (1 if r < total % P else 0)
                pos += count
            assert pos == total


def test_rejects_foreign_and_truncated_files(pkg, tmp_path):
    B = pkg.binding
    from sph_exa_b200 import restart
    p = str(tmp_path / "x.bin")
    with pytest.raises(pkg.SphError, match="cannot open"):
        restart.RestartFile(p)
    open(p, "wb").write(b"not a restart file at all, just bytes")
    with pytest.raises(pkg.SphError, match="not a restart file"):
        restart.RestartFile(p)
    write_file(pkg, p, make_fields(B, 300, np.float64), B.F64)
    data = open(p, "rb").read()
    open(p, "wb").write(data[:len(data) - 64])
    with pytest.raises(pkg.SphError, match="truncated"):
        restart.RestartFile(p)
    open(p, "wb").write(data[:40])
    with pytest.raises(pkg.SphError, match="truncated"):
        restart.RestartFile(p)


@pytest.mark.gpu
def test_turbulence_state_in_file(pkg, ctx, tmp_path):
    """(needs the device: the turbulence object keeps its mode table there)  OU phases and the text form of the random engine travel with the dump, under the reference's names"""
    B = pkg.binding
    from sph_exa_b200 import restart, turbulence
    t = turbulence.Turbulence(**turbulence.DEFAULTS)
    p = str(tmp_path / "t.bin")
    write_file(pkg, p, make_fields(B, 10, np.float64), B.F64, turb=t.h)
    with restart.RestartFile(p) as f:
        assert f.info.hasTurbulence == 1
        nm = int(f.attribute("turbulence::numModes", np.int64)[0])
        assert nm == t.numModes
        st = t.state()
        assert np.array_equal(f.attribute("turbulence::phases"), st["phases"])
        assert np.array_equal(f.attribute("turbulence::modes"), st["modes"].ravel())
        assert np.array_equal(f.attribute("turbulence::amplitudes"), st["amplitudes"])
        assert f.attribute("turbulence::variance")[0] == st["variance"]
        rng = bytes(f.attribute("rngEngineState", np.uint8)).rstrip(b"\0").decode()
        assert rng == t.rngState()


def test_block_time_step_state_round_trip(pkg, tmp_path):
    """ts::numRungs, ts::dt_m1 (Timestep::loadOrStore, sph/include/sph/timestep.h:29-33) and the byte field "rung"
    (conserved field of HydroVeBdtProp, ve_hydro_bdt.hpp:94)"""
    B = pkg.binding
    from sph_exa_b200 import restart

    n = 777
    arrays = make_fields(B, n, np.float64)
    rung = (np.arange(n) % 4).astype(np.uint8)
    info = B.RestartInfo()
    info.iteration, info.numParticlesGlobal, info.numParticles, info.dtype = 3, n, n, B.F64
    info.box = pkg.make_box([0, 1, 0, 1, 0, 1], [1, 1, 1])
    hc = B.HydroConst()
    f = B.Fields()
    for name in B.FIELD_NAMES:
        setattr(f, name, None)
    for name, a in arrays.items():
        setattr(f, name, a.ctypes.data)
    bdt = B.RestartBdt(3, (C.c_float * 4)(1e-3, 2e-3, 4e-3, 0.0), rung.ctypes.data)
    p = str(tmp_path / "bdt.bin")
    B.check(B.lib.sphb200_restart_write_bdt(p.encode(), C.byref(info), C.byref(hc), None, 0, 1, C.byref(f),
                                            C.byref(bdt)))
    with restart.RestartFile(p) as rf:
        assert rf.attribute("ts::numRungs", np.int64)[0] == 3
        assert list(rf.attribute("ts::dt_m1", np.float32)) == list(np.float32([1e-3, 2e-3, 4e-3, 0.0]))
        assert rf.has_field("rung")
        got = np.full(100, 255, np.uint8)
        rf.read_bytes("rung", 600, 100, got)
        assert np.array_equal(got, rung[600:700])
        with pytest.raises(B.SphError):
            rf.read_bytes("rung", 700, 100, got)
    # a dump of the global-time-step propagator has no rung field
    q = str(tmp_path / "plain.bin")
    write_file(pkg, q, arrays, B.F64)
    with restart.RestartFile(q) as rf:
        assert not rf.has_field("rung")
        assert rf.attribute("ts::numRungs", np.int64) is None


def test_find_files_follows_the_rank_count_of_the_dump(pkg, tmp_path):
    """left-over files of an earlier dump with more ranks are ignored, a dump with missing rank files is refused"""
    B = pkg.binding
    from sph_exa_b200 import restart
    base = str(tmp_path / "dump.bin")
    arrays = make_fields(B, 10, np.float64)
    # an old 3-rank dump, then a 2-rank dump under the same name
    for r in range(3):
        write_file(pkg, f"{base}.{r}", arrays, B.F64, n_global=30, rank=r, num_ranks=3)
    assert restart.find_files(base) == [f"{base}.{r}" for r in range(3)]
    for r in range(2):
        write_file(pkg, f"{base}.{r}", arrays, B.F64, iteration=9, n_global=20, rank=r, num_ranks=2)
    assert restart.find_files(base) == [f"{base}.0", f"{base}.1"]
    os.remove(f"{base}.1")
    os.remove(f"{base}.2")
    with pytest.raises(B.SphError):
        restart.find_files(base)
    # a single-file dump is found under its own name
    write_file(pkg, base, arrays, B.F64)
    assert restart.find_files(base) == [base]
