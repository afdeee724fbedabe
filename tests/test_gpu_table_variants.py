"""The two kernel-table variants of the pair operators (table through L1 / table staged in shared memory with bulk
async copies, hydro.cu) must return the same bits: which one runs is decided from timings, so results may not
depend on it."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FIELDS = ["x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "du", "ax", "ay", "az", "xm", "kx", "gradh", "c11",
          "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "nc"]


def run(pkg, case, side, real, variant, av_clean, steps=3):
    from sph_exa_b200 import sim as S
    f, box, s = pkg.ic.CASES[case](side, "f64" if real == np.float64 else "mixed")
    ctx = pkg.Context(0)
    ctx.setTableVariant(variant)
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, dtype=pkg.F64 if real == np.float64 else pkg.MIXED,
                       av_clean=av_clean)
    gs.upload(f)
    for _ in range(steps):
        gs.step()
    gs.computeForces()
    out = gs.download(names=FIELDS)
    out["minDt"] = np.array([gs.minDt])
    gs.close()
    ctx.close()
    return out


@pytest.mark.parametrize("case,side", [("sedov", 30), ("noh", 30), ("turbulence", 24), ("kh", 32)])
@pytest.mark.parametrize("real", [np.float64, np.float32])
@pytest.mark.parametrize("av_clean", [False, True])
def test_variants_return_the_same_bits(pkg, case, side, real, av_clean):
    a = run(pkg, case, side, real, 0, av_clean)
    b = run(pkg, case, side, real, 1, av_clean)
    for k in a:
        assert np.array_equal(a[k], b[k]), (case, k, float(np.abs(a[k].astype(np.float64) - b[k]).max()))


def test_timed_choice_returns_the_same_bits(pkg):
    """variant -1 times both variants on the first launches; sized above the tuner's minimum so that it really runs"""
    a = run(pkg, "turbulence", 72, np.float64, 0, False, steps=4)
    b = run(pkg, "turbulence", 72, np.float64, -1, False, steps=4)
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_host_resident_step_uses_both_variants_alike(pkg):
    """sphb200_sim_step_host: AVswitches with the Courant reduction + momentumEnergy in ranges"""
    import torch
    from sph_exa_b200 import sim as S, binding as B
    f, box, s = pkg.ic.CASES["sedov"](30)
    n = f["x"].size

    def struct(d):
        st = B.Fields()
        for k in B.FIELD_NAMES:
            setattr(st, k, d[k].data_ptr() if k in d else None)
        return st

    outs = []
    for variant in (0, 1):
        ctx = pkg.Context(0)
        ctx.setTableVariant(variant)
        gs = S.HydroVeProp(ctx, box, s, capacity=n)
        bufs = [{k: torch.empty(n, dtype=torch.float64).pin_memory() for k in B.CONSERVED} for _ in range(2)]
        for k in B.CONSERVED:
            bufs[0][k].numpy()[:] = f[k]
        for _ in range(3):
            gs.step_host(struct(bufs[0]), struct(bufs[1]), n)
            bufs.reverse()
        outs.append({k: bufs[0][k].numpy().copy() for k in B.CONSERVED})
        gs.close()
        ctx.close()
    for k in outs[0]:
        assert np.array_equal(outs[0][k], outs[1][k]), k


def test_rejects_unknown_variant(pkg):
    ctx = pkg.Context(0)
    with pytest.raises(pkg.SphError):
        ctx.setTableVariant(2)
    ctx.close()
