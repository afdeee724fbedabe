"""CPU test of the host half of the turbulence driver (csrc/turbulence.cu) against sph::TurbulenceData compiled from
the reference headers (oracle/_ref): the mode table, the initial Ornstein-Uhlenbeck state and the OU / projected phase
streams must be bit-identical (same arithmetic, same std::mt19937 / std::normal_distribution sequence).  Creating the
driver and advancing it need no device; the stirring kernel is tested in tests/test_gpu_turbulence.py."""
import numpy as np
import pytest


@pytest.mark.parametrize("spect", [1, 0, 2])
def test_mode_table_and_ou_stream(pkg, ref, spect):
    from sph_exa_b200.turbulence import Turbulence
    kw = dict(stSpectForm=spect, rngSeed=4711)
    rt, gt = ref.RefTurbulence(**kw), Turbulence(**kw)
    assert gt.numModes == rt.numModes and gt.numModes > 0
    a, b = gt.state(), rt.state()
    for k in ("variance", "decayTime", "solWeight", "solWeightNorm"):
        assert a[k] == b[k], k
    for k in ("modes", "amplitudes", "phases"):
        assert np.array_equal(a[k], b[k]), k
    # three OU steps on the host alone; the reference advances inside driveTurbulence (one particle is enough)
    x = np.zeros(1)
    acc = [np.zeros(1) for _ in range(3)]
    for step, dt in enumerate((1e-4, 3e-4, 2e-4)):
        rt.drive(0, 1, x, x, x, *acc, dt)
        gt.advance(None, dt)
        a, b = gt.state(), rt.state()
        for k in ("phases", "phasesReal", "phasesImag"):
            assert np.array_equal(a[k], b[k]), (step, k)
    rt.close()
    gt.close()
