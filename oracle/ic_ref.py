"""TEST INFRASTRUCTURE ONLY: the Sedov initial condition for the CPU reference arm of bench.py, so that the arm never
imports the product package.  Formulas: main/src/init/grid.hpp:102-131 (regularGrid), sedov_init.hpp:48-92,
sedov_constants.hpp:11-21 of the reference; the product's own copy lives in sph-exa_b200/ic.py and the two are
compared in tests/test_oracle.py."""
import math

import numpy as np


def sedov(side, real=np.float64):
    r, width, ng0, gamma, mui = 0.5, 0.1, 100, 5.0 / 3.0, 10.0
    n = side ** 3
    step = (2.0 * r) / side
    c = (-r + 0.5 * step) + np.arange(side, dtype=np.float64) * step
    z, y, x = np.meshgrid(c, c, c, indexing="ij")
    x, y, z = x.ravel().copy(), y.ravel().copy(), z.ravel().copy()
    ener0 = 1.0 / math.pow(math.pi, 1.5) / 1.0 / math.pow(width, 3.0)
    h = np.cbrt(3.0 / (4 * math.pi) * ng0 * math.pow(2 * r, 3) / n) * 0.5
    T = np.dtype(real).type
    cv = float(T(float(T(8.317e7) / T(mui)) / (gamma - float(T(1)))))  # idealGasCv in Tmass (sph/eos.hpp:13-19)
    temp = (ener0 * np.exp(-((x * x + y * y + z * z) / (width * width))) + 1e-8) / cv
    zero = lambda: np.zeros(n, real)
    f = dict(x=x, y=y, z=z, h=np.full(n, h, real), m=np.full(n, 1.0 / n, real), vx=zero(), vy=zero(), vz=zero(),
             x_m1=zero(), y_m1=zero(), z_m1=zero(), temp=temp, du_m1=zero(), alpha=np.full(n, 0.05, real))
    box = ([-r, r, -r, r, -r, r], [1, 1, 1])
    settings = dict(gamma=gamma, mui=mui, minDt=1e-6, minDt_m1=1e-6, ng0=ng0, ngmax=150, Kcour=0.2, Krho=0.06,
                    numParticlesGlobal=n)
    return f, box, settings
