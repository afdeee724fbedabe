"""Turbulence stirring driver (sph::TurbulenceData + driveTurbulence, sph/include/sph/hydro_turb/*) over the C ABI."""
import ctypes as C

import numpy as np

from . import binding as B
from .binding import lib, check

#: TurbulenceConstants() of main/src/init/turbulence_init.hpp:48-71 (the stirring-related ones)
DEFAULTS = dict(solWeight=0.5, stMaxModes=100000, Lbox=1.0, stEnergyPrefac=5.0e-3, stMachVelocity=0.3, epsilon=1e-15,
                rngSeed=251299, stSpectForm=1, powerLawExp=5.0 / 3.0, anglesExp=2.0)


class Turbulence:
    """host-side mode table + Ornstein-Uhlenbeck phases; `drive` adds the stirring accelerations on the device"""

    def __init__(self, **constants):
        c = dict(DEFAULTS)
        c.update({k: v for k, v in constants.items() if k in DEFAULTS})
        cfg = B.TurbConfig()
        for k in ("solWeight", "Lbox", "stEnergyPrefac", "stMachVelocity", "epsilon", "powerLawExp", "anglesExp"):
            setattr(cfg, k, float(c[k]))
        cfg.stMaxModes, cfg.rngSeed, cfg.stSpectForm = int(c["stMaxModes"]), int(c["rngSeed"]), int(c["stSpectForm"])
        self.h = C.c_void_p()
        check(lib.sphb200_turb_create(C.byref(cfg), C.byref(self.h)))
        self.numModes = lib.sphb200_turb_num_modes(self.h)

    def state(self):
        n = self.numModes
        sc = np.zeros(4)
        out = dict(modes=np.zeros(3 * n), amplitudes=np.zeros(n), phases=np.zeros(6 * n), phasesReal=np.zeros(3 * n),
                   phasesImag=np.zeros(3 * n))
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        check(lib.sphb200_turb_get(self.h, p(sc), p(out["modes"]), p(out["amplitudes"]), p(out["phases"]),
                                   p(out["phasesReal"]), p(out["phasesImag"])))
        out.update(variance=sc[0], decayTime=sc[1], solWeight=sc[2], solWeightNorm=sc[3], numModes=n)
        return out

    def rngState(self):
        n = lib.sphb200_turb_rng_state(self.h, None, C.c_size_t(0))
        buf = C.create_string_buffer(n + 1)
        lib.sphb200_turb_rng_state(self.h, buf, C.c_size_t(n + 1))
        return buf.value.decode()

    def setState(self, phases=None, rng_state=None):
        ph = np.ascontiguousarray(phases, np.float64) if phases is not None else None
        check(lib.sphb200_turb_set_state(self.h, ph.ctypes.data_as(C.c_void_p) if ph is not None else None,
                                         rng_state.encode() if rng_state is not None else None))

    def drive(self, ctx, first, last, d, dt):
        """driveTurbulence (driver.hpp:101-128) on torch tensors d[x,y,z,ax,ay,az]"""
        import torch
        dtype = B.F64 if d["ax"].dtype == torch.float64 else B.MIXED
        ptr = lambda t: C.c_void_p(t.data_ptr())
        check(lib.sphb200_turb_drive(self.h, ctx.h, dtype, C.c_uint32(first), C.c_uint32(last), ptr(d["x"]), ptr(d["y"]),
                                     ptr(d["z"]), ptr(d["ax"]), ptr(d["ay"]), ptr(d["az"]), C.c_double(dt)))

    def advance(self, ctx, dt):
        """the host half of driveTurbulence: one Ornstein-Uhlenbeck step by dt, projection, upload of the phases"""
        check(lib.sphb200_turb_advance(self.h, ctx.h if ctx is not None else None, C.c_double(dt)))

    def stir(self, ctx, first, last, d):
        """the device half: stirring accelerations of the current phases added on [first, last)"""
        import torch
        dtype = B.F64 if d["ax"].dtype == torch.float64 else B.MIXED
        ptr = lambda t: C.c_void_p(t.data_ptr())
        check(lib.sphb200_turb_stir(self.h, ctx.h, dtype, C.c_uint32(first), C.c_uint32(last), ptr(d["x"]), ptr(d["y"]),
                                    ptr(d["z"]), ptr(d["ax"]), ptr(d["ay"]), ptr(d["az"])))

    def close(self):
        if self.h:
            lib.sphb200_turb_destroy(self.h)
            self.h = None


def computeStirring(ctx, first, last, d, modes, phaseReal, phaseImag, amplitudes, solWeightNorm):
    """computeStirringGpu (stirring_gpu.cu:43-93); mode arrays are float64 device tensors"""
    import torch
    dtype = B.F64 if d["ax"].dtype == torch.float64 else B.MIXED
    ptr = lambda t: C.c_void_p(t.data_ptr())
    check(lib.sphb200_stirring(ctx.h, dtype, C.c_uint32(first), C.c_uint32(last), ptr(d["x"]), ptr(d["y"]), ptr(d["z"]),
                               ptr(d["ax"]), ptr(d["ay"]), ptr(d["az"]), int(amplitudes.numel()), ptr(modes),
                               ptr(phaseReal), ptr(phaseImag), ptr(amplitudes), C.c_double(solWeightNorm)))
