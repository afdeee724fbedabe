"""CUDA-event times of the five pair operators on the state of a running simulation (tier-1 entry points), and the
fields they produce (for a bitwise comparison between SPHB200_TABLE=l1 and SPHB200_TABLE=smem runs):

  SPHB200_TABLE=l1 OUT=/tmp/a.npy python profiles/microbench/pair_operator_times.py turbulence 200
"""
import ctypes as C, os, sys, numpy as np, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import sphexa_b200_loader
pkg = sphexa_b200_loader.load()
from sph_exa_b200 import sim as S, binding as B
from sph_exa_b200.binding import lib, check
case, side = sys.argv[1], int(sys.argv[2])
f, box, s = pkg.ic.CASES[case](side)
n = f["x"].size
ctx = pkg.Context(0)
gs = S.HydroVeProp(ctx, box, s, capacity=n)
gs.upload(f)
for _ in range(3): gs.step()
gs.computeForces()
cfg = B.SimConfig(); lib.sphb200_sim_get_config(gs.h, C.byref(cfg))
fd = gs.deviceFields()
bx = B.Box(); lib.sphb200_sim_get_box(gs.h, C.byref(bx))
nb = C.c_void_p(lib.sphb200_sim_neighbors(gs.h))
a, b = gs.startIndex, gs.endIndex
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ops = {"xmass": lambda: check(lib.sphb200_xmass(ctx.h, cfg.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(bx), C.byref(cfg.hc), nb, C.byref(fd))),
       "gradh": lambda: check(lib.sphb200_ve_def_gradh(ctx.h, cfg.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(bx), C.byref(cfg.hc), nb, C.byref(fd))),
       "iad": lambda: check(lib.sphb200_iad_divv_curlv(ctx.h, cfg.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(bx), C.byref(cfg.hc), nb, C.byref(fd))),
       "av": lambda: check(lib.sphb200_av_switches(ctx.h, cfg.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(bx), C.byref(cfg.hc), nb, C.byref(fd))),
       "momentum": lambda: check(lib.sphb200_momentum_energy(ctx.h, cfg.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(bx), C.byref(cfg.hc), nb, C.byref(fd), None))}
res = {}
for k, fn in ops.items():
    res[k] = round(timed(fn), 3)
out = gs.download(names=["xm", "kx", "gradh", "c11", "divv", "alpha", "ax", "du"])
np.save(os.environ.get("OUT", "/tmp/out.npy"), np.stack([out[k] for k in sorted(out)]))
print(case, side, "SPHB200_TABLE=" + os.environ.get("SPHB200_TABLE", "timed"), res, flush=True)
