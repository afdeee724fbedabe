"""Host-side initial conditions and constants for the hot path (numpy; no GPU work).

Restates, formula by formula, the pieces of the reference front-end that feed the
hot path: the tabulated sinc^n kernel + normalisation (sph/include/sph/sph_kernel_tables.hpp),
the Sedov lattice (main/src/init/{grid.hpp:102-131,sedov_init.hpp:48-92,sedov_constants.hpp}),
and seeded synthetic stand-ins for the glass-based cases (Noh: noh_init.hpp:66-102,
turbulence: turbulence_init.hpp:48-97, Kelvin-Helmholtz: kelvin_helmholtz_init.hpp) whose
glass files are not available.
"""
import math
import numpy as np

TABLE_SIZE = 20000


def _wharmonic(v):
    if v == 0.0:
        return 1.0
    pv = (math.pi / 2) * v
    return math.sin(pv) / pv


def _wharmonic_derivative(v):
    if v == 0.0:
        return 0.0
    pv = (math.pi / 2) * v
    sincv = math.sin(pv) / pv
    return sincv * (math.pi / 2) * ((math.cos(pv) / math.sin(pv)) - 1.0 / pv)


def kernel_tables(dtype=np.float64, sinc_index=6.0):
    """(wh, whd, K): tabulateFunction<T,20000> of sinc^n and its derivative on [0,2] and kernel_3D_k
    (sph_kernel_tables.hpp:28-101, particles_data.hpp:364-371)"""
    n = TABLE_SIZE
    dx = (2.0 - 0.0) / (n - 1)
    wh = np.empty(n, np.float64)
    whd = np.empty(n, np.float64)
    for i in range(n):
        x = 0.0 + i * dx
        wh[i] = math.pow(_wharmonic(x), sinc_index)
        whd[i] = sinc_index * math.pow(_wharmonic(x), sinc_index - 1) * _wharmonic_derivative(x)
    # mixed mode: the reference tabulates in float arithmetic; here the double table is rounded to float
    # (differs by <= 1 float ulp, inside the mixed-mode tolerance)
    wh = wh.astype(dtype)
    whd = whd.astype(dtype)

    # kernel_3D_k: 1 / simpson(0, 2, 2000, 4 pi x^2 W(x)) with sorted partial sums
    def f(x):
        return 4.0 * math.pi * x * x * math.pow(_wharmonic(x), sinc_index)

    a, b, m = 0.0, 2.0, 2000
    num_odd = m // 2
    num_even = num_odd - 1
    h = (b - a) / float(m)
    odd = sorted(f(a + float(2 * (i + 1) - 1) * h) for i in range(num_odd))
    even = sorted(f(a + float(2 * (i + 1)) * h) for i in range(num_even))
    so = 0.0
    for v in odd:
        so += v
    se = 0.0
    for v in even:
        se += v
    K = 1.0 / (h / 3.0 * (f(a) + f(b) + 4.0 * so + 2.0 * se))
    return wh, whd, K


def ideal_gas_cv(mui, gamma, tm=np.float64):
    """idealGasCv (sph/eos.hpp:13-19): evaluated in the type of mui (Tmass)"""
    T = np.dtype(tm).type
    R = T(8.317e7)
    return T(float(R / T(mui)) / (gamma - float(T(1))))


def regular_grid(r, side):
    """regularGrid (init/grid.hpp:102-131): x fastest, then y, then z"""
    step = (2.0 * r) / side
    r_ini = -r + 0.5 * step
    c = r_ini + np.arange(side, dtype=np.float64) * step
    z, y, x = np.meshgrid(c, c, c, indexing="ij")
    return x.ravel().copy(), y.ravel().copy(), z.ravel().copy()


SEDOV = dict(gamma=5.0 / 3.0, r1=0.5, mTotal=1.0, energyTotal=1.0, width=0.1, u0=1e-8, minDt=1e-6, minDt_m1=1e-6,
             ng0=100, ngmax=150, mui=10.0)


def _real(dtype):
    return np.float64 if dtype in (0, "f64", np.float64) else np.float32


def _sedov_fields(x, y, z, n_glob, T):
    """initSedovFields (sedov_init.hpp:48-92) for given coordinates; n_glob = numParticlesGlobal"""
    c = dict(SEDOV)
    ener0 = c["energyTotal"] / math.pow(math.pi, 1.5) / 1.0 / math.pow(c["width"], 3.0)
    r = c["r1"]
    n = x.size
    total_volume = math.pow(2 * r, 3)
    h_init = np.cbrt(3.0 / (4 * math.pi) * c["ng0"] * total_volume / n_glob) * 0.5
    m_part = c["mTotal"] / n_glob
    width2 = c["width"] * c["width"]
    cv = float(ideal_gas_cv(c["mui"], c["gamma"], T))
    r2 = x * x + y * y + z * z
    temp = (ener0 * np.exp(-(r2 / width2)) + c["u0"]) / cv
    f = dict(x=x, y=y, z=z, h=np.full(n, h_init, T), m=np.full(n, m_part, T), vx=np.zeros(n, T), vy=np.zeros(n, T),
             vz=np.zeros(n, T), x_m1=np.zeros(n, T), y_m1=np.zeros(n, T), z_m1=np.zeros(n, T), temp=temp,
             du_m1=np.zeros(n, T), alpha=np.full(n, 0.05, T))
    box = ([-r, r, -r, r, -r, r], [1, 1, 1])
    settings = dict(c, Kcour=0.2, Krho=0.06, numParticlesGlobal=n_glob)
    return f, box, settings


def sedov(side, dtype="f64"):
    """Sedov blast wave on a side^3 lattice, periodic box [-0.5,0.5]^3 (sedov_init.hpp:48-133).

    returns (fields dict of conserved arrays, box (lim, bc), settings)"""
    x, y, z = regular_grid(SEDOV["r1"], side)
    return _sedov_fields(x, y, z, side ** 3, _real(dtype))


def sedov_range(side, first, last, dtype="f64"):
    """lattice sites [first,last) of the side^3 Sedov IC (one rank's slab, sedov_init.hpp:106-133 + partitionRange)"""
    T = _real(dtype)
    c = dict(SEDOV)
    ener0 = c["energyTotal"] / math.pow(math.pi, 1.5) / 1.0 / math.pow(c["width"], 3.0)
    n_glob = side ** 3
    r = c["r1"]
    step = (2.0 * r) / side
    r_ini = -r + 0.5 * step
    idx = np.arange(first, last, dtype=np.int64)
    k = idx % side
    j = (idx // side) % side
    i = idx // (side * side)
    x = r_ini + k.astype(np.float64) * step
    y = r_ini + j.astype(np.float64) * step
    z = r_ini + i.astype(np.float64) * step
    n = idx.size
    h_init = np.cbrt(3.0 / (4 * math.pi) * c["ng0"] * math.pow(2 * r, 3) / n_glob) * 0.5
    cv = float(ideal_gas_cv(c["mui"], c["gamma"], T))
    r2 = x * x + y * y + z * z
    temp = (ener0 * np.exp(-(r2 / (c["width"] * c["width"]))) + c["u0"]) / cv
    f = dict(x=x, y=y, z=z, h=np.full(n, h_init, T), m=np.full(n, c["mTotal"] / n_glob, T), vx=np.zeros(n, T),
             vy=np.zeros(n, T), vz=np.zeros(n, T), x_m1=np.zeros(n, T), y_m1=np.zeros(n, T), z_m1=np.zeros(n, T),
             temp=temp, du_m1=np.zeros(n, T), alpha=np.full(n, 0.05, T))
    box = ([-r, r, -r, r, -r, r], [1, 1, 1])
    settings = dict(c, Kcour=0.2, Krho=0.06, numParticlesGlobal=n_glob)
    return f, box, settings


def _jittered_lattice(side, lo, hi, jitter, seed):
    """seeded jittered lattice replacing the unavailable glass blocks"""
    rng = np.random.Generator(np.random.MT19937(seed))
    step = (hi - lo) / side
    c = lo + (np.arange(side, dtype=np.float64) + 0.5) * step
    z, y, x = np.meshgrid(c, c, c, indexing="ij")
    out = []
    for a in (x, y, z):
        a = a.ravel() + rng.uniform(-jitter * step, jitter * step, side ** 3)
        out.append(np.clip(a, np.nextafter(lo, hi), np.nextafter(hi, lo)))
    return out


def _noh_fields(x, y, z, n_glob, T):
    """initNohFields (noh_init.hpp:66-102) for given coordinates inside the sphere; n_glob = numParticlesGlobal"""
    gamma, mui, r1 = 5.0 / 3.0, 10.0, 0.5
    rad = np.sqrt(x * x + y * y + z * z)
    n = x.size
    total_volume = 4.0 * math.pi / 3.0 * r1 ** 3
    h_init = np.cbrt(3.0 / (4 * math.pi) * 100 * total_volume / n_glob) * 0.5
    cv = float(ideal_gas_cv(mui, gamma, T))
    rs = np.maximum(rad, 1e-10)
    vy = (-(y / rs)).astype(T)
    vz = (-(z / rs)).astype(T)
    dt0 = 1e-4
    f = dict(x=x, y=y, z=z, h=np.full(n, h_init, T), m=np.full(n, 1.0 / n_glob, T), vx=(-(x / rs)).astype(T), vy=vy,
             vz=vz, x_m1=(-(x / rs) * dt0).astype(T), y_m1=(-(y / rs) * dt0).astype(T),
             z_m1=(-(z / rs) * dt0).astype(T), temp=np.full(n, 1e-20 / cv), du_m1=np.zeros(n, T),
             alpha=np.full(n, 0.05, T))
    box = ([-r1, r1, -r1, r1, -r1, r1], [0, 0, 0])
    settings = dict(gamma=gamma, mui=mui, minDt=dt0, minDt_m1=dt0, ng0=100, ngmax=150, Kcour=0.2, Krho=0.06,
                    numParticlesGlobal=n_glob)
    return f, box, settings


def noh(side, dtype="f64", seed=42):
    """Noh implosion (noh_init.hpp:66-102) on a jittered lattice cut to the sphere r <= 0.5; open box"""
    r1 = 0.5
    x, y, z = _jittered_lattice(side, -r1, r1, 0.2, seed)
    keep = np.sqrt(x * x + y * y + z * z) <= r1
    x, y, z = x[keep], y[keep], z[keep]
    return _noh_fields(x, y, z, x.size, _real(dtype))


# ---------------------------------------------------------------- glass-block initial conditions (init/grid.hpp)
def equi_distant_sfc_segments(rank, num_ranks, leaves_per_rank=100):
    """equiDistantSfcSegments (main/src/init/grid.hpp:80-88): this rank's slice of [0, 2^63), boundaries rounded down
    to octree nodes of level log8ceil(leaves_per_rank * num_ranks) (initialDomainSplits, domaindecomp.hpp:243-256)"""
    n = leaves_per_rank * num_ranks
    level = 0
    while 8 ** level < n:
        level += 1
    full = 1 << 63
    delta = full // num_ranks
    mask = ~((1 << (3 * (21 - level))) - 1)
    bounds = [0] + [(i * delta) & mask for i in range(1, num_ranks)] + [full]
    return bounds[rank], bounds[rank + 1]


def write_template_block(path, x, y, z):
    """a glass block (coordinates in [0,1)^3) in the container of csrc/checkpoint.cu; the reference keeps its blocks
    in H5Part files with the datasets x, y, z (readTemplateBlock, init/utils.hpp:71-85)"""
    from . import binding as B
    import ctypes as C
    n = len(x)
    arrays = {k: np.zeros(n, np.float64) for k in B.CONSERVED}
    arrays["x"], arrays["y"], arrays["z"] = (np.ascontiguousarray(a, np.float64) for a in (x, y, z))
    info = B.RestartInfo()
    info.numParticlesGlobal = info.numParticles = n
    info.dtype = B.F64
    info.box = B.make_box([0, 1, 0, 1, 0, 1], [0, 0, 0])
    hc = B.HydroConst()
    f = B.Fields()
    for name in B.FIELD_NAMES:
        setattr(f, name, None)
    for name, a in arrays.items():
        setattr(f, name, a.ctypes.data)
    B.check(B.lib.sphb200_restart_write(str(path).encode(), C.byref(info), C.byref(hc), None, 0, 1, C.byref(f)))


def read_template_block(path):
    """readTemplateBlock (init/utils.hpp:71-85): x, y, z of the block"""
    from . import restart
    with restart.RestartFile(str(path)) as rf:
        n = int(rf.info.numParticles)
        out = {k: np.empty(n, np.float64) for k in "xyz"}
        rf.read_fields(0, n, rf.info.dtype, out)
    return out["x"], out["y"], out["z"]


def _assemble(ctx, block, box, multiplicity, rank, num_ranks):
    import torch
    from . import ops, binding as B
    key_start, key_end = equi_distant_sfc_segments(rank, num_ranks)
    bx = B.make_box(*box)
    dev = [torch.from_numpy(np.ascontiguousarray(a, np.float64)).cuda() for a in block]
    return ops.assembleCuboid(ctx, key_start, key_end, bx, multiplicity, *dev)


def sedov_glass(ctx, block, cbrt_num_part, rank=0, num_ranks=1, dtype="f64"):
    """SedovGlass::init (sedov_init.hpp:155-190): the periodic box tiled with rint(cbrtNumPart / cbrt(blockSize))^3
    copies of the glass block, this rank's SFC segment of it; block = (x, y, z) in [0,1)^3"""
    multi = int(np.rint(cbrt_num_part / np.cbrt(len(block[0]))))
    r = SEDOV["r1"]
    box = ([-r, r, -r, r, -r, r], [1, 1, 1])
    x, y, z = (a.cpu().numpy() for a in _assemble(ctx, block, box, [multi] * 3, rank, num_ranks))
    return _sedov_fields(x, y, z, multi ** 3 * len(block[0]), _real(dtype))


def noh_glass(ctx, block, cbrt_num_part, rank=0, num_ranks=1, dtype="f64", n_glob=None):
    """NohGlassSphere::init (noh_init.hpp:110-150): tiled open box cut to the sphere r <= r1.  numParticlesGlobal is
    the sum of the ranks' counts (MPI_Allreduce in the reference): pass n_glob when num_ranks > 1"""
    from . import ops
    multi = int(np.rint(cbrt_num_part / np.cbrt(len(block[0]))))
    r1 = 0.5
    box = ([-r1, r1, -r1, r1, -r1, r1], [0, 0, 0])
    x, y, z = _assemble(ctx, block, box, [multi] * 3, rank, num_ranks)
    x, y, z = (a.cpu().numpy() for a in ops.cutSphere(ctx, r1, x, y, z))
    return _noh_fields(x, y, z, x.size if n_glob is None else n_glob, _real(dtype))


def noh_range_coords(side, first, last, seed=42):
    """lattice sites [first, last) of the side^3 jittered Noh lattice that lie inside the sphere r <= 0.5 (one rank's
    share; the jitter stream is seeded per slice).  The fields need the global in-sphere count: noh_range_fields"""
    r1 = 0.5
    step = 2 * r1 / side
    idx = np.arange(first, last, dtype=np.int64)
    rng = np.random.Generator(np.random.MT19937([seed, first]))
    out = []
    for c in (idx % side, (idx // side) % side, idx // (side * side)):
        a = -r1 + (c.astype(np.float64) + 0.5) * step + rng.uniform(-0.2 * step, 0.2 * step, idx.size)
        out.append(np.clip(a, np.nextafter(-r1, r1), np.nextafter(r1, -r1)))
    x, y, z = out
    keep = np.sqrt(x * x + y * y + z * z) <= r1
    return x[keep], y[keep], z[keep]


def noh_range_fields(x, y, z, n_glob, dtype="f64"):
    return _noh_fields(x, y, z, n_glob, _real(dtype))


def turbulence(side, dtype="f64", seed=7):
    """subsonic turbulence IC (turbulence_init.hpp:48-97) on a jittered lattice; periodic unit box; stirring off"""
    T = _real(dtype)
    gamma, mui = 1.001, 0.62
    x, y, z = _jittered_lattice(side, -0.5, 0.5, 0.2, seed)
    n = x.size
    h_init = np.cbrt(3.0 / (4 * math.pi) * 100 * 1.0 / n) * 0.5
    cv = float(ideal_gas_cv(mui, gamma, T))
    f = dict(x=x, y=y, z=z, h=np.full(n, h_init, T), m=np.full(n, 1.0 / n, T), vx=np.zeros(n, T), vy=np.zeros(n, T),
             vz=np.zeros(n, T), x_m1=np.zeros(n, T), y_m1=np.zeros(n, T), z_m1=np.zeros(n, T),
             temp=np.full(n, 1000.0 / cv), du_m1=np.zeros(n, T), alpha=np.full(n, 0.05, T))
    box = ([-0.5, 0.5] * 3, [1, 1, 1])
    settings = dict(gamma=gamma, mui=mui, minDt=1e-4, minDt_m1=1e-4, ng0=100, ngmax=150, Kcour=0.4, Krho=0.06,
                    numParticlesGlobal=n)
    return f, box, settings


def turbulence_range(side, first, last, dtype="f64", seed=7):
    """lattice sites [first, last) of the side^3 turbulence IC for one rank of a weak-scaling run: same lattice, h, m
    and temperature as turbulence(); the jitter stream is seeded per slice (seed, first), so the particle set depends
    on the partition — a synthetic stand-in for the glass blocks either way"""
    T = _real(dtype)
    gamma, mui = 1.001, 0.62
    n_glob = side ** 3
    step = 1.0 / side
    idx = np.arange(first, last, dtype=np.int64)
    rng = np.random.Generator(np.random.MT19937([seed, first]))
    out = []
    for c in (idx % side, (idx // side) % side, idx // (side * side)):
        a = -0.5 + (c.astype(np.float64) + 0.5) * step + rng.uniform(-0.2 * step, 0.2 * step, idx.size)
        out.append(np.clip(a, np.nextafter(-0.5, 0.5), np.nextafter(0.5, -0.5)))
    x, y, z = out
    n = idx.size
    h_init = np.cbrt(3.0 / (4 * math.pi) * 100 * 1.0 / n_glob) * 0.5
    cv = float(ideal_gas_cv(mui, gamma, T))
    f = dict(x=x, y=y, z=z, h=np.full(n, h_init, T), m=np.full(n, 1.0 / n_glob, T), vx=np.zeros(n, T),
             vy=np.zeros(n, T), vz=np.zeros(n, T), x_m1=np.zeros(n, T), y_m1=np.zeros(n, T), z_m1=np.zeros(n, T),
             temp=np.full(n, 1000.0 / cv), du_m1=np.zeros(n, T), alpha=np.full(n, 0.05, T))
    box = ([-0.5, 0.5] * 3, [1, 1, 1])
    settings = dict(gamma=gamma, mui=mui, minDt=1e-4, minDt_m1=1e-4, ng0=100, ngmax=150, Kcour=0.4, Krho=0.06,
                    numParticlesGlobal=n_glob)
    return f, box, settings


def kelvin_helmholtz(side, dtype="f64", seed=11):
    """Kelvin-Helmholtz slab (kelvin_helmholtz_init.hpp:60-181), synthetic: periodic (0,1)x(0,1)x(0,1/16),
    inner band 0.25<y<0.75 at twice the particle density, shear +-0.5 with exp(-|dy|/0.025) blend"""
    T = _real(dtype)
    gamma, mui = 5.0 / 3.0, 10.0
    lz = 0.0625
    nz = max(2, int(round(side * lz)))
    rng = np.random.Generator(np.random.MT19937(seed))

    def block(nx, ny, nzz, y0, y1):
        sx, sy, sz = 1.0 / nx, (y1 - y0) / ny, lz / nzz
        gx = (np.arange(nx) + 0.5) * sx
        gy = y0 + (np.arange(ny) + 0.5) * sy
        gz = (np.arange(nzz) + 0.5) * sz
        zz, yy, xx = np.meshgrid(gz, gy, gx, indexing="ij")
        pts = [a.ravel() + rng.uniform(-0.2, 0.2, a.size) * s for a, s in ((xx, sx), (yy, sy), (zz, sz))]
        return pts

    s2 = 2.0 ** (1.0 / 3.0)
    ni = int(round(side * s2))
    nzi = max(2, int(round(nz * s2)))
    xo1, yo1, zo1 = block(side, side // 4, nz, 0.0, 0.25)
    xo2, yo2, zo2 = block(side, side // 4, nz, 0.75, 1.0)
    xi, yi, zi = block(ni, ni // 2, nzi, 0.25, 0.75)
    x = np.concatenate([xo1, xi, xo2])
    y = np.concatenate([yo1, yi, yo2])
    z = np.concatenate([zo1, zi, zo2])
    x = np.clip(x, 1e-12, 1 - 1e-12)
    y = np.clip(y, 1e-12, 1 - 1e-12)
    z = np.clip(z, 1e-12, lz - 1e-12)
    n = x.size
    n_inner = xi.size
    rho_int, rho_ext = 2.0, 1.0
    m_part = 0.5 * lz * rho_int / n_inner
    h_int = 0.5 * np.cbrt(3.0 * 100 * m_part / (4 * math.pi * rho_int))
    h_ext = 0.5 * np.cbrt(3.0 * 100 * m_part / (4 * math.pi * rho_ext))
    inner = (y > 0.25) & (y < 0.75)
    dist = np.minimum(np.abs(y - 0.25), np.abs(y - 0.75))
    w = np.clip(dist / (2 * h_ext), 0.0, 1.0)
    h_mid = 0.5 * (h_int + h_ext)
    h = np.where(inner, h_mid + (h_int - h_mid) * w, h_mid + (h_ext - h_mid) * w)
    ls = 0.025
    vx = np.where(y < 0.5, -0.5 + 1.0 / (1.0 + np.exp(-(y - 0.25) / ls)) * 1.0,
                  0.5 - 1.0 / (1.0 + np.exp(-(y - 0.75) / ls)) * 1.0)
    vy = 0.01 * np.sin(4 * math.pi * x)
    p = 2.5
    rho = np.where(inner, rho_int, rho_ext)
    cv = float(ideal_gas_cv(mui, gamma, T))
    u = p / ((gamma - 1.0) * rho)
    dt0 = 1e-7
    f = dict(x=x, y=y, z=z, h=h.astype(T), m=np.full(n, m_part, T), vx=vx.astype(T), vy=vy.astype(T),
             vz=np.zeros(n, T), x_m1=(vx * dt0).astype(T), y_m1=(vy * dt0).astype(T), z_m1=np.zeros(n, T),
             temp=u / cv, du_m1=np.zeros(n, T), alpha=np.full(n, 1.0, T))
    box = ([0.0, 1.0, 0.0, 1.0, 0.0, lz], [1, 1, 1])
    settings = dict(gamma=gamma, mui=mui, minDt=dt0, minDt_m1=dt0, ng0=100, ngmax=150, Kcour=0.4, Krho=0.06,
                    numParticlesGlobal=n)
    return f, box, settings


CASES = dict(sedov=sedov, noh=noh, turbulence=turbulence, kh=kelvin_helmholtz)


def conserved_totals(f, settings):
    """conserved quantities (main/src/observables/conserved_quantities.hpp:50-110): mass, momentum, E_kin, E_int"""
    m = f["m"].astype(np.float64)
    vx, vy, vz = (f[k].astype(np.float64) for k in ("vx", "vy", "vz"))
    cv = float(ideal_gas_cv(settings["mui"], settings["gamma"], f["m"].dtype))
    ekin = 0.5 * np.sum(m * (vx * vx + vy * vy + vz * vz))
    eint = np.sum(m * cv * f["temp"])
    return dict(mass=m.sum(), px=np.sum(m * vx), py=np.sum(m * vy), pz=np.sum(m * vz), ekin=ekin, eint=eint,
                etot=ekin + eint)
