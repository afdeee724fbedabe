"""BASELINE.json configs[1] at its full size: Sedov 200^3 = 8 M particles, f64, one B200.

Where the reference is affordable at this size it is the checker (all 8 M SFC keys, the sorted keys, the bucket-64
leaf array and its counts: oracle/_ref, the reference's own headers); the rest is checked through size-independent
properties: the ordering is a permutation and sorting is idempotent, the tree update has a fixed point, neighbor sets
of sampled targets equal an all-to-all search (the reference's own test oracle, domain/test/unit/neighbors/
all_to_all.hpp), neighborhoods are symmetric for equal h, mass / momentum / energy are conserved over steps.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu

SIDE = 200
N = SIDE ** 3


@pytest.fixture(scope="module")
def sedov200(pkg):
    return pkg.ic.sedov(SIDE)


def test_keys_sort_and_tree_at_full_size(pkg, ctx, ref, sedov200):
    import torch
    from sph_exa_b200 import ops
    f, box, s = sedov200
    bx = pkg.make_box(*box)
    x, y, z = (to_dev(f[k]) for k in "xyz")
    keys_d = ops.computeSfcKeysGpu(ctx, x, y, z, bx)
    want = ref.compute_sfc_keys(f["x"], f["y"], f["z"], box)
    assert np.array_equal(to_host(keys_d, np.uint64), want)          # every one of the 8 M keys

    work = keys_d.clone()
    perm = ops.sortByKeyGpu(ctx, work)
    assert bool((work[1:] >= work[:-1]).all())                       # keys < 2^63: signed compare is the key order
    want_sorted = np.sort(want, kind="stable")
    assert np.array_equal(to_host(work, np.uint64), want_sorted)
    assert bool((keys_d[perm.long()] == work).all())
    assert int(torch.bincount(perm.long(), minlength=N).max()) == 1  # a permutation of 0..N-1
    # lattice keys are unique, so the stable order is the only order: equal to the reference's ordering
    assert np.array_equal(to_host(perm, np.uint32), np.argsort(want, kind="stable").astype(np.uint32))
    again = work.clone()
    perm2 = ops.sortByKeyGpu(ctx, again)                             # idempotence
    assert bool((again == work).all())
    assert bool((perm2.long() == torch.arange(N, device="cuda")).all())

    # bucket-64 leaf array from scratch (computeOctree, csarray.hpp:454-468) against the reference on the same keys
    leaves, counts, _ = ops.computeOctreeGpu(ctx, work, 64)
    wl, wc, _ = ref.compute_octree(want_sorted, 64, capacity=400_000)
    gl, gc = to_host(leaves, np.uint64), to_host(counts, np.uint32)
    assert np.array_equal(gl, wl) and np.array_equal(gc, wc)
    assert gc.sum() == N and gc.max() <= 64 and gl[0] == 0 and gl[-1] == 1 << 63
    assert gl.size - 1 == 262144                                     # 64^3 leaves at level 6


def pbc_neighbors(x, y, z, i, radius):
    d2 = np.zeros(x.size)
    for a in (x, y, z):
        d = a - a[i]
        d -= np.rint(d)  # box length 1
        d2 += d * d
    nb = np.nonzero(d2 < radius * radius)[0]
    return nb[nb != i]


def test_step_at_full_size(pkg, ctx, sedov200):
    """three steps of the propagator: neighbor sets of sampled targets against an all-to-all search, symmetry of the
    neighborhoods, tree fixed point, conservation"""
    import torch
    from sph_exa_b200 import sim as S, binding as B
    f, box, s = sedov200
    gs = S.HydroVeProp(ctx, box, s, capacity=N)
    gs.upload(f)
    t0 = pkg.ic.conserved_totals(f, s)

    gs.computeForces()
    assert gs.numLeaves == 262144 and int(gs.leafCounts().sum()) == N
    st = gs.download(names=["x", "y", "z", "h", "nc"], with_keys=True)
    keys = st["keys"]
    assert np.all(keys[1:] > keys[:-1])
    nc = st["nc"]
    assert gs.totalNeighbors() == int(nc.sum(dtype=np.int64)) == 93 * N   # 92 lattice neighbors + self, everywhere
    ngmax = s["ngmax"]
    rm_d = torch.empty((N, ngmax), dtype=torch.int32, device="cuda")
    assert B.lib.sphb200_neighbors_to_rowmajor(ctx.h, C.c_void_p(B.lib.sphb200_sim_neighbors(gs.h)),
                                               C.c_void_p(to_dev(nc).data_ptr()), C.c_uint32(N), C.c_uint(ngmax),
                                               C.c_void_p(rm_d.data_ptr())) == 0
    rng = np.random.Generator(np.random.MT19937(200))
    sample = np.concatenate([[0, 1, N - 1], rng.integers(0, N, 60)])
    rows = rm_d[torch.from_numpy(sample).cuda()].cpu().numpy().view(np.uint32)
    for i, row in zip(sample, rows):
        want = pbc_neighbors(st["x"], st["y"], st["z"], i, 2 * st["h"][i])
        got = row[:nc[i] - 1]
        assert np.all(got[1:] > got[:-1])                           # ascending j: the canonical summation order
        assert np.array_equal(got, want), i
    # equal h everywhere: j in N(i) <=> i in N(j), so the sum of (j - i) over all pairs vanishes
    cnt = torch.from_numpy((nc.astype(np.int64) - 1)).cuda()
    valid = torch.arange(ngmax, device="cuda").unsqueeze(0) < cnt.unsqueeze(1)
    tot = 0
    for a in range(0, N, 1_000_000):                                # chunks: keep the temporaries small
        b = min(N, a + 1_000_000)
        j = rm_d[a:b].long()
        i = torch.arange(a, b, device="cuda").unsqueeze(1)
        tot += int(torch.where(valid[a:b], j - i, torch.zeros_like(j)).sum())
    assert tot == 0
    del rm_d, valid, cnt

    gs.integrate()
    leaves1 = gs.leaves()
    for _ in range(2):
        gs.step()
    # the blast moves few particles across leaf boundaries in three steps: the tree update is at its fixed point
    assert np.array_equal(gs.leaves(), leaves1)
    out = gs.download()
    t1 = pkg.ic.conserved_totals(out, s)
    assert t1["mass"] == t0["mass"]
    pscale = float(np.sum(np.abs(out["m"] * out["vx"]))) + 1e-300
    for k in ("px", "py", "pz"):
        assert abs(t1[k]) / pscale < 1e-10, k
    assert abs(t1["etot"] - t0["etot"]) <= 1e-6 * abs(t0["etot"])
    for k in "xyz":
        assert out[k].min() >= -0.5 and out[k].max() < 0.5
    assert out["h"].min() > 0 and gs.minDt > 0
    gs.close()


# ---------------------------------------------------------------------------------------------------------------
# fields against the reference at the BASELINE sizes (field parity used to stop at 48^3)
FIELDS_FULL = ["x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "xm", "kx", "gradh", "prho", "c", "c11", "c12",
               "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay", "az", "du"]
SCALE_GROUP = {"ax": ("ax", "ay", "az"), "ay": ("ax", "ay", "az"), "az": ("ax", "ay", "az"),
               "c12": ("c11", "c22", "c33"), "c13": ("c11", "c22", "c33"), "c23": ("c11", "c22", "c33"),
               "vx": ("vx", "vy", "vz"), "vy": ("vx", "vy", "vz"), "vz": ("vx", "vy", "vz")}


def _one_reference_step(pkg, ctx, ref, f, box, s, tol=1e-10, turb=False):
    """one computeForces + integrate of the reference's CPU operators (all host threads) against the GPU propagator:
    keys and neighbor counts exact, every field <= tol (scaled by the magnitude of the field / its vector / tensor)"""
    import os
    from oracle.ref_pipeline import RefSim
    from sph_exa_b200 import sim as S
    ref.set_num_threads(len(os.sched_getaffinity(0)))
    n = f["x"].size
    gs = S.HydroVeProp(ctx, box, s, capacity=n)
    gs.upload(f)
    rs = RefSim(f, box, s)
    gs.computeForces()
    rs.computeForces()
    out = gs.download(names=FIELDS_FULL + ["nc"], with_keys=True)
    assert np.array_equal(out["keys"], rs.keys)
    assert np.array_equal(out["nc"], rs.nc)
    worst = {}
    for k in FIELDS_FULL:
        scale = max(max(np.abs(rs.f[g]).max() for g in SCALE_GROUP.get(k, (k,))), 1e-300)
        worst[k] = float(np.abs(out[k] - rs.f[k]).max() / scale)
    bad = {k: v for k, v in worst.items() if not v < tol}
    assert not bad, f"fields beyond {tol}: {bad}"
    gs.integrate()
    rs.integrate()
    assert abs(gs.minDt - rs.minDt) <= 1e-12 * rs.minDt
    out = gs.download(names=["x", "y", "z", "h", "vx", "vy", "vz", "temp", "x_m1", "du_m1"])
    for k, v in out.items():
        scale = max(np.abs(rs.f[k]).max(), 1e-300)
        assert np.abs(v - rs.f[k]).max() / scale < tol, k
    gs.close()
    return worst


def test_fields_equal_reference_sedov_200(pkg, ctx, ref, sedov200):
    """BASELINE.json configs[1], all 8 M particles: every field of one full step against the reference"""
    f, box, s = sedov200
    _one_reference_step(pkg, ctx, ref, f, box, s)


def test_fields_equal_reference_noh_share(pkg, ctx, ref):
    """BASELINE.json configs[2]: the per-GPU share of Noh 300^3 over 8 GPUs (150^3 lattice cut to the sphere, 1.77 M
    particles, open box)"""
    f, box, s = pkg.ic.noh(150)
    _one_reference_step(pkg, ctx, ref, f, box, s)


def test_fields_equal_reference_turbulence_250(pkg, ctx, ref):
    """BASELINE.json configs[3] per GPU: 250^3 = 15.6 M jittered particles (the h iteration and the Atwood ramp are
    active here, unlike on the Sedov lattice)"""
    f, box, s = pkg.ic.turbulence(250)
    _one_reference_step(pkg, ctx, ref, f, box, s)
