"""Multi-rank Domain + HydroVeProp of the C ABI (sphb200_domain_*, csrc/domain.cu) on ONE GPU: P virtual ranks through
the loopback transport, compared with the reference's CPU pipeline (oracle/_ref, RefSim) on the same initial conditions.

Parity: the owned particles of all ranks, concatenated in rank order, ARE the single-rank SFC-sorted state -- keys and
neighbor counts bit-exact, every field <= 1e-10 (fp64).  Halo discovery is pinned against cstone::findHalos.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import to_dev, to_host, rel_err

pytestmark = pytest.mark.gpu

FIELDS = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha", "xm", "kx",
          "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax", "ay", "az", "du"]
# lattice accelerations and off-diagonal c_ij are cancellation residuals: scaled by the vector / tensor magnitude
GROUP_SCALE = {"ax": ("ax", "ay", "az"), "ay": ("ax", "ay", "az"), "az": ("ax", "ay", "az"),
               "c12": ("c11", "c22", "c33"), "c13": ("c11", "c22", "c33"), "c23": ("c11", "c22", "c33"),
               "x_m1": ("x_m1", "y_m1", "z_m1"), "y_m1": ("x_m1", "y_m1", "z_m1"), "z_m1": ("x_m1", "y_m1", "z_m1"),
               "vx": ("vx", "vy", "vz"), "vy": ("vx", "vy", "vz"), "vz": ("vx", "vy", "vz")}


def make_group(pkg, ctx, case, side, P, capacity_factor=3.0, shuffle_seed=None, shares=None):
    from sph_exa_b200 import domain as D
    f, box, s = pkg.ic.CASES[case](side)
    n = f["x"].size
    s = dict(s)
    s["numParticlesGlobal"] = n
    comm = D.Comm.loopback(P)
    ranks = []
    order = np.arange(n)
    if shuffle_seed is not None:
        order = np.random.default_rng(shuffle_seed).permutation(n)
    bucket = max(64, n // (100 * P))
    cuts = [(n * r) // P for r in range(P + 1)] if shares is None else [int(round(n * c)) for c in np.cumsum([0] + shares)]
    for r in range(P):
        first, last = cuts[r], cuts[r + 1]
        mine = {k: v[order[first:last]] for k, v in f.items()}
        # every virtual rank has its own context (scratch, gather records, time-step scalars) on the shared stream
        rctx = pkg.Context(0)
        sim = D.DomainRank(rctx, box, s, int(capacity_factor * n / P) + 4096, comm, r, bucket_size=bucket)
        sim.own_ctx = True
        sim.upload(mine)
        ranks.append(sim)
    return D.DistHydroVe(ranks), comm, f, box, s


def compare(group, ref_sim, names, tol=1e-10):
    out = group.download(names=[k for k in names if k != "keys"])
    assert np.array_equal(out["keys"], ref_sim.keys), "SFC keys of the owned particles differ from the reference"
    assert np.all(np.diff(out["keys"].astype(np.int64)) >= 0)
    worst = {}
    for k in names:
        if k in ("keys",):
            continue
        want = ref_sim.nc if k == "nc" else ref_sim.f[k]
        if k == "nc":
            assert np.array_equal(out[k], want), "neighbor counts differ from the reference"
            continue
        scale = max(max(np.abs(ref_sim.f[g]).max() for g in GROUP_SCALE.get(k, (k,))), 1e-300)
        err = np.abs(out[k] - want).max() / scale
        worst[k] = err
        assert err < tol, f"{k}: {err:.3e}"
    return worst


@pytest.mark.parametrize("case,side,P,steps", [("sedov", 20, 2, 3), ("sedov", 24, 3, 2), ("sedov", 32, 8, 2),
                                               ("noh", 30, 2, 3), ("noh", 30, 3, 2), ("kh", 40, 2, 2),
                                               ("turbulence", 24, 2, 3), ("turbulence", 30, 8, 2)])
def test_virtual_ranks_equal_reference(pkg, ctx, ref, case, side, P, steps):
    """P virtual ranks: every owned particle's key / nc / fields equal the reference's single-rank CPU run"""
    from oracle.ref_pipeline import RefSim
    group, comm, f, box, s = make_group(pkg, ctx, case, side, P)
    try:
        rs = RefSim(f, box, s)
        for step in range(steps):
            group.computeForces()
            rs.computeForces()
            compare(group, rs, FIELDS + ["nc"])
            infos = [r.info() for r in group.ranks]
            assert sum(i.endIndex - i.startIndex for i in infos) == f["x"].size
            # halo traffic is symmetric: what q receives from r is what r sends to q
            for a in infos:
                for b in infos:
                    assert a.haloSend[b.rank] == b.haloRecv[a.rank]
            # identical replicated assignment, particles inside their key range
            for i in infos:
                assert list(i.assignment[:P + 1]) == list(infos[0].assignment[:P + 1])
            group.integrate()
            rs.integrate()
            assert abs(group.ranks[0].minDt - rs.minDt) <= 1e-12 * rs.minDt
            for r in group.ranks:
                assert r.minDt == group.ranks[0].minDt
        group.sync()
        rs.sync()
        compare(group, rs, ["x", "y", "z", "h", "vx", "vy", "vz", "temp", "alpha", "du_m1"])
    finally:
        group.close()
        comm.close()


def test_migration_from_a_shuffled_start(pkg, ctx, ref):
    """ranks start with random particles: the first sync moves ~ (P-1)/P of them (exchangeParticles), later syncs move
    what crosses the boundaries"""
    from oracle.ref_pipeline import RefSim
    group, comm, f, box, s = make_group(pkg, ctx, "turbulence", 24, 3, shuffle_seed=5)
    try:
        rs = RefSim(f, box, s)
        group.computeForces()
        rs.computeForces()
        compare(group, rs, FIELDS + ["nc"])
        infos = [r.info() for r in group.ranks]
        assert sum(i.migratedOut for i in infos) == sum(i.migratedIn for i in infos) > f["x"].size // 2
        group.integrate()
        rs.integrate()
        group.computeForces()
        rs.computeForces()
        compare(group, rs, FIELDS + ["nc"])
    finally:
        group.close()
        comm.close()


def test_ranks_that_start_empty_or_overfull(pkg, ctx, ref):
    """the file reader of one rank holds everything, another holds nothing: the first sync hands every rank its share
    (the reference's `domain.sync` accepts any initial distribution, domain.hpp:143-160)"""
    from oracle.ref_pipeline import RefSim
    group, comm, f, box, s = make_group(pkg, ctx, "sedov", 20, 3, capacity_factor=3.5, shares=[0.0, 1.0, 0.0])
    try:
        rs = RefSim(f, box, s)
        for _ in range(2):
            group.step()
            rs.step()
        group.computeForces()
        rs.computeForces()
        compare(group, rs, FIELDS + ["nc"])
        own = [i.endIndex - i.startIndex for i in (r.info() for r in group.ranks)]
        assert min(own) > 0.25 * f["x"].size and sum(own) == f["x"].size
    finally:
        group.close()
        comm.close()


def test_conserved_quantities_over_ranks(pkg, ctx, ref):
    group, comm, f, box, s = make_group(pkg, ctx, "sedov", 20, 3)
    try:
        group.step()
        group.computeForces()
        c = group.conserved()
        own = group.download(names=["x", "y", "z", "vx", "vy", "vz", "m", "temp"])
        want = ref.conserved_quantities(own["x"], own["y"], own["z"], own["vx"], own["vy"], own["vz"], own["m"],
                                        own["temp"], s["mui"], s["gamma"])
        got = [c["ecin"], c["eint"], c["egrav"], c["linmom_x"], c["linmom_y"], c["linmom_z"], c["angmom_x"],
               c["angmom_y"], c["angmom_z"]]
        pscale = float(np.abs(own["m"] * np.abs(own["vx"])).sum())
        scale = [abs(want[0]), abs(want[1]), 1.0] + [pscale] * 6
        for k in range(9):
            assert abs(got[k] - want[k]) <= 1e-10 * max(scale[k], 1e-300), k
    finally:
        group.close()
        comm.close()


@pytest.mark.parametrize("case,side", [("sedov", 24), ("noh", 30), ("kh", 40)])
def test_find_halos_equals_reference(pkg, ctx, ref, case, side):
    """sphb200_halo_radii + sphb200_find_halos against Halos::discover's radii and cstone::findHalos on the same linked
    octree, for several assigned leaf ranges (periodic, open and anisotropic boxes)"""
    import torch
    from sph_exa_b200 import ops
    from sph_exa_b200.binding import lib, ptr, check
    f, box, s = pkg.ic.CASES[case](side)
    lim, bc = list(box[0]), list(box[1])
    for d, k in enumerate("xyz"):
        if bc[d] != 1:
            lim[2 * d], lim[2 * d + 1] = float(f[k].min()), float(f[k].max())
    box = (lim, bc)
    keys = ref.compute_sfc_keys(f["x"], f["y"], f["z"], box)
    keys, perm = ref.sort_by_key(keys)
    h = np.ascontiguousarray(f["h"][perm], np.float64)
    leaves, counts, _ = ref.compute_octree(keys, 64)
    tree = ref.build_octree(leaves)
    L = counts.size
    layout = np.zeros(L + 1, np.uint32)
    np.cumsum(counts, out=layout[1:])
    b = pkg.make_box(*box)
    d_pre, d_child, d_i2l = to_dev(tree["prefixes"]), to_dev(tree["childOffsets"]), to_dev(tree["internalToLeaf"])
    d_leaves, d_layout, d_h = to_dev(leaves), to_dev(layout), to_dev(h)
    for first, last, ext in [(0, L // 2, 1.0), (L // 3, 2 * L // 3, 1.0), (L - L // 5, L, 1.0), (L // 4, L // 2, 1.5)]:
        want_r = ref.halo_radii(h, layout, first, last, ext)
        want = ref.find_halos(tree, leaves, want_r, box, first, last)
        radii = torch.zeros(L, dtype=torch.float32, device="cuda")
        check(lib.sphb200_halo_radii(ctx.h, ptr(d_h), pkg.F64, ptr(d_layout), first, last, C.c_float(ext), ptr(radii)))
        assert np.array_equal(to_host(radii), want_r)
        flags = torch.zeros(L, dtype=torch.int32, device="cuda")
        nb = C.c_int(0)
        check(lib.sphb200_find_halos(ctx.h, ptr(d_pre), ptr(d_child), ptr(d_i2l), ptr(d_leaves), ptr(radii), C.byref(b),
                                     first, last, ptr(flags), C.byref(nb)))
        got = to_host(flags)
        assert np.array_equal(got != 0, want != 0), f"{case} [{first},{last}) ext {ext}"
        assert want[first:last].sum() == 0 and want.sum() > 0 and 0 < nb.value <= last - first


@pytest.mark.parametrize("case,side,P", [("sedov", 24, 3), ("noh", 30, 2), ("turbulence", 24, 4)])
def test_halo_sets_equal_find_halos(pkg, ctx, ref, case, side, P):
    """the owner-side discovery sends exactly the particles of the leaves cstone::findHalos flags for the requester on
    the global bucket-64 tree: the reference's `Halos:` count per rank"""
    group, comm, f, box, s = make_group(pkg, ctx, case, side, P)
    try:
        group.sync()
        infos = [r.info() for r in group.ranks]
        own = group.download(names=["x", "y", "z", "h"])
        keys = own["keys"]
        lim, bc = list(box[0]), list(box[1])
        lim = group.ranks[0].boxLimits()
        gbox = (lim, bc)
        assert np.array_equal(ref.compute_sfc_keys(own["x"], own["y"], own["z"], gbox), keys)
        leaves, counts, _ = ref.compute_octree(keys, 64)
        tree = ref.build_octree(leaves)
        layout = np.zeros(counts.size + 1, np.uint32)
        np.cumsum(counts, out=layout[1:])
        bounds = np.array(infos[0].assignment[:P + 1], np.uint64)
        for i in infos:
            first = int(np.searchsorted(leaves, bounds[i.rank], side="left"))
            last = int(np.searchsorted(leaves, bounds[i.rank + 1], side="left"))
            assert leaves[first] == bounds[i.rank] and leaves[last] == bounds[i.rank + 1]
            radii = ref.halo_radii(own["h"], layout, first, last, 1.0)
            flags = ref.find_halos(tree, leaves, radii, gbox, first, last)
            want_halos = int(counts[flags != 0].sum())
            got_halos = i.nWithHalos - (i.endIndex - i.startIndex)
            assert got_halos == want_halos, f"rank {i.rank}: {got_halos} halos, reference {want_halos}"
            # per peer as well
            for q in range(P):
                a = int(np.searchsorted(leaves, bounds[q], side="left"))
                b = int(np.searchsorted(leaves, bounds[q + 1], side="left"))
                assert int(counts[a:b][flags[a:b] != 0].sum()) == i.haloRecv[q]
    finally:
        group.close()
        comm.close()


@pytest.mark.parametrize("case,side,P,stir", [("sedov", 24, 2, False), ("turbulence", 24, 3, True), ("noh", 30, 2, False)])
def test_step_host_equals_upload_step_download(pkg, ctx, ref, case, side, P, stir):
    """sphb200_domain_step_host (transfers staged beside the kernels, late fields reordered right before IAD, early
    Courant step, MomentumEnergy in 16 ranges) == device-resident steps, bit for bit, over several steps"""
    import torch
    from sph_exa_b200 import binding as B
    from sph_exa_b200 import domain as D
    ga, comm_a, f, box, s = make_group(pkg, ctx, case, side, P)
    gb, comm_b, _, _, _ = make_group(pkg, ctx, case, side, P)
    turbs = []
    if stir:
        from sph_exa_b200.turbulence import Turbulence
        for grp in (ga, gb):
            for r in grp.ranks:
                t = Turbulence()
                r.attachTurbulence(t)
                turbs.append(t)
    try:
        cap = ga.ranks[0].capacity
        n = f["x"].size
        n_h = [(n * (r + 1)) // P - (n * r) // P for r in range(P)]
        hin = [{k: torch.empty(cap, dtype=torch.float64).pin_memory() for k in B.CONSERVED} for _ in range(P)]
        hout = [{k: torch.empty(cap, dtype=torch.float64).pin_memory() for k in B.CONSERVED} for _ in range(P)]
        for r in range(P):
            first = (n * r) // P
            for k in B.CONSERVED:
                hin[r][k].numpy()[:n_h[r]] = f[k][first:first + n_h[r]]

        def structs(bufs):
            out = []
            for r in range(P):
                fs = B.Fields()
                for k in B.FIELD_NAMES:
                    setattr(fs, k, bufs[r][k].data_ptr() if k in bufs[r] else None)
                out.append(fs)
            return out

        for step in range(3):
            ga.step()
            want = [r.download() for r in ga.ranks]
            n_h = D.step_host(gb, n_h, structs(hin), structs(hout), cap)
            for r in range(P):
                assert n_h[r] == want[r]["x"].size
                for k in B.CONSERVED:
                    assert np.array_equal(hout[r][k].numpy()[:n_h[r]], want[r][k]), (step, r, k)
                assert gb.ranks[r].minDt == ga.ranks[r].minDt
            hin, hout = hout, hin
    finally:
        ga.close()
        gb.close()
        comm_a.close()
        comm_b.close()
        for t in turbs:
            t.close()
