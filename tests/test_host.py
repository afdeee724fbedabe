"""CPU tests of the host-side logic: IC formulas, SFC assignment, halo layout, and a world_size-2 gloo run of the
multi-rank bookkeeping (counts all-reduce -> identical assignment on every rank -> migration counts)."""
import math
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sedov_ic_formulas(pkg):
    """sedov_init.hpp:48-92, sedov_constants.hpp:11-21, grid.hpp:102-131"""
    f, box, s = pkg.ic.sedov(50)
    n = 125000
    assert f["x"].size == n
    assert abs(f["h"][0] - 0.5 * np.cbrt(3 / (4 * math.pi) * 100 / n)) < 1e-16
    assert abs(f["h"][0] - 0.02879) < 1e-5
    assert f["m"][0] == 1.0 / n
    assert f["x"][0] == -0.5 + 0.5 / 50 and f["x"][1] == -0.5 + 0.5 / 50 + 1.0 / 50
    assert f["y"][49] == f["y"][0] and f["y"][50] != f["y"][0]      # x fastest
    cv = 8.317e7 / 10.0 / (5.0 / 3.0 - 1.0)
    ener0 = 1.0 / math.pi ** 1.5 / 0.1 ** 3
    r2 = f["x"] ** 2 + f["y"] ** 2 + f["z"] ** 2
    assert np.allclose(f["temp"], (ener0 * np.exp(-r2 / 0.01) + 1e-8) / cv, rtol=1e-14)
    assert box == ([-0.5, 0.5] * 3, [1, 1, 1]) and s["minDt"] == 1e-6
    # a rank's slab equals the corresponding slice of the global IC
    g, _, _ = pkg.ic.sedov_range(50, 1000, 5000)
    for k in f:
        assert np.array_equal(g[k], f[k][1000:5000]), k


def test_equi_distant_sfc_segments(pkg, ref):
    """the SFC slice a rank assembles its glass-block IC in (main/src/init/grid.hpp:80-88), against the reference"""
    for num_ranks in (1, 2, 3, 7, 8, 64, 100, 1000):
        for rank in sorted({0, 1 % num_ranks, num_ranks // 2, num_ranks - 1}):
            assert pkg.ic.equi_distant_sfc_segments(rank, num_ranks) == ref.equi_distant_sfc_segment(rank, num_ranks)
    assert pkg.ic.equi_distant_sfc_segments(0, 1) == (0, 1 << 63)
    assert pkg.ic.equi_distant_sfc_segments(2, 5, 10) == ref.equi_distant_sfc_segment(2, 5, 10)


def test_per_rank_slices_of_the_weak_scaling_ics(pkg):
    """turbulence_range / noh_range_coords: a rank's lattice slab for the multi-GPU runs of configs[2] and [3]"""
    side, P = 20, 3
    n = side ** 3
    from sph_exa_b200.dist import partition_range
    tot, xs = 0, []
    for r in range(P):
        a, b = partition_range(n, r, P)
        f, box, s = pkg.ic.turbulence_range(side, a, b)
        g, _, _ = pkg.ic.turbulence_range(side, a, b)
        assert all(np.array_equal(f[k], g[k]) for k in f)            # seeded
        assert f["x"].size == b - a and s["numParticlesGlobal"] == n
        for k in "xyz":
            assert f[k].min() > -0.5 and f[k].max() < 0.5
        assert f["m"][0] == 1.0 / n
        tot += f["x"].size
        xs.append(f["x"])
    assert tot == n
    whole, _, sw = pkg.ic.turbulence(side)
    assert whole["h"][0] == f["h"][0] and sw["Kcour"] == s["Kcour"]      # same h, m, constants as the global IC
    cnt = 0
    for r in range(P):
        a, b = partition_range(n, r, P)
        x, y, z = pkg.ic.noh_range_coords(side, a, b)
        assert np.all(np.sqrt(x * x + y * y + z * z) <= 0.5)
        cnt += x.size
    assert 0.45 * n < cnt < 0.6 * n                                    # pi/6 of the cube
    f, box, s = pkg.ic.noh_range_fields(x, y, z, cnt)
    assert f["m"][0] == 1.0 / cnt and box[1] == [0, 0, 0] and s["numParticlesGlobal"] == cnt
    r = np.sqrt(x * x + y * y + z * z)
    assert np.allclose(f["vx"], -x / r) and np.allclose(f["x_m1"], f["vx"] * 1e-4)


def test_hilbert_transition_table(pkg, ref):
    """the 24-state, three-levels-per-lookup table of the GPU key kernel (csrc/sfc.cu), walked on the host: seven
    lookups reproduce iHilbert (domain/include/cstone/sfc/hilbert.hpp:58-110) for random and extreme coordinates"""
    import ctypes as C
    B = pkg.binding
    tab = np.zeros(24 * 512, np.uint16)
    assert B.lib.sphb200_hilbert_table(tab.ctypes.data_as(C.c_void_p), C.c_size_t(tab.size)) == 24
    assert B.lib.sphb200_hilbert_table(tab.ctypes.data_as(C.c_void_p), C.c_size_t(100)) == -1
    rng = np.random.Generator(np.random.MT19937(21))
    n = 200_000
    ix, iy, iz = (rng.integers(0, 1 << 21, n, dtype=np.uint32) for _ in range(3))
    for k, v in enumerate([0, (1 << 21) - 1, 1 << 20, (1 << 20) - 1]):
        ix[k], iy[k], iz[k] = v, v, v
        ix[4 + k], iy[4 + k], iz[4 + k] = v, 0, (1 << 21) - 1 - v
    key = np.zeros(n, np.uint64)
    state = np.zeros(n, np.uint32)  # entry & 0xFE00 of the previous lookup = next state << 9
    for g in range(6, -1, -1):
        idx = state | (((ix >> (3 * g)) & 7) << 6) | (((iy >> (3 * g)) & 7) << 3) | ((iz >> (3 * g)) & 7)
        e = tab[idx ^ ((idx >> 5) & 0x3E)].astype(np.uint32)
        key = (key << np.uint64(9)) | (e & 511).astype(np.uint64)
        state = e & 0xFE00
    assert (state >> 9).max() < 24
    assert np.array_equal(key, ref.hilbert_encode(ix, iy, iz))


def test_span_sfc_range(pkg, ref):
    """the octree cells a rank's glass-block assembly works through, against cstone::spanSfcRange"""
    import ctypes as C
    B = pkg.binding
    full = 1 << 63
    rng = np.random.Generator(np.random.MT19937(8))
    cases = [(0, full), (0, 1), (1, 2), (full - 1, full), (0o1234567 << 42, 0o1234571 << 42),
             (12345678901234567, 0o7 << 60), (0o3 << 60, (0o3 << 60) + (1 << 30))]
    for _ in range(200):
        a, b = sorted(int(v) for v in rng.integers(0, full, 2, dtype=np.uint64))
        if a < b:
            cases.append((a, b))
    for num_ranks in (2, 3, 7, 8):
        cases += [pkg.ic.equi_distant_sfc_segments(r, num_ranks) for r in range(num_ranks)]
    out = np.zeros(256, np.uint64)
    for a, b in cases:
        n = B.lib.sphb200_span_sfc_range(C.c_uint64(a), C.c_uint64(b), out.ctypes.data_as(C.c_void_p), C.c_size_t(256))
        want = ref.span_sfc_range(a, b)
        assert n == want.size - 1, (a, b)
        assert np.array_equal(out[:n + 1], want), (a, b)
        assert out[0] == a and out[n] == b and np.all(out[1:n + 1] > out[:n])
    assert B.lib.sphb200_span_sfc_range(C.c_uint64(5), C.c_uint64(5), out.ctypes.data_as(C.c_void_p), C.c_size_t(256)) == 0


def test_other_ics_are_seeded_and_inside_box(pkg):
    for case, side in (("noh", 20), ("turbulence", 16), ("kh", 32)):
        a, box, s = pkg.ic.CASES[case](side)
        b, _, _ = pkg.ic.CASES[case](side)
        for k in a:
            assert np.array_equal(a[k], b[k]), (case, k)
        for d, k in enumerate("xyz"):
            assert a[k].min() >= box[0][2 * d] and a[k].max() <= box[0][2 * d + 1]
        assert len({v.size for v in a.values()}) == 1
    noh, _, _ = pkg.ic.noh(20)
    r = np.sqrt(noh["x"] ** 2 + noh["y"] ** 2 + noh["z"] ** 2)
    assert r.max() <= 0.5
    assert np.allclose(noh["vx"] * noh["x"] + noh["vy"] * noh["y"] + noh["vz"] * noh["z"], -r, atol=1e-6)


def test_uniform_bins_and_assignment(pkg):
    """uniformBins / makeSfcAssignment (domaindecomp.hpp:49-125), cases of unit/domain/domaindecomp.cpp"""
    from sph_exa_b200 import dist as D
    bins, cnt = D.uniform_bins([4, 3, 4, 3, 4, 3, 4, 3, 4, 3], 2)
    assert list(bins) == [0, 5, 10] and list(cnt) == [18, 17]
    bins, cnt = D.uniform_bins([1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1], 4)
    assert list(bins) == [0, 4, 8, 12, 16] and list(cnt) == [4, 4, 4, 4]
    leaves = (np.arange(9, dtype=np.uint64) << np.uint64(60))
    b, c = D.make_sfc_assignment(2, [10, 0, 0, 0, 0, 0, 0, 30], leaves)
    assert b[0] == 0 and b[-1] == leaves[-1] and c.sum() == 40
    assert list(D.key_splits(np.array([1, 5, 9, 13], np.uint64), np.array([0, 6, 20], np.uint64))) == [0, 2, 4]


def test_halo_layout(pkg):
    from sph_exa_b200 import dist as D
    assert D.halo_layout(0, 10, [0, 4, 2]) == (0, 10, 16, [-1, 10, 14])
    assert D.halo_layout(1, 10, [3, 0, 2]) == (3, 13, 15, [0, -1, 13])
    assert D.halo_layout(2, 10, [3, 4, 0]) == (7, 17, 17, [0, 3, -1])
    assert D.partition_range(10, 0, 3) == (0, 4) and D.partition_range(10, 2, 3) == (7, 10)


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import sphexa_b200_loader
    pkg = sphexa_b200_loader.load()
    from sph_exa_b200 import dist as D
    from oracle import orc
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # each rank holds a slab of a Sedov lattice; global tree counts are summed over ranks (update_mpi.hpp:81-104)
    f, box, _ = pkg.ic.sedov(16)
    n = f["x"].size
    a, b = D.partition_range(n, rank, world)
    keys, _ = orc.sort_by_key(orc.sfc_keys(f["x"][a:b], f["y"][a:b], f["z"][a:b], box))
    all_keys, _ = orc.sort_by_key(orc.sfc_keys(f["x"], f["y"], f["z"], box))
    leaves, gcounts, _ = orc.compute_octree(all_keys, 64)
    local = orc.lib() and np.zeros(gcounts.size, np.uint32)
    orc.lib().orc_node_counts(orc.P(leaves), gcounts.size, orc.P(keys), __import__("ctypes").c_size_t(keys.size),
                              orc.P(local))
    t = torch.from_numpy(local.astype(np.int64))
    dist.all_reduce(t)
    assert np.array_equal(t.numpy(), gcounts.astype(np.int64)), "summed local counts != global counts"
    bounds, bc = D.make_sfc_assignment(world, t.numpy(), leaves)
    splits = D.key_splits(keys, bounds)
    send = torch.tensor([int(splits[q + 1] - splits[q]) for q in range(world)])
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send) if dist.get_backend() != "gloo" else None
    gathered = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(gathered, send)
    recv = torch.stack(gathered)[:, rank]
    # after migration every rank owns exactly the particles of its key range
    own = int(recv.sum())
    want = int(np.searchsorted(all_keys, bounds[rank + 1]) - np.searchsorted(all_keys, bounds[rank]))
    start, end, total, off = D.halo_layout(rank, own, [0] * world)
    q.put((rank, own, want, int(bc[rank]), (start, end, total)))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_bookkeeping_gloo(pkg):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    total = 0
    for rank, own, want, assigned, lay in res:
        assert own == want == assigned
        assert lay == (0, own, own)
        total += own
    assert total == 16 ** 3


def _migrate_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import sphexa_b200_loader
    sphexa_b200_loader.load()
    from sph_exa_b200 import dist as D
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def a2a(recvbuf, sendbuf, out_splits, in_splits):
        # all_to_all_single written with all_gather_object (gloo has no all-to-all)
        blocks, off = [], 0
        for c in in_splits:
            blocks.append(sendbuf[off:off + c].clone())
            off += c
        everyone = [None] * world
        dist.all_gather_object(everyone, blocks)
        off = 0
        for src in range(world):
            part = everyone[src][rank]
            assert part.numel() == out_splits[src]
            recvbuf[off:off + part.numel()] = part
            off += part.numel()

    # rank r holds sorted keys; boundaries move so that particles leave in both directions
    rng = np.random.Generator(np.random.MT19937(100 + rank))
    n = 1000 + 37 * rank
    keys = np.sort(rng.integers(0, 3000, n)).astype(np.int64)
    bounds = np.array([0, 900, 2100, 3000])          # key range of rank q: [bounds[q], bounds[q+1])
    splits = [int(np.searchsorted(keys, b)) for b in bounds[:-1]] + [n]
    send = [splits[k + 1] - splits[k] for k in range(world)]
    mat = [None] * world
    dist.all_gather_object(mat, send)
    recv = [mat[src][rank] for src in range(world)]
    f = {"key": torch.from_numpy(keys), "val": torch.from_numpy(keys * 2.5 + rank * 0.125),
         "tag": torch.full((n,), rank, dtype=torch.int32), "w": torch.from_numpy((keys % 7).astype(np.float64))}
    out = D.migrate_fields(torch, f, splits, send, recv, rank, a2a)
    k = out["key"].numpy()
    assert k.size == sum(recv)
    assert np.all((k >= bounds[rank]) & (k < bounds[rank + 1]))
    # the kept range comes first and is untouched; every arrival kept its own fields together
    kept = splits[rank + 1] - splits[rank]
    assert np.array_equal(k[:kept], keys[splits[rank]:splits[rank + 1]])
    tag = out["tag"].numpy()
    assert np.all(tag[:kept] == rank) and np.all(tag[kept:] != rank)
    assert np.array_equal(out["val"].numpy(), k * 2.5 + tag * 0.125)
    assert np.array_equal(out["w"].numpy(), (k % 7).astype(np.float64))
    totals = [None] * world
    dist.all_gather_object(totals, (n, int(k.size), int(k.sum())))
    q.put((rank, totals, int(keys.sum())))
    dist.barrier()
    dist.destroy_process_group()


def test_migration_moves_only_the_leaving_ranges_gloo(pkg):
    """dist.migrate_fields on 3 ranks (CPU tensors, all-to-all emulated over gloo): every particle ends on the rank
    owning its key, fields of a particle stay together, nothing is lost or duplicated"""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 90
    world = 3
    procs = [ctx.Process(target=_migrate_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    totals = res[0][1]
    assert sum(t[0] for t in totals) == sum(t[1] for t in totals)          # particle count conserved
    assert sum(t[2] for t in totals) == sum(r[2] for r in res)              # key checksum conserved
