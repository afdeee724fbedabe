"""Glass-block initial conditions on the GPU against the reference's own code (oracle/_ref).

Reference: main/src/init/grid.hpp (`assembleCuboid` :238-264 with `spanSfcRange`, `createFpBox`, `gridIntersection`,
`extractBlock`/`scaleBlockToGlobal` :152-222; `cutSphere` :267-300; `equiDistantSfcSegments` :80-88) as used by the
glass initialisers (sedov_init.hpp:155-190, noh_init.hpp:110-150).  Coordinates and their order are compared bit for
bit.
"""
import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu

FULL = 1 << 63
BOXES = [
    ([-0.5, 0.5, -0.5, 0.5, -0.5, 0.5], [1, 1, 1]),
    ([0.0, 1.0, 0.0, 1.0, 0.0, 0.0625], [1, 1, 1]),  # Kelvin-Helmholtz slab
    ([-1.2, 1.5, 0.3, 0.9, -2.0, 4.0], [0, 1, 0]),
]


def block(n, seed):
    rng = np.random.Generator(np.random.MT19937(seed))
    return [np.ascontiguousarray(rng.random(n)) for _ in range(3)]


def assemble(pkg, ctx, a, b, box, m, blk):
    from sph_exa_b200 import ops
    out = ops.assembleCuboid(ctx, a, b, pkg.make_box(*box), m, *[to_dev(v) for v in blk])
    return [to_host(v, np.float64) for v in out]


@pytest.mark.parametrize("box", BOXES)
@pytest.mark.parametrize("m", [(1, 1, 1), (3, 3, 3), (2, 5, 1)])
def test_assemble_whole_box(pkg, ctx, ref, box, m):
    blk = block(1000, 3)
    want = ref.assemble_cuboid(0, FULL, box, m, *blk)
    got = assemble(pkg, ctx, 0, FULL, box, m, blk)
    assert want[0].size == 1000 * m[0] * m[1] * m[2]
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


@pytest.mark.parametrize("num_ranks", [2, 3, 8, 13])
def test_assemble_rank_segments(pkg, ctx, ref, num_ranks):
    """every rank's equiDistantSfcSegments slice equals the reference's, and the slices partition the tiled box"""
    box, m = BOXES[0], (4, 4, 4)
    blk = block(777, 5)
    total, seen = 0, []
    prev_end = 0
    for rank in range(num_ranks):
        a, b = pkg.ic.equi_distant_sfc_segments(rank, num_ranks)
        assert a == prev_end and b > a
        prev_end = b
        want = ref.assemble_cuboid(a, b, box, m, *blk)
        got = assemble(pkg, ctx, a, b, box, m, blk)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        total += got[0].size
        seen.append(np.stack(got, axis=1))
    assert prev_end == FULL
    assert total == 777 * 64
    allp = np.concatenate(seen)
    assert np.unique(allp, axis=0).shape[0] == total


@pytest.mark.parametrize("a,b", [(0o1234567 << 42, 0o1234571 << 42), (12345678901234567, 0o7 << 60),
                                 (1, 2), (0o3 << 60, (0o3 << 60) + (1 << 30))])
def test_assemble_arbitrary_key_ranges(pkg, ctx, ref, a, b):
    """ranges that need many small cells (spanSfcRange, sfc/common.hpp:362-400), down to a single finest-level cell"""
    box, m = BOXES[2], (3, 2, 4)
    blk = block(2000, 7)
    want = ref.assemble_cuboid(a, b, box, m, *blk)
    got = assemble(pkg, ctx, a, b, box, m, blk)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_assemble_capacity_and_empty(pkg, ctx):
    import ctypes as C
    from sph_exa_b200 import binding as B
    blk = [to_dev(v) for v in block(100, 1)]
    bx = pkg.make_box(*BOXES[0])
    m = (C.c_int * 3)(2, 2, 2)
    count = C.c_uint64(0)
    rc = B.lib.sphb200_assemble_cuboid(ctx.h, 0, C.c_uint64(FULL), C.byref(bx), m, B.ptr(blk[0]), B.ptr(blk[1]),
                                       B.ptr(blk[2]), 100, None, None, None, 0, C.byref(count))
    assert rc == B.ERR_CAPACITY and count.value == 800
    rc = B.lib.sphb200_assemble_cuboid(ctx.h, C.c_uint64(5), C.c_uint64(5), C.byref(bx), m, B.ptr(blk[0]),
                                       B.ptr(blk[1]), B.ptr(blk[2]), 100, None, None, None, 0, C.byref(count))
    assert rc == 0 and count.value == 0


@pytest.mark.parametrize("n", [1, 255, 256, 100_003])
def test_cut_sphere(pkg, ctx, ref, n):
    from sph_exa_b200 import ops
    rng = np.random.Generator(np.random.MT19937(n))
    x, y, z = (np.ascontiguousarray(rng.uniform(-0.5, 0.5, n)) for _ in range(3))
    x[0], y[0], z[0] = 0.5, 0.0, 0.0  # on the sphere: kept (<=)
    want = ref.cut_sphere(0.5, x, y, z)
    got = [to_host(v, np.float64) for v in ops.cutSphere(ctx, 0.5, to_dev(x), to_dev(y), to_dev(z))]
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_glass_initialisers_run(pkg, ctx, ref, tmp_path):
    """SedovGlass / NohGlassSphere from a block file: particle counts, and one step of the propagator on the result"""
    from sph_exa_b200 import sim as S
    blk = block(512, 11)
    path = tmp_path / "glass.bin"
    pkg.ic.write_template_block(path, *blk)
    back = pkg.ic.read_template_block(path)
    for a, b in zip(back, blk):
        assert np.array_equal(a, b)

    f, box, s = pkg.ic.sedov_glass(ctx, back, cbrt_num_part=16)
    assert f["x"].size == 8 * 512 and s["numParticlesGlobal"] == 8 * 512
    want = ref.assemble_cuboid(0, FULL, box, (2, 2, 2), *blk)
    assert np.array_equal(f["x"], want[0])
    parts = [pkg.ic.sedov_glass(ctx, back, 16, rank=r, num_ranks=2)[0]["x"].size for r in range(2)]
    assert sum(parts) == 8 * 512

    f, box, s = pkg.ic.noh_glass(ctx, back, cbrt_num_part=24)
    n = f["x"].size
    assert 0.45 * 27 * 512 < n < 0.6 * 27 * 512  # pi/6 of the cube
    assert np.all(np.sqrt(f["x"] ** 2 + f["y"] ** 2 + f["z"] ** 2) <= 0.5)
    p = S.HydroVeProp(ctx, box, s, capacity=n)
    p.upload(f)
    p.step()
    assert p.minDt > 0
    p.close()
