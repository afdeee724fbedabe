"""Block-time-step building blocks on the GPU against the numpy restatement in oracle/bdt.py (the reference has only
CUDA kernels for these: sph/include/sph/ts_groups.cu, cstone/traversal/groups.cuh; the same library functions are
compared with those kernels themselves, built for sm_100a, in tests/test_gpu_reference_kernels.py)."""
import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu


def sorted_case(pkg, ctx, ref, n, box, seed, gaussian, bucket=64):
    from sph_exa_b200 import ops
    from test_gpu_domain import random_coords
    x, y, z = random_coords(n, box, seed, gaussian)
    keys, perm = ref.sort_by_key(ref.compute_sfc_keys(x, y, z, box))
    x, y, z = x[perm], y[perm], z[perm]
    leaves, counts, _ = ref.compute_octree(keys, bucket)
    layout = np.zeros(counts.size + 1, np.uint32)
    np.cumsum(counts, out=layout[1:])
    tree = ops.Octree(ctx, to_dev(leaves), pkg.make_box(*box), to_dev(counts))
    return x, y, z, leaves, layout, tree


@pytest.mark.parametrize("gaussian", [False, True])
@pytest.mark.parametrize("rng_", [(0, None), (1000, 54321)])
def test_compute_groups(pkg, ctx, ref, gaussian, rng_):
    from oracle import bdt
    from sph_exa_b200 import ops
    from test_gpu_domain import BOXES
    n = 100_003
    for box in (BOXES[1], BOXES[2]):
        x, y, z, leaves, layout, tree = sorted_case(pkg, ctx, ref, n, box, 3, gaussian)
        first, last = rng_[0], (n if rng_[1] is None else rng_[1])
        want = bdt.compute_groups(first, last, x, y, z, leaves, layout, box[0])
        g = ops.computeGroups(ctx, first, last, to_dev(x), to_dev(y), to_dev(z), tree, pkg.make_box(*box))
        start, end = to_host(g.start, np.uint32), to_host(g.end, np.uint32)
        assert np.array_equal(start, want[:-1])
        assert np.array_equal(end, want[1:])
        sizes = end - start
        assert sizes.min() >= 1 and sizes.max() <= 32 and sizes.sum() == last - first
        if gaussian:
            assert g.numGroups > (last - first + 31) // 32  # the tails of the distribution do get split


def test_group_timesteps_rungs_extract_targets(pkg, ctx, ref):
    import torch
    from oracle import bdt
    from sph_exa_b200 import ops
    from test_gpu_domain import BOXES
    n = 20_011
    box = BOXES[1]
    x, y, z, leaves, layout, tree = sorted_case(pkg, ctx, ref, n, box, 5, True)
    first, last = 100, n - 7
    g = ops.computeGroups(ctx, first, last, to_dev(x), to_dev(y), to_dev(z), tree, pkg.make_box(*box))
    start, end = to_host(g.start, np.uint32), to_host(g.end, np.uint32)
    rng = np.random.Generator(np.random.MT19937(9))
    for real in (np.float64, np.float32):
        divv = rng.normal(size=n).astype(real)
        ax, ay, az = (rng.normal(size=n).astype(real) * 10 for _ in range(3))
        dtp = rng.uniform(1e-5, 1e-3, size=n).astype(real)
        gdt0 = np.full(g.numGroups, np.finfo(np.float32).max, np.float32)
        gdt = to_dev(gdt0)
        ops.groupDivvTimestep(ctx, 0.06, g.view(), to_dev(divv), gdt)
        w1 = bdt.group_divv_timestep(0.06, start, end, divv, gdt0)
        assert np.array_equal(to_host(gdt), w1)
        ops.groupAccTimestep(ctx, 0.3, g.view(), to_dev(ax), to_dev(ay), to_dev(az), gdt)
        w2 = bdt.group_acc_timestep(0.3, start, end, ax, ay, az, w1)
        got = to_host(gdt)
        assert np.allclose(got, w2, rtol=3e-7, atol=0)  # |a|^2 may contract to FMAs on the device
        ops.groupMinTimestep(ctx, g.view(), to_dev(dtp), gdt)
        assert np.array_equal(to_host(gdt), bdt.group_min_timestep(start, end, dtp, got))
    # a sliced view only touches its groups
    gdt = to_dev(gdt0)
    ops.groupMinTimestep(ctx, g.view(10, 20), to_dev(dtp), gdt[10:])
    got = to_host(gdt)
    assert np.all(got[:10] == gdt0[:10]) and np.all(got[20:] == gdt0[20:]) and np.all(got[10:20] < 1)
    # rungs, extraction, target lists
    rungs = torch.zeros(n, dtype=torch.uint8, device="cuda")
    order = rng.permutation(g.numGroups).astype(np.uint32)
    sub = ops.extractGroups(ctx, g, to_dev(order), 5, 205)
    assert np.array_equal(to_host(sub.start, np.uint32), start[order[5:205]])
    assert np.array_equal(to_host(sub.end, np.uint32), end[order[5:205]])
    ops.storeRung(ctx, sub.view(), 3, rungs)
    want = np.zeros(n, np.uint8)
    member = np.zeros(n, bool)
    for s, e in zip(start[order[5:205]], end[order[5:205]]):
        want[s:e] = 3
        member[s:e] = True
    assert np.array_equal(rungs.cpu().numpy(), want)
    t = ops.targetsFromGroups(ctx, sub.view())
    assert np.array_equal(to_host(t, np.uint32), np.nonzero(member)[0])
    assert ops.targetsFromGroups(ctx, sub.view(0, 0)).numel() == 0
    allt = ops.targetsFromGroups(ctx, g.view())
    assert np.array_equal(to_host(allt, np.uint32), np.arange(first, last))
