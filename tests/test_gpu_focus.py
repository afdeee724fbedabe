"""The focus tree (csrc/focus.cu) against the reference's own CPU code compiled into
oracle/_ref (oracle/ref_focus.cpp): computeMinMacR2, markMacs, rebalanceDecisionEssential, enforceKeys,
protectAncestors operator by operator on the same linked octree, and the tree object update by update against
cstone::FocusedOctreeSingleNode (focus/octree_focus.hpp:219-288) -- leaves, counts and the convergence flag of EVERY
update, bit for bit -- for the SFC segments of 2 to 8 ranks with their peers' boundaries as mandatory keys, moving
particles, a moving focus and keys that have to be injected."""
import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu

from test_gpu_domain import BOXES, random_coords  # noqa: E402

ROOT_END = np.uint64(1 << 63)


def sorted_keys(ref, n, box, seed, gaussian):
    x, y, z = random_coords(n, box, seed, gaussian)
    return np.sort(ref.compute_sfc_keys(x, y, z, box))


def linked(pkg, ctx, ref, keys, bucket):
    from sph_exa_b200 import ops
    leaves, counts, _ = ref.compute_octree(keys, bucket)
    host = ref.build_octree(leaves)
    dev = ops.Octree(ctx, to_dev(leaves))
    assert np.array_equal(to_host(dev.prefixes, np.uint64), host["prefixes"])
    return leaves, counts, host, dev


def node_counts(host, leaves, counts):
    """counts of all nodes of a linked octree (numpy; only used to build inputs)"""
    layout = np.zeros(counts.size + 1, np.int64)
    np.cumsum(counts, out=layout[1:])
    p = host["prefixes"]
    plen = np.array([int(v).bit_length() - 1 for v in p])
    start = np.array([((int(v) ^ (1 << int(l))) << (63 - int(l))) for v, l in zip(p, plen)], np.uint64)
    end = start + np.array([1 << (63 - int(l)) for l in plen], np.uint64)
    a = np.searchsorted(leaves, start)
    b = np.searchsorted(leaves, end)
    return (layout[b] - layout[a]).astype(np.uint32)


def prefix_start(prefix):
    """decodePlaceholderBit"""
    p = int(prefix)
    length = p.bit_length() - 1
    return (p ^ (1 << length)) << (63 - length)


def segments(ref, leaves, counts, num_ranks):
    bounds, _ = ref.sfc_assignment(num_ranks, counts, leaves)
    return bounds


@pytest.mark.parametrize("gaussian", [False, True])
def test_mac_operators_equal_reference(pkg, ctx, ref, gaussian):
    import torch
    from sph_exa_b200 import focus as F
    n = 60_000
    for bi, box in enumerate(BOXES):
        keys = sorted_keys(ref, n, box, 11 + bi, gaussian)
        leaves, counts, host, dev = linked(pkg, ctx, ref, keys, 32)
        nn = host["numNodes"]
        b = pkg.make_box(*box)
        bounds = segments(ref, leaves, counts, 5)
        for theta in (0.4, 0.75, 1.0):
            inv = F.invThetaMinMac(theta)
            want_c = ref.min_mac_r2(host["prefixes"], inv, box)
            c4 = torch.empty((nn, 4), dtype=torch.float64, device="cuda")
            F.minMacR2(ctx, dev.prefixes, inv, b, c4)
            assert np.array_equal(to_host(c4), want_c)
            for r in range(5):
                fs, fe = bounds[r], bounds[r + 1]
                want = ref.mark_macs(host, want_c, box, fs, fe)
                got = torch.full((nn,), 7, dtype=torch.int8, device="cuda")
                F.markMacs(ctx, dev.prefixes, dev.parents, c4, b, fs, fe, got)
                assert np.array_equal(to_host(got), want), (box, theta, r)
                assert want[0] == 1 and 0 < want.sum() < nn
        # a focus that is a single finest-level cell, the whole key range, and an unaligned range
        for fs, fe in ((keys[100], keys[100] + np.uint64(1)), (np.uint64(0), ROOT_END),
                       (keys[500] | np.uint64(5), keys[40_000] | np.uint64(3))):
            want = ref.mark_macs(host, want_c, box, fs, fe)
            got = torch.empty((nn,), dtype=torch.int8, device="cuda")
            F.markMacs(ctx, dev.prefixes, dev.parents, c4, b, fs, fe, got)
            assert np.array_equal(to_host(got), want)


@pytest.mark.parametrize("gaussian", [False, True])
def test_rebalance_enforce_protect_equal_reference(pkg, ctx, ref, gaussian):
    import torch
    from sph_exa_b200 import focus as F
    rng = np.random.Generator(np.random.MT19937(5))
    n = 80_000
    box = BOXES[1]
    keys = sorted_keys(ref, n, box, 3, gaussian)
    # a tree that is too fine in places and too coarse in others for bucket 64
    leaves, counts, host, dev = linked(pkg, ctx, ref, np.sort(np.concatenate([keys[::8], keys[:4000]])), 16)
    leaves_counts = np.zeros(leaves.size - 1, np.uint32)
    ref_counts = np.searchsorted(keys, leaves)
    leaves_counts[:] = np.diff(ref_counts)
    counts_all = node_counts(host, leaves, leaves_counts)
    nn = host["numNodes"]
    bounds = segments(ref, leaves, leaves_counts, 4)
    inv = F.invThetaMinMac(0.5)
    c4 = ref.min_mac_r2(host["prefixes"], inv, box)
    statuses, kinds = set(), set()
    for r in range(4):
        fs, fe = bounds[r], bounds[r + 1]
        macs = ref.mark_macs(host, c4, box, fs, fe)
        for bucket in (64, 8):
            want = ref.rebalance_decision_essential(host, counts_all, macs, fs, fe, bucket)
            ops = torch.full((nn,), -1, dtype=torch.int32, device="cuda")
            F.rebalanceDecisionEssential(ctx, dev.prefixes, dev.childOffsets, dev.parents, to_dev(counts_all),
                                         to_dev(macs), fs, fe, bucket, ops)
            assert np.array_equal(to_host(ops), want)
            kinds |= set(np.unique(want).tolist())
            # mandatory keys: the segment, all rank boundaries, keys one level below a leaf, keys deep inside a
            # coarse far-away leaf (status failed) and the ends of the key range
            mand = [np.array([fs, fe], np.uint64), np.concatenate([[fs, fe], bounds]).astype(np.uint64)]
            far = (r + 2) % 4
            i = int(np.searchsorted(leaves, bounds[far])) + 3
            one_below = leaves[i] + (leaves[i + 1] - leaves[i]) // np.uint64(8) * np.uint64(3)
            mand.append(np.array([fs, fe, one_below], np.uint64))
            deep = leaves[i] + (leaves[i + 1] - leaves[i]) // np.uint64(512) * np.uint64(77)
            mand.append(np.array([fs, fe, deep, one_below, 0, 1 << 63], np.uint64))
            mand.append(rng.choice(keys, 20).astype(np.uint64))
            merged = np.nonzero(want == 0)[0]
            if merged.size:  # keys of nodes that are about to be merged away: ResolutionStatus::cancelMerge
                mand.append(np.array([prefix_start(host["prefixes"][i]) for i in merged[[1, merged.size // 2, -1]]],
                                     np.uint64))
            for mk in mand:
                w = want.copy()
                st_want = ref.enforce_keys(mk, host, w)
                g = to_dev(want.copy())
                st = F.enforceKeys(ctx, to_dev(mk), dev.prefixes, dev.childOffsets, dev.parents, g)
                assert st == st_want
                assert np.array_equal(to_host(g), w)
                statuses.add(st)
                conv_want = ref.protect_ancestors(host, w)
                conv = F.protectAncestors(ctx, dev.prefixes, dev.parents, g)
                assert conv == conv_want
                assert np.array_equal(to_host(g), w)
    assert kinds == {0, 1, 8}
    assert statuses == {F.CONVERGED, F.CANCEL_MERGE, F.REBALANCE, F.FAILED}


def run_both(pkg, ctx, ref, box, bucket, theta, schedule):
    """schedule: list of (keys, focusStart, focusEnd, mandatoryKeys); returns the number of updates compared"""
    from sph_exa_b200 import focus as F
    theirs = ref.FocusedOctreeSingleNode(bucket, theta)
    ours = F.FocusedOctree(ctx, bucket, theta)
    b = pkg.make_box(*box)
    flags = []
    for step, (keys, fs, fe, mand) in enumerate(schedule):
        dkeys = to_dev(keys)
        want_conv = theirs.update(box, keys, fs, fe, mand)
        conv = ours.update(b, dkeys, fs, fe, mand)
        wl, wc = theirs.leaves_counts()
        gl, gc = ours.treeLeaves(), ours.leafCounts()
        assert np.array_equal(gl, wl), (step, gl.size, wl.size)
        assert np.array_equal(gc, wc), step
        assert conv == want_conv, step
        flags.append(conv)
    ours.close()
    return flags, wl, wc


@pytest.mark.parametrize("case", [(1, False, 0.5, 64, 4), (1, True, 0.5, 64, 8), (0, True, 0.75, 16, 3),
                                  (2, False, 1.0, 64, 2), (3, True, 0.6, 32, 5), (4, False, 0.5, 64, 4)])
def test_focus_tree_equals_reference_update_by_update(pkg, ctx, ref, case):
    """every rank's focus tree from the root node to convergence: FocusedOctreeSingleNode::update each time"""
    from sph_exa_b200 import focus as F
    bi, gaussian, theta, bucket, num_ranks = case
    box = BOXES[bi]
    keys = sorted_keys(ref, 120_000, box, 40 + bi, gaussian)
    gl, gcnt, _ = ref.compute_octree(keys, max(bucket, keys.size // (100 * num_ranks)))
    bounds = segments(ref, gl, gcnt, num_ranks)
    b = pkg.make_box(*box)
    for r in range(num_ranks):
        fs, fe = bounds[r], bounds[r + 1]
        peers = F.findPeersMac(r, bounds, gl, b, F.invThetaMinMac(theta))
        mand = np.array([bounds[p + d] for p in peers for d in (0, 1)], np.uint64)
        flags, leaves, counts = run_both(pkg, ctx, ref, box, bucket, theta, [(keys, fs, fe, mand)] * 9)
        assert flags[-1] and not flags[0]
        assert int(counts.sum()) == keys.size
        i0, i1 = np.searchsorted(leaves, [fs, fe])
        assert leaves[i0] == fs and leaves[i1] == fe
        # inside the focus the tree resolves the bucket size; far outside (clustered particles) it is coarser than that
        assert counts[i0:i1].max() <= bucket
        if gaussian:
            assert np.delete(counts, np.s_[i0:i1]).max() > bucket


def test_focus_tree_follows_moving_particles_focus_and_injected_keys(pkg, ctx, ref):
    box = BOXES[1]
    rng = np.random.Generator(np.random.MT19937(77))
    n = 100_000
    x, y, z = random_coords(n, box, 9, True)
    schedule = []
    third = np.uint64((1 << 63) // 3)
    fs, fe = third & ~np.uint64((1 << 45) - 1), (np.uint64(2) * third) & ~np.uint64((1 << 45) - 1)
    for step in range(24):
        keys = np.sort(ref.compute_sfc_keys(x, y, z, box))
        mand = []
        if step == 8:   # the focus moves by a sixth of the key range (focusTransfer happens in the caller)
            fs, fe = fs + (fe - fs) // np.uint64(2), fe + (fe - fs) // np.uint64(2)
            fs, fe = fs & ~np.uint64((1 << 45) - 1), fe & ~np.uint64((1 << 45) - 1)
        if step in (5, 12, 13):  # finest-level keys far from the focus: ResolutionStatus::failed -> injectKeys
            mand = [int(keys[10]) | 1, int(keys[n - 5]) & ~7, int(keys[n // 2])]
        schedule.append((keys, fs, fe, np.array(mand, np.uint64)))
        for a, lim in zip((x, y, z), (box[0][0:2], box[0][2:4], box[0][4:6])):
            a += rng.normal(0, 0.004, n)
            np.clip(a, lim[0], np.nextafter(lim[1], lim[0]), out=a)
    flags, leaves, counts = run_both(pkg, ctx, ref, box, 64, 0.5, schedule)
    assert len(flags) == 24 and int(counts.sum()) == n


def test_staged_calls_and_reduced_counts_equal_update(pkg, ctx, ref):
    """updateTree + updateMinMac + updateCounts == update; leaf counts summed over 3 'ranks' and handed in with
    setLeafCounts (what a P-rank run does after its all-reduce) give the same tree as counting all keys at once"""
    import torch
    from sph_exa_b200 import focus as F, ops
    box = BOXES[2]
    b = pkg.make_box(*box)
    keys = sorted_keys(ref, 90_000, box, 21, True)
    dkeys = to_dev(keys)
    parts = [to_dev(np.ascontiguousarray(keys[i::3])) for i in range(3)]
    fs, fe = np.uint64(1 << 61), np.uint64(3 << 60) + np.uint64(1 << 61)
    one, staged, reduced = (F.FocusedOctree(ctx, 64, 0.5) for _ in range(3))
    for step in range(12):
        c1 = one.update(b, dkeys, fs, fe)
        c2 = staged.updateTree(fs, fe)
        staged.updateMinMac(b, fs, fe)
        staged.updateCounts(dkeys)
        c3 = reduced.updateTree(fs, fe)
        reduced.updateMinMac(b, fs, fe)
        v = reduced.view()
        total = torch.zeros(v.numLeaves, dtype=torch.int32, device="cuda")
        lv = to_dev(reduced.treeLeaves())
        for p in parts:
            total += ops.computeNodeCountsGpu(ctx, lv, p)
        reduced.setLeafCounts(total)
        assert c1 == c2 == c3
        for other in (staged, reduced):
            assert np.array_equal(other.treeLeaves(), one.treeLeaves())
            assert np.array_equal(other.leafCounts(), one.leafCounts())
            a, o = one.nodeArrays(), other.nodeArrays()
            for k in ("counts", "macs", "centers", "prefixes"):
                assert np.array_equal(a[k], o[k]), k
    assert c1
    with pytest.raises(pkg.SphError):
        bad = F.FocusedOctree(ctx, 64, 0.5)
        bad.updateTree(fs, fe)
        bad.updateTree(fs, fe)  # counts and MACs of the new tree are missing


@pytest.mark.parametrize("case,side,P,theta", [("sedov", 24, 3, 1.0), ("noh", 30, 2, 0.5), ("turbulence", 24, 4, 1.0),
                                               ("sedov", 32, 8, 0.6)])
def test_ranks_of_a_domain_build_the_reference_focus_trees(pkg, ctx, ref, case, side, P, theta):
    """sphb200_domain_update_focus on P virtual ranks: after the first call (FocusedOctree::converge) every rank's
    tree, with counts summed over the ranks, IS the tree cstone::FocusedOctreeSingleNode converges to on the global
    particle set for that rank's segment and peers; over the following steps (moving particles, shifting assignment)
    it follows the device object driven in Domain::sync's order on the global keys"""
    from sph_exa_b200 import focus as F
    from test_gpu_domain_dist import make_group
    group, comm, f, box, s = make_group(pkg, ctx, case, side, P)
    try:
        group.sync()
        conv = group.updateFocus(theta)
        assert conv
        infos = [r.info() for r in group.ranks]
        bounds = np.array(infos[0].assignment[:P + 1], np.uint64)
        keys = group.download(names=["x"])["keys"]
        gbox = (group.ranks[0].boxLimits(), list(box[1]))
        b = pkg.make_box(*gbox)
        gleaves = group.ranks[0].globalLeaves()
        twins = []
        for r, rank in enumerate(group.ranks):
            leaves, counts, peers = rank.focusTree()
            assert peers == ref.find_peers_mac(r, bounds, gleaves, gbox, F.invThetaMinMac(theta))
            mand = np.array([bounds[p + d] for p in peers for d in (0, 1)], np.uint64)
            theirs = ref.FocusedOctreeSingleNode(64, theta)
            for _ in range(40):
                if theirs.update(gbox, keys, bounds[r], bounds[r + 1], mand):
                    break
            wl, wc = theirs.leaves_counts()
            assert np.array_equal(leaves, wl), (r, leaves.size, wl.size)
            assert np.array_equal(counts, wc), r
            assert int(counts.sum()) == keys.size
            # a single-GPU twin of the rank's tree, brought to the same state with the global keys
            twin = F.FocusedOctree(ctx, 64, theta)
            dk = to_dev(keys)
            for _ in range(40):
                twin.updateMinMac(b, bounds[r], bounds[r + 1])
                done = twin.updateTree(bounds[r], bounds[r + 1], mand)
                twin.updateCounts(dk)
                if done:
                    break
            assert np.array_equal(twin.treeLeaves(), leaves)
            twins.append(twin)
        for step in range(3):
            group.step()
            group.sync()
            group.updateFocus(theta)
            infos = [r.info() for r in group.ranks]
            bounds = np.array(infos[0].assignment[:P + 1], np.uint64)
            keys = group.download(names=["x"])["keys"]
            gbox = (group.ranks[0].boxLimits(), list(box[1]))
            b = pkg.make_box(*gbox)
            dk = to_dev(keys)
            for r, rank in enumerate(group.ranks):
                leaves, counts, peers = rank.focusTree()
                mand = np.array([bounds[p + d] for p in peers for d in (0, 1)], np.uint64)
                twins[r].updateMinMac(b, bounds[r], bounds[r + 1])
                twins[r].updateTree(bounds[r], bounds[r + 1], mand)
                twins[r].updateCounts(dk)
                assert np.array_equal(leaves, twins[r].treeLeaves()), (step, r)
                assert np.array_equal(counts, twins[r].leafCounts()), (step, r)
        for t in twins:
            t.close()
    finally:
        group.close()
        comm.close()


@pytest.mark.parametrize("kind", ["empty", "one", "coincident", "whole_range"])
def test_focus_tree_edge_cases(pkg, ctx, ref, kind):
    """no particles, one particle, 200 coincident particles (the tree descends to the deepest level, one level per
    update) and a focus that is the whole key range (one rank): every update equals the reference's"""
    box = BOXES[1]
    full = sorted_keys(ref, 30_000, box, 5, True)
    if kind == "empty":
        keys, fs, fe = np.zeros(0, np.uint64), np.uint64(1 << 60), np.uint64(1 << 61)
    elif kind == "one":
        keys, fs, fe = full[:1].copy(), np.uint64(0), np.uint64(1 << 62)
    elif kind == "coincident":
        keys = np.sort(np.concatenate([full[:2000], np.full(200, full[777], np.uint64)]))
        fs, fe = np.uint64(0), np.uint64(1 << 62)
    else:
        keys, fs, fe = full, np.uint64(0), ROOT_END
    updates = 26 if kind == "coincident" else 8
    flags, leaves, counts = run_both(pkg, ctx, ref, box, 64, 0.5, [(keys, fs, fe, np.zeros(0, np.uint64))] * updates)
    assert int(counts.sum()) == keys.size
    if kind == "coincident":
        assert np.diff(leaves).min() == 1 and flags[-1]  # a finest-level cell holds the coincident particles
    if kind == "whole_range":
        assert flags[-1] and counts.max() <= 64
