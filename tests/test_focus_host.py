"""CPU test of the host part of the focus tree (csrc/focus.cu): sphb200_find_peers_mac against the reference's own
findPeersMac (domain/include/cstone/traversal/peers.hpp:63-120) compiled into oracle/_ref, on the trees and assignments
the domain produces (makeSfcAssignment on uniform, clustered and coarse global trees; periodic, open and anisotropic
boxes; opening angles from 0.3 to 1.0)."""
import numpy as np
import pytest

BOXES = [
    ([0.0, 1.0, 0.0, 1.0, 0.0, 1.0], [0, 0, 0]),
    ([-0.5, 0.5, -0.5, 0.5, -0.5, 0.5], [1, 1, 1]),
    ([0.0, 1.0, 0.0, 2.0, 0.0, 3.0], [1, 0, 1]),
    ([0.0, 1.0, 0.0, 1.0, 0.0, 0.0625], [1, 1, 1]),
]


def _tree(rng, ref, kind, n, bucket):
    if kind == "uniform":
        keys = rng.integers(0, 1 << 63, size=n, dtype=np.uint64)
    elif kind == "clustered":
        keys = rng.integers(0, 1 << 63, size=n, dtype=np.uint64)
        keys[: n // 2] = rng.integers(3 << 60, (3 << 60) + (1 << 54), size=n // 2, dtype=np.uint64)
    else:  # coarse: the tree a GlobalAssignment starts from
        return ref.initial_global_tree(8), None
    keys = np.sort(keys)
    leaves, counts, _ = ref.compute_octree(keys, bucket)
    return leaves, counts


@pytest.mark.parametrize("kind", ["uniform", "clustered", "coarse"])
@pytest.mark.parametrize("num_ranks", [2, 5, 8, 16])
def test_find_peers_mac_equals_reference(pkg, ref, kind, num_ranks):
    from sph_exa_b200 import focus as F
    rng = np.random.default_rng(7 * num_ranks + len(kind))
    leaves, counts = _tree(rng, ref, kind, 30000, 64)
    if counts is None:
        counts = rng.integers(1, 100, size=leaves.size - 1).astype(np.uint32)
    bounds, _ = ref.sfc_assignment(num_ranks, counts, leaves)
    mutual = 0
    for box in BOXES:
        for theta in (0.3, 0.5, 1.0):
            inv = F.invThetaMinMac(theta)
            lists = []
            for rank in range(num_ranks):
                got = F.findPeersMac(rank, bounds, leaves, pkg.make_box(*box), inv)
                want = ref.find_peers_mac(rank, bounds, leaves, box, inv)
                assert got == want, (kind, num_ranks, box, theta, rank)
                assert rank not in got
                lists.append(got)
            # the reference guarantees mutuality (peers.hpp:57-58)
            for a in range(num_ranks):
                for b in lists[a]:
                    assert a in lists[b]
                    mutual += 1
    assert mutual > 0


def test_find_peers_mac_degenerate_cases(pkg, ref):
    from sph_exa_b200 import focus as F
    box = BOXES[1]
    leaves = ref.initial_global_tree(4)
    bounds = np.array([0, 1 << 63], np.uint64)
    assert F.findPeersMac(0, bounds, leaves, pkg.make_box(*box), 2.5) == []
    # an empty segment has no peers
    counts = np.ones(leaves.size - 1, np.uint32)
    b4, _ = ref.sfc_assignment(4, counts, leaves)
    b5 = np.concatenate([b4[:2], b4[1:]])
    assert F.findPeersMac(1, b5, leaves, pkg.make_box(*box), 2.5) == []
    # root-only tree
    root = np.array([0, 1 << 63], np.uint64)
    assert F.findPeersMac(0, np.array([0, 1 << 63], np.uint64), root, pkg.make_box(*box), 2.5) == \
        ref.find_peers_mac(0, np.array([0, 1 << 63], np.uint64), root, box, 2.5)
