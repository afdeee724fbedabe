"""CPU tests of the host logic of the multi-rank domain (csrc/domain.cu) against the reference's own functions compiled
into oracle/_ref: makeSfcAssignment + limitBoundaryShifts (domain/include/cstone/domain/domaindecomp.hpp:115-181) and
the tree a GlobalAssignment starts from (assignment.hpp:62-69)."""
import numpy as np
import pytest


def _random_tree(rng, ref, n=20000, bucket=16):
    keys = np.sort(rng.integers(0, 1 << 63, size=n, dtype=np.uint64))
    # cluster a third of the keys so that the leaf counts are uneven
    keys[: n // 3] = np.sort(rng.integers(1 << 61, (1 << 61) + (1 << 55), size=n // 3, dtype=np.uint64))
    keys = np.sort(keys)
    leaves, counts, _ = ref.compute_octree(keys, bucket)
    return leaves, counts


@pytest.mark.parametrize("num_ranks", [1, 2, 3, 5, 8, 13, 32])
def test_initial_global_tree_equals_reference(pkg, ref, num_ranks):
    from sph_exa_b200 import domain as D
    assert np.array_equal(D.initial_global_tree(num_ranks), ref.initial_global_tree(num_ranks))


@pytest.mark.parametrize("num_ranks", [2, 3, 8, 17])
def test_sfc_assignment_equals_reference(pkg, ref, num_ranks):
    from sph_exa_b200 import domain as D
    rng = np.random.default_rng(num_ranks)
    leaves, counts = _random_tree(rng, ref)
    b, c = D.sfc_assignment(num_ranks, counts, leaves)
    rb, rc = ref.sfc_assignment(num_ranks, counts, leaves)
    assert np.array_equal(b, rb) and np.array_equal(c, rc)
    assert b[0] == 0 and b[-1] == np.uint64(1 << 63) and int(c.sum()) == int(counts.sum())


@pytest.mark.parametrize("num_ranks", [3, 8])
def test_limit_boundary_shifts_equals_reference(pkg, ref, num_ranks):
    """the previous assignment pins every boundary between its two old neighbours (domaindecomp.hpp:140-181)"""
    from sph_exa_b200 import domain as D
    rng = np.random.default_rng(100 + num_ranks)
    leaves, counts = _random_tree(rng, ref)
    triggered = 0
    for trial in range(20):
        # an old assignment from a very different particle distribution: some new boundaries have to be clamped
        old_counts = counts.copy()
        a = int(rng.integers(0, counts.size - counts.size // 10))
        old_counts[a:a + counts.size // 10] *= np.uint32(200)  # most of the old particles sat in a tenth of the leaves
        old, _ = ref.sfc_assignment(num_ranks, old_counts, leaves)
        b, c = D.sfc_assignment(num_ranks, counts, leaves, old)
        rb, rc = ref.sfc_assignment(num_ranks, counts, leaves, old)
        assert np.array_equal(b, rb) and np.array_equal(c, rc)
        free, _ = ref.sfc_assignment(num_ranks, counts, leaves)
        triggered += int(not np.array_equal(free, rb))
        for r in range(1, num_ranks):
            assert old[r - 1] <= b[r] <= old[r + 1]
    assert triggered > 0, "no trial exercised the clamp"
