"""sphb200_layout_noalloc: the exclusive scan of sphb200_layout with its tile sums in a static __device__ array
(no cudaMalloc on the call path).  Collected last on purpose: a failure here must not hide
other tests.  Compared with the existing entry point through the C ABI, bit for bit (integer work), on calls in a row
that share the static scratch."""
import ctypes as C

import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu

MAX = 16777216  # SPHB200_LAYOUT_NOALLOC_MAX


def _both(pkg, ctx, counts_h):
    import torch
    from sph_exa_b200.binding import lib, ptr, check
    n = counts_h.size
    counts = to_dev(counts_h)
    a = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda")
    b = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda")
    check(lib.sphb200_layout(ctx.h, ptr(counts), n, ptr(a)))
    check(lib.sphb200_layout_noalloc(ctx.h, ptr(counts), n, ptr(b)))
    ctx.sync()
    return to_host(a, np.uint32), to_host(b, np.uint32)


def test_layout_noalloc_equals_layout_on_calls_in_a_row(pkg, ctx):
    rng = np.random.default_rng(7)
    # two (and more) calls in a row on the shared scratch: large, then smaller ones that leave stale tile sums behind,
    # tile-boundary sizes, one tile, empty
    for n in (1_000_003, 300_000, 4096 * 3, 4097, 4096, 31, 1, 0, 2_500_001):
        counts = rng.integers(0, 65, size=n, dtype=np.uint32)
        a, b = _both(pkg, ctx, counts)
        assert np.array_equal(a, b), f"n={n}"
        ref = np.zeros(n + 1, np.uint64)
        np.cumsum(counts, out=ref[1:])
        assert np.array_equal(b.astype(np.uint64), ref), f"n={n}"
    # the same input twice in a row gives the same output (nothing depends on what the scratch held)
    counts = rng.integers(0, 1000, size=777_777, dtype=np.uint32)
    _, b1 = _both(pkg, ctx, counts)
    _, b2 = _both(pkg, ctx, counts)
    assert np.array_equal(b1, b2)


def test_layout_noalloc_rejects_sizes_beyond_the_static_scratch(pkg, ctx):
    from sph_exa_b200.binding import lib
    rc = lib.sphb200_layout_noalloc(ctx.h, None, MAX + 1, None)
    assert rc != 0
    msg = lib.sphb200_last_error().decode()
    assert "exceeds the static scratch limit" in msg and str(MAX) in msg
    rc = lib.sphb200_layout_noalloc(ctx.h, None, -1, None)
    assert rc != 0
