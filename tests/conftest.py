import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    """a plain `pytest tests` on a machine without CUDA skips the gpu-marked tests instead of failing them"""
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    if have_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def pkg():
    import sphexa_b200_loader
    return sphexa_b200_loader.load()


@pytest.fixture(scope="session")
def ref():
    from oracle import ref as r
    if not r.available():
        pytest.skip("oracle/_ref/libsphexa_ref.so not built (run `make -C oracle ref` where /root/reference exists)")
    return r


@pytest.fixture(scope="session")
def ctx(pkg):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    c = pkg.Context(0)
    yield c
    c.close()


def to_dev(a):
    import torch
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint64:
        a = a.view(np.int64)
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    return torch.from_numpy(a).cuda()


def to_host(t, dtype=None):
    a = t.detach().cpu().numpy()
    if dtype is not None:
        a = a.view(dtype)
    return a


def slab_from_rowmajor(ctx, nb, nc, ngmax):
    """row-major neighbor lists [n, ngmax] (reference CPU layout) -> device list buffer of the CUDA kernels"""
    from sph_exa_b200 import ops
    rm = to_dev(np.ascontiguousarray(nb[:, :ngmax], np.uint32))
    return ops.neighborsFromRowMajor(ctx, rm, to_dev(np.ascontiguousarray(nc, np.uint32)), ngmax)


def sorted_sets_equal(nb_a, nc_a, nb_b, nc_b, ngmax):
    """neighbor lists equal as sets per particle (the reference's own tests sort before comparing,
    domain/test/unit/neighbors/findneighbors.cpp:61,92-95)"""
    if not np.array_equal(nc_a, nc_b):
        return False
    cnt = np.minimum(nc_a.astype(np.int64) - 1, ngmax)
    k = np.arange(ngmax)[None, :]
    mask = k < cnt[:, None]
    big = np.iinfo(np.uint32).max
    a = np.where(mask, nb_a, big)
    b = np.where(mask, nb_b, big)
    a.sort(axis=1)
    b.sort(axis=1)
    return np.array_equal(a, b)


def rel_err(a, b):
    """max |a-b| scaled by the field's magnitude (forces on a lattice cancel to ~0, so a per-element
    relative error is meaningless there)"""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    scale = max(np.abs(b).max(), 1e-300)
    return np.abs(a - b).max() / scale
