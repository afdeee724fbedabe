"""Block-time-step group kernels against the reference's OWN GPU kernels on the same device.

cstone::computeGroupSplits (domain/include/cstone/traversal/groups.cuh:150-310, what sph::computeSpatialGroups calls,
sph/include/sph/groups.cu:30-47) and sph::groupDivvTimestepGpu / groupAccTimestepGpu (sph/include/sph/ts_groups.cu)
exist in the reference only as CUDA code, so tests/test_gpu_groups.py could only compare with the numpy restatement
of oracle/bdt.py.  oracle/build_ref_gpu.sh compiles those reference sources, where they lie, for sm_100a into
baseline/_ref/libsphexa_ref_gpu.so (test infrastructure, git-ignored, travels to the GPU box); here the library's
kernels and the reference's run on the same device arrays and must agree bit for bit.  This pins the last
"parity unpinned" pieces of oracle/bdt.py: the restatement and the reference agree wherever both are compared with
the library.
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import to_dev, to_host

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "baseline", "_ref", "libsphexa_ref_gpu.so")


@pytest.fixture(scope="module")
def refgpu():
    if not os.path.exists(LIB):
        pytest.skip("baseline/_ref/libsphexa_ref_gpu.so not built (oracle/build_ref_gpu.sh needs /root/reference)")
    lib = C.CDLL(LIB)
    for name in ("refgpu_group_splits_f64", "refgpu_group_splits_f32"):
        getattr(lib, name).restype = C.c_int
    return lib


def vp(t):
    return C.c_void_p(t.data_ptr())


def reference_splits(refgpu, first, last, x, y, z, h, leaves, layout, box, tol=2.0):
    import torch
    lim = (C.c_double * 6)(*box[0])
    bc = (C.c_int * 3)(*box[1])
    cap = (last - first) + 2
    out = torch.empty(cap, dtype=torch.int32, device="cuda")
    fn = refgpu.refgpu_group_splits_f64 if h.dtype == torch.float64 else refgpu.refgpu_group_splits_f32
    n = fn(C.c_uint32(first), C.c_uint32(last), vp(x), vp(y), vp(z), vp(h), vp(leaves), C.c_int(leaves.numel() - 1),
           vp(layout), lim, bc, C.c_float(tol), vp(out), C.c_uint32(cap))
    assert n > 0
    return to_host(out[:n], np.uint32)


@pytest.mark.parametrize("gaussian", [False, True])
@pytest.mark.parametrize("rng_", [(0, None), (1000, 54321)])
def test_group_splits_equal_the_reference_kernel(pkg, ctx, ref, refgpu, gaussian, rng_):
    import torch
    from sph_exa_b200 import ops
    from test_gpu_domain import BOXES
    from test_gpu_groups import sorted_case
    n = 100_003
    for box in (BOXES[0], BOXES[1], BOXES[2]):
        for bucket in (64, 16):
            x, y, z, leaves, layout, tree = sorted_case(pkg, ctx, ref, n, box, 3, gaussian, bucket)
            first, last = rng_[0], (n if rng_[1] is None else rng_[1])
            dx, dy, dz = to_dev(x), to_dev(y), to_dev(z)
            g = ops.computeGroups(ctx, first, last, dx, dy, dz, tree, pkg.make_box(*box))
            torch.cuda.synchronize()
            for real in (np.float64, np.float32):
                h = to_dev(np.full(n, 0.01, real))
                want = reference_splits(refgpu, first, last, dx, dy, dz, h, to_dev(leaves), to_dev(layout), box)
                assert np.array_equal(to_host(g.start, np.uint32), want[:-1])
                assert np.array_equal(to_host(g.end, np.uint32), want[1:])


def test_group_timesteps_equal_the_reference_kernels(pkg, ctx, ref, refgpu):
    import torch
    from sph_exa_b200 import ops
    from test_gpu_domain import BOXES
    from test_gpu_groups import sorted_case
    n = 50_021
    box = BOXES[1]
    x, y, z, leaves, layout, tree = sorted_case(pkg, ctx, ref, n, box, 7, True)
    first, last = 64, n - 5
    g = ops.computeGroups(ctx, first, last, to_dev(x), to_dev(y), to_dev(z), tree, pkg.make_box(*box))
    rng = np.random.Generator(np.random.MT19937(11))
    fmax = np.finfo(np.float32).max
    for real, tag in ((np.float64, "f64"), (np.float32, "f32")):
        divv = to_dev((rng.normal(size=n) * 3).astype(real))
        ax, ay, az = (to_dev((rng.normal(size=n) * 10 ** rng.uniform(-3, 3)).astype(real)) for _ in range(3))
        ours = to_dev(np.full(g.numGroups, fmax, np.float32))
        theirs = to_dev(np.full(g.numGroups, fmax, np.float32))
        ops.groupDivvTimestep(ctx, 0.06, g.view(), divv, ours)
        torch.cuda.synchronize()
        getattr(refgpu, "refgpu_group_divv_" + tag)(C.c_float(0.06), C.c_uint32(first), C.c_uint32(last),
                                                    C.c_uint32(g.numGroups), vp(g.start), vp(g.end), vp(divv),
                                                    vp(theirs))
        assert np.array_equal(to_host(ours), to_host(theirs))
        assert to_host(ours).max() < fmax
        ops.groupAccTimestep(ctx, 0.3, g.view(), ax, ay, az, ours)
        torch.cuda.synchronize()
        getattr(refgpu, "refgpu_group_acc_" + tag)(C.c_float(0.3), C.c_uint32(first), C.c_uint32(last),
                                                   C.c_uint32(g.numGroups), vp(g.start), vp(g.end), vp(ax), vp(ay),
                                                   vp(az), vp(theirs))
        a, b = to_host(ours), to_host(theirs)
        # |a|^2 may contract to FMAs differently in the two builds: one float ulp
        assert np.allclose(a, b, rtol=2.5e-7, atol=0), np.abs(a / b - 1).max()
