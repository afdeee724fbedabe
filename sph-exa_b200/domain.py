"""Python mirror of the multi-rank Domain / HydroVeProp of the C ABI (sphb200_domain_*, csrc/domain.cu).

Everything that runs per step is C++/CUDA inside the library: the sync (global tree with all-reduced counts,
assignment with limitBoundaryShifts, migration, focus tree, halo discovery, layout, halo exchanges), the forces with
the four exchanges in between, and the MIN all-reduce of the time step.  This file only creates the handles and
marshals host buffers, like sim.py does for one rank.

Transports (csrc/comm.cu): `Comm.loopback(P)` runs P virtual ranks in this process on one GPU (tests on the
single-GPU box); `Comm.nccl(rank, P, device, bcast)` is the one-process-per-GPU transport, NCCL called from C++.
"""
import ctypes as C

import numpy as np

from . import binding as B
from .binding import lib, check, vp
from .sim import HydroVeProp, _host_fields

FIELD_ID = {n: i for i, n in enumerate(B.FIELD_NAMES)}


class DomainInfo(C.Structure):
    _fields_ = [("rank", C.c_int), ("numRanks", C.c_int), ("startIndex", C.c_uint32), ("endIndex", C.c_uint32),
                ("nWithHalos", C.c_uint32), ("numGlobalLeaves", C.c_int), ("numFocusLeaves", C.c_int),
                ("numSearchLeaves", C.c_int), ("numSurfaceBoxes", C.c_uint32), ("numSendRanges", C.c_uint32),
                ("migratedIn", C.c_uint32), ("migratedOut", C.c_uint32), ("haloBytes", C.c_uint64),
                ("haloSend", C.c_uint32 * 32), ("haloRecv", C.c_uint32 * 32), ("assignment", C.c_uint64 * 33)]


lib.sphb200_comm_create_loopback.argtypes = [C.c_int, vp]
lib.sphb200_comm_unique_id.argtypes = [vp]
lib.sphb200_comm_create_nccl.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp]
lib.sphb200_comm_destroy.argtypes = [vp]
lib.sphb200_comm_destroy.restype = None
lib.sphb200_domain_create.argtypes = [vp, vp, vp, vp]
for _n in ("sphb200_domain_sync", "sphb200_domain_compute_forces", "sphb200_domain_integrate", "sphb200_domain_step"):
    getattr(lib, _n).argtypes = [vp, C.c_int]
lib.sphb200_domain_exchange_halos.argtypes = [vp, C.c_int, vp, C.c_int]
lib.sphb200_domain_conserved.argtypes = [vp, C.c_int, vp]
lib.sphb200_domain_set_halo_factor.argtypes = [vp, C.c_float]
lib.sphb200_domain_set_halo_factor.restype = None
lib.sphb200_domain_info.argtypes = [vp, vp]
lib.sphb200_sfc_assignment.argtypes = [C.c_int, vp, vp, C.c_int, vp, vp, vp]
lib.sphb200_initial_global_tree.argtypes = [C.c_int, vp, C.c_int]
lib.sphb200_find_halos.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp, vp]
lib.sphb200_halo_radii.argtypes = [vp, vp, C.c_int, vp, C.c_int, C.c_int, C.c_float, vp]


class Comm:
    """sphb200_comm: the transport of a multi-rank run"""

    def __init__(self, handle, num_ranks, rank=None):
        self.h, self.num_ranks, self.rank = handle, num_ranks, rank

    @classmethod
    def loopback(cls, num_ranks):
        h = vp()
        check(lib.sphb200_comm_create_loopback(num_ranks, C.byref(h)))
        return cls(h, num_ranks)

    @classmethod
    def nccl(cls, rank, num_ranks, device, bcast):
        """bcast(bytearray of 128) -> the same 128 bytes on every rank, taken from rank 0 (MPI_Bcast in the reference's
        front-end; torch.distributed.broadcast in bench.py)"""
        buf = (C.c_char * 128)()
        if rank == 0:
            check(lib.sphb200_comm_unique_id(buf))
        raw = bcast(bytearray(buf.raw))
        buf2 = (C.c_char * 128).from_buffer_copy(bytes(raw))
        h = vp()
        check(lib.sphb200_comm_create_nccl(device, rank, num_ranks, buf2, C.byref(h)))
        return cls(h, num_ranks, rank)

    def close(self):
        if self.h:
            lib.sphb200_comm_destroy(self.h)
            self.h = vp()


class DomainRank(HydroVeProp):
    """one rank of a multi-rank run: a sphb200_sim created by sphb200_domain_create (the communicator attached)"""

    def __init__(self, ctx, box, settings, capacity, comm, rank, dtype=B.F64, bucket_size=None, bucket_size_focus=64,
                 av_clean=False, tables=None):
        self._comm = comm

        def create(ctx_h, cfg_ref, out_ref):
            return lib.sphb200_domain_create(ctx_h, cfg_ref, comm.h, out_ref)

        super().__init__(ctx, box, settings, capacity, dtype=dtype, rank=rank, num_ranks=comm.num_ranks,
                         bucket_size=bucket_size, bucket_size_focus=bucket_size_focus, av_clean=av_clean,
                         tables=tables, creator=create)

    def info(self):
        out = DomainInfo()
        check(lib.sphb200_domain_info(self.h, C.byref(out)))
        return out

    def setHaloFactor(self, f):
        lib.sphb200_domain_set_halo_factor(self.h, C.c_float(f))


class DistHydroVe:
    """HydroVeProp over the ranks this process drives (one with NCCL, all P with the loopback transport)"""

    def __init__(self, ranks):
        self.ranks = list(ranks)
        self._arr = (vp * len(self.ranks))(*[r.h for r in self.ranks])
        self.n = len(self.ranks)

    def sync(self):
        check(lib.sphb200_domain_sync(self._arr, self.n))

    def computeForces(self):
        check(lib.sphb200_domain_compute_forces(self._arr, self.n))

    def integrate(self):
        check(lib.sphb200_domain_integrate(self._arr, self.n))

    def step(self):
        check(lib.sphb200_domain_step(self._arr, self.n))

    def exchangeHalos(self, names):
        ids = (C.c_int * len(names))(*[FIELD_ID[n] for n in names])
        check(lib.sphb200_domain_exchange_halos(self._arr, self.n, ids, len(names)))

    def conserved(self):
        from . import ops
        out = (C.c_double * 10)()
        check(lib.sphb200_domain_conserved(self._arr, self.n, out))
        return ops.conserved_dict(list(out))

    def download(self, names=None):
        """owned particles of all local ranks, concatenated in rank (= SFC) order"""
        parts = [r.download(names=names, with_keys=True) for r in self.ranks]
        return {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}

    def close(self):
        for r in self.ranks:
            r.close()
            if getattr(r, "own_ctx", False):
                r.ctx.close()


def sfc_assignment(num_ranks, counts, leaves, old_bounds=None):
    """makeSfcAssignment + limitBoundaryShifts on the host (no device): (bounds[P+1], rank counts[P])"""
    counts = np.ascontiguousarray(counts, np.uint32)
    leaves = np.ascontiguousarray(leaves, np.uint64)
    bounds = np.zeros(num_ranks + 1, np.uint64)
    rc = np.zeros(num_ranks, np.uint32)
    old = np.ascontiguousarray(old_bounds, np.uint64) if old_bounds is not None else None
    check(lib.sphb200_sfc_assignment(num_ranks, counts.ctypes.data, leaves.ctypes.data, counts.size,
                                     old.ctypes.data if old is not None else None, bounds.ctypes.data, rc.ctypes.data))
    return bounds, rc


def initial_global_tree(num_ranks):
    buf = np.zeros(1 << 16, np.uint64)
    n = lib.sphb200_initial_global_tree(num_ranks, buf.ctypes.data, buf.size)
    if n <= 0:
        raise B.SphError("initial_global_tree failed")
    return buf[:n + 1].copy()
