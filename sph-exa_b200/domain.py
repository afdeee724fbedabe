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
                ("haloSend", C.c_uint32 * 32), ("haloRecv", C.c_uint32 * 32), ("assignment", C.c_uint64 * 33),
                ("haloExchangeMs", C.c_float * 6)]


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
lib.sphb200_domain_update_focus.argtypes = [vp, C.c_int, C.c_float, vp]
lib.sphb200_domain_focus_view.argtypes = [vp, vp, vp]


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

    def focusTree(self):
        """(leaves, leaf counts, peer ranks) of this rank's MAC-coarsened focus tree (after DistHydroVe.updateFocus)"""
        from .focus import FocusView
        v = FocusView()
        peers = (C.c_int * self._comm.num_ranks)()
        check(lib.sphb200_domain_focus_view(self.h, C.byref(v), peers))
        leaves = self._d2h(v.leaves, v.numLeaves + 1, np.uint64)
        counts = self._d2h(v.leafCounts, v.numLeaves, np.uint32)
        return leaves, counts, [q for q in range(self._comm.num_ranks) if peers[q]]


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

    def updateFocus(self, theta=1.0):
        """one round of the reference's focus-tree update on every rank (sphb200_domain_update_focus); the default
        theta is the reference's without self-gravity (sphexa.cpp:127)"""
        conv = C.c_int(0)
        check(lib.sphb200_domain_update_focus(self._arr, self.n, C.c_float(theta), C.byref(conv)))
        return bool(conv.value)

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


lib.sphb200_domain_step_host.argtypes = [vp, C.c_int, vp, vp, vp, vp, C.c_size_t]


def step_host(group, n_in, fin, fout, capacity):
    """sphb200_domain_step_host: fin / fout are binding.Fields of pinned host pointers (one per local rank, or a single
    one for a group with one local rank); returns the number of particles every local rank owns after the step"""
    single = not isinstance(n_in, (list, tuple))
    if single:
        n_in, fin, fout = [n_in], [fin], [fout]
    k = group.n
    nin = (C.c_size_t * k)(*n_in)
    nout = (C.c_size_t * k)()
    fi = (B.Fields * k)(*fin)
    fo = (B.Fields * k)(*fout)
    check(lib.sphb200_domain_step_host(group._arr, k, nin, fi, fo, nout, C.c_size_t(capacity)))
    out = [int(v) for v in nout]
    return out[0] if single else out


def bind_to_gpu_numa_node(local_rank):
    """pin this process to the CPUs closest to its GPU (nvmlDeviceSetCpuAffinity) so that the pinned host buffers of the
    host-resident step are first-touched on the GPU's own NUMA node -- what `numactl` / the batch system does for a
    production rank.  Returns the number of CPUs in the new affinity mask (None when NVML is unavailable)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        index = int(visible.split(",")[local_rank]) if visible and visible.split(",")[local_rank].isdigit() else local_rank
        handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        return len(os.sched_getaffinity(0))
    except Exception:
        return None


def nccl_comm(rank, world, local_rank):
    """NCCL transport for a torchrun launch: the unique id travels through torch.distributed (the reference's front-end
    would MPI_Bcast it)"""
    import torch
    import torch.distributed as dist

    def bcast(raw):
        t = torch.tensor(list(raw), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        return bytes(t.cpu().tolist())

    return Comm.nccl(rank, world, local_rank, bcast)


# ---------------------------------------------------------------------------------------------- bench entry (N > 1)
def bench_main(args, pkg, METRIC, UNIT, ALGO_BYTES, load_peaks, ClockSampler):
    """python bench.py --gpus N under torchrun: one rank per GPU, the C++ multi-rank domain over NCCL"""
    import json
    import os
    import time

    import torch
    import torch.distributed as dist
    from .dist import partition_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})")
    torch.cuda.set_device(local_rank)
    numa_cpus = bind_to_gpu_numa_node(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dtype = B.F64 if args.dtype == "f64" else B.MIXED

    # weak scaling: args.side^3 particles per GPU.  The global lattice side is kept even: an odd side puts a plane of
    # particles exactly on the mid-planes of the box, which are octree (and rank) boundaries, and round-off flips their
    # keys across it every step (a few thousand migrating particles per step that no real initial condition has)
    side = int(round(args.side * world ** (1.0 / 3.0)))
    side += side % 2
    n_global = side ** 3
    first, last = partition_range(n_global, rank, world)
    case = getattr(args, "case", "sedov")
    if case == "turbulence":
        f, box, s = pkg.ic.turbulence_range(side, first, last, args.dtype)
    elif case == "noh":
        xyz = pkg.ic.noh_range_coords(side, first, last)
        cnt = torch.tensor([xyz[0].size], dtype=torch.int64, device="cuda")
        dist.all_reduce(cnt)
        n_global = int(cnt.item())
        f, box, s = pkg.ic.noh_range_fields(*xyz, n_global, args.dtype)
    elif case == "sedov":
        f, box, s = pkg.ic.sedov_range(side, first, last, args.dtype)
    else:
        raise SystemExit("--case on more than one GPU: sedov, noh or turbulence")
    s = dict(s)
    s["numParticlesGlobal"] = n_global
    n_mine = f["x"].size
    ctx = pkg.Context(local_rank)
    comm = nccl_comm(rank, world, local_rank)
    # room for the balanced share plus halos; the slabs the ranks start from can be larger than that (Noh: cut sphere)
    capacity = int(max(1.3 * n_global / world, 1.05 * n_mine)) + 65536
    sim = DomainRank(ctx, box, s, capacity, comm, rank, dtype=dtype)
    turb = None
    if case == "turbulence":
        from .turbulence import Turbulence
        turb = Turbulence()
        sim.attachTurbulence(turb)
    sim.upload(f)
    del f
    group = DistHydroVe([sim])

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        group.step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches
    halo0 = sim.info().haloBytes
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        group.step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    launches = ctx.launches - launches0
    info = sim.info()
    halo_bytes = (info.haloBytes - halo0) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    own = torch.tensor([info.endIndex - info.startIndex, info.nWithHalos, info.numFocusLeaves, info.numSearchLeaves,
                        info.numSurfaceBoxes], dtype=torch.int64, device="cuda")
    gathered = [torch.zeros_like(own) for _ in range(world)]
    dist.all_gather(gathered, own)

    # per-phase device times (cudaEvents recorded by the library; a halo exchange is counted with the operator after it)
    phase_ms = {}
    reps = 3
    for _ in range(reps):
        group.step()
        for k, v in sim.phaseTimes().items():
            phase_ms[k] = phase_ms.get(k, 0.0) + v / reps

    e2e = None
    if not getattr(args, "no_e2e", False):
        real = np.float64 if args.dtype == "f64" else np.float32
        e2e_steps = max(2, min(args.steps, 5))
        cap = capacity
        state = sim.download()
        n_h = state["x"].size
        host_in, host_out = {}, {}
        for k in B.CONSERVED:
            want = torch.float64 if (k in B.ALWAYS_F64 or real == np.float64) else torch.float32
            host_in[k] = torch.empty(cap, dtype=want).pin_memory()
            host_out[k] = torch.empty(cap, dtype=want).pin_memory()
            host_in[k].numpy()[:n_h] = state[k]
        del state

        def structs():
            fi, fo = B.Fields(), B.Fields()
            for k in B.FIELD_NAMES:
                setattr(fi, k, host_in[k].data_ptr() if k in host_in else None)
                setattr(fo, k, host_out[k].data_ptr() if k in host_out else None)
            return fi, fo

        esz = sum(host_in[k].element_size() for k in B.CONSERVED)
        h2d = d2h = 0
        e2e_warm = 3  # untimed: first touch of the pinned buffers, the timed choice of the kernel-table variants
        for it in range(e2e_warm + e2e_steps):
            if it == e2e_warm:
                barrier()
                h2d = d2h = 0
                t0 = time.perf_counter()
            fi, fo = structs()
            h2d += n_h * esz
            n_h = step_host(group, n_h, fi, fo, cap)
            d2h += n_h * esz
            host_in, host_out = host_out, host_in
        barrier()
        dt_e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(dt_e, op=dist.ReduceOp.MAX)
        dt_e = float(dt_e.item())
        e2e = {"value": n_global * e2e_steps / dt_e, "unit": UNIT, "h2d_bytes_per_step": h2d // e2e_steps,
               "d2h_bytes_per_step": d2h // e2e_steps, "ms_per_step": dt_e / e2e_steps * 1e3, "steps": e2e_steps,
               "note": "bytes are rank 0's; every rank moves its own share through sphb200_domain_step_host",
               "cpus_in_numa_affinity_rank0": numa_cpus}

    if rank == 0:
        peak, peak_src = load_peaks()
        value = n_global * args.steps / (total_ms * 1e-3)
        kernels = {k: v for k, v in phase_ms.items() if k not in ("domain::sync",)}
        dom = max(kernels, key=kernels.get)
        col = 0 if args.dtype == "f64" else 1
        n_own = int(gathered[0][0])
        algo = ALGO_BYTES[dom][col] * n_own
        achieved = algo / (phase_ms[dom] * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": {"turbulence": "Stirred subsonic turbulence (turbulence-ve)",
                                    "noh": "Noh implosion (lattice cut to the sphere, open box)"}.get(
                                        case, "Sedov blast wave") +
                                   f" {side}^3 ({n_global} particles) SFC-partitioned over {world} B200 "
                                   f"({args.side}^3 per GPU), multi-rank domain in C++ (sphb200_domain_*), NCCL halo "
                                   f"exchange and reductions issued from C++, full step",
                       "l2": "inputs larger than L2", "parallelism": f"sfc{world}"},
            "find_neighbors_ms": phase_ms.get("FindNeighbors"), "phase_ms": phase_ms,
            "particles_per_rank": [[int(g[0]), int(g[1])] for g in gathered],
            "focus_leaves_per_rank": [int(g[2]) for g in gathered],
            "search_tree_leaves_per_rank": [int(g[3]) for g in gathered],
            "halo_boxes_published_per_rank": [int(g[4]) for g in gathered],
            "halo_bytes_per_step_rank0": halo_bytes,
            "halo_exchange_ms_rank0": {k: round(float(v), 4) for k, v in zip(
                ("sync_xyzhm", "before_gradh", "before_iad", "before_av", "before_momentum", "sum"),
                sim.info().haloExchangeMs)}, "clocks": clocks, "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "kernel_ms": phase_ms[dom]},
            "e2e": e2e, "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
    barrier()
    group.close()
    comm.close()
    if turb is not None:
        turb.close()
    dist.destroy_process_group()
    return 0
