"""SFC-partitioned multi-GPU SPH-VE: one process per GPU, torch.distributed (NCCL) for the exchanges.

Follows the reference's multi-rank flow (domain/include/cstone/domain/domain.hpp:196-244,
assignment.hpp:84-168, halos/halos.hpp, main/src/propagator/ve_hydro.hpp:132-217):

  sync:   global box (Allreduce MIN/MAX) -> keys -> sort -> replicated global tree (counts Allreduce SUM)
          -> SFC assignment (uniformBins) -> particle migration (all-to-all) -> local bucket-64 tree
          -> halo discovery -> halo exchange of x,y,z,h,m -> search tree over owned + halo particles
  forces: neighbors+XMass | halo(xm) | Gradh+EOS | halo(v,prho,c,kx) | IAD | halo(c_ij,divv) | AV | halo(alpha) | momentum
  integrate: local dt -> Allreduce MIN -> positions, h

Differences to the reference, by design: halo discovery is owner-side on query boxes (a
superset of the reference's halo leaves), halos of rank q are stored in rank order so that [halos<me | owned | halos>me]
is SFC-sorted without another sort, and the search tree outside the owned key range is a plain bucket-64 tree instead of
the MAC-coarsened focus tree.  Neighbor sets and all fields of owned particles equal the single-rank run.

Host-only helpers (uniform_bins, key_splits, halo_layout) are pure numpy and are tested on CPU with gloo.
"""
import ctypes as C
import json
import math
import os
import sys
import time

import numpy as np

MAX_QUERY_BOXES = 512


# ---------------------------------------------------------------------------------------------- host logic (CPU-testable)
def uniform_bins(counts, num_bins):
    """uniformBins (domain/include/cstone/domain/domaindecomp.hpp:49-72): leaf index boundaries of num_bins bins of
    (nearly) equal particle count; returns (bins[num_bins+1], bin_counts[num_bins])"""
    scan = np.zeros(len(counts) + 1, np.uint64)
    np.cumsum(np.asarray(counts, np.uint64), out=scan[1:])
    bin_count = float(scan[-1]) / num_bins
    bins = np.zeros(num_bins + 1, np.int64)
    bins[-1] = len(counts)
    for i in range(1, num_bins):
        target = np.uint64(int(i * bin_count))  # computed in double, then truncated (domaindecomp.hpp:63)
        bins[i] = np.searchsorted(scan, target, side="left")
    bin_counts = (scan[bins[1:]] - scan[bins[:-1]]).astype(np.int64)
    return bins, bin_counts


def make_sfc_assignment(num_ranks, counts, leaves):
    """makeSfcAssignment (domaindecomp.hpp:115-125): rank r owns keys in [b[r], b[r+1])"""
    bins, bin_counts = uniform_bins(counts, num_ranks)
    return np.asarray(leaves, np.uint64)[bins], bin_counts


def key_splits(sorted_keys, boundaries):
    """createSendRanges (domaindecomp.hpp:217-231): index ranges of the sorted local keys per destination rank"""
    return np.searchsorted(np.asarray(sorted_keys, np.uint64), np.asarray(boundaries, np.uint64), side="left")


def migrate_fields(t, f, splits, send, recv, me, a2a):
    """exchangeParticles (domaindecomp_mpi.hpp:85-173) on key-sorted arrays: rank q gets the range
    [splits[q], splits[q+1]) of every array in `f`.  Only the leaving particles travel — the ranges below and above
    the one this rank keeps; fields of one element type go through ONE all-to-all as peer-major [fields x count_q]
    blocks (the self block is empty).  Returns the new arrays: kept range first, arrivals behind it in peer order (the
    caller sorts by key afterwards).  a2a(recvbuf, sendbuf, output_split_sizes, input_split_sizes) is the collective."""
    P = len(send)
    lo_keep, hi_keep = splits[me], splits[me + 1]
    send_x = [0 if q == me else int(send[q]) for q in range(P)]
    recv_x = [0 if q == me else int(recv[q]) for q in range(P)]
    groups = {}
    for k, a in f.items():
        groups.setdefault(a.dtype, []).append(k)
    out = {}
    for dt_, names in groups.items():
        nf = len(names)
        stacked = t.stack([t.cat([f[k][:lo_keep], f[k][hi_keep:]]) for k in names])  # [nf, leaving]
        off, blocks = 0, []
        for q in range(P):
            blocks.append(stacked[:, off:off + send_x[q]].reshape(-1))
            off += send_x[q]
        sendbuf = t.cat(blocks)
        recvbuf = t.empty(nf * sum(recv_x), dtype=dt_, device=stacked.device)
        a2a(recvbuf, sendbuf, [nf * c for c in recv_x], [nf * c for c in send_x])
        off, parts = 0, []
        for q in range(P):
            parts.append(recvbuf[off:off + nf * recv_x[q]].reshape(nf, recv_x[q]))
            off += nf * recv_x[q]
        got = t.cat(parts, dim=1)  # [nf, arriving]
        for i, k in enumerate(names):
            out[k] = t.cat([f[k][lo_keep:hi_keep], got[i]])
    return out


def halo_layout(rank, n_own, recv_counts):
    """[halos from ranks < me | owned | halos from ranks > me]: (start, end, n_total, offset of every peer's halos)"""
    recv = list(int(c) for c in recv_counts)
    before = sum(recv[:rank])
    start, end = before, before + n_own
    off, cur = [], 0
    for q, c in enumerate(recv):
        if q == rank:
            off.append(-1)
            cur = end
            continue
        off.append(cur)
        cur += c
    return start, end, cur, off


def partition_range(n, rank, num_ranks):
    """partitionRange (main/src/init/utils.hpp): contiguous slab of the global lattice per rank"""
    per = n // num_ranks
    rem = n % num_ranks
    first = rank * per + min(rank, rem)
    return first, first + per + (1 if rank < rem else 0)


# ---------------------------------------------------------------------------------------------- device side
class DistHydroVeProp:
    """HydroVeProp for one rank of a P-rank run; fields are torch tensors, kernels come from the C ABI"""

    CONSERVED = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "x_m1", "y_m1", "z_m1", "temp", "du_m1", "alpha"]
    DEPENDENT = ["xm", "kx", "gradh", "prho", "c", "c11", "c12", "c13", "c22", "c23", "c33", "divv", "curlv", "ax",
                 "ay", "az"]

    def __init__(self, pkg, ctx, box, settings, dtype, rank, world, n_global, bucket_focus=64, group=None):
        import torch
        import torch.distributed as dist
        from . import ops, binding as B
        self.torch, self.dist, self.ops, self.B, self.pkg = torch, dist, ops, B, pkg
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        self.dtype = dtype
        self.real_t = torch.float64 if dtype == B.F64 else torch.float32
        self.s = dict(settings)
        self.lim, self.bc = list(box[0]), list(box[1])
        self.box = pkg.make_box(self.lim, self.bc)
        self.n_global = n_global
        self.bucket = max(bucket_focus, n_global // (100 * world))  # sphexa.cpp:133-135
        self.bucket_focus = bucket_focus
        self.ng0, self.ngmax = int(settings.get("ng0", 100)), int(settings.get("ngmax", 150))
        wh, whd, K = pkg.ic.kernel_tables(np.float64 if dtype == B.F64 else np.float32)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.dev = dev
        self.wh = torch.from_numpy(wh).to(dev)
        self.whd = torch.from_numpy(whd).to(dev)
        self.K = K
        self.minDt, self.minDt_m1, self.ttot = settings["minDt"], settings["minDt_m1"], 0.0
        self.f = {}
        self.gtree = None      # replicated global tree (leaves, global counts)
        self.ftree = None      # bucket-64 tree over owned keys
        self.stree = None      # bucket-64 tree over owned + halo keys (search tree)
        self.assignment = None
        self.times = {}
        self.halo_bytes = 0
        self.halo_factor = 1.0
        self.turb = None
        self._sync_profile = bool(int(os.environ.get("SPHB200_SYNC_PROFILE", "0")))
        self.sync_times = {}

    # ---- helpers
    def upload(self, arrays):
        t = self.torch
        for k in self.CONSERVED:
            want = np.float64 if k in self.B.ALWAYS_F64 else (np.float64 if self.dtype == self.B.F64 else np.float32)
            self.f[k] = t.from_numpy(np.ascontiguousarray(arrays[k], dtype=want)).to(self.dev)

    def _hc(self):
        s = self.s
        return self.ops.hydro_const(self.K, self.wh, self.whd, ng0=self.ng0, ngmax=self.ngmax,
                                    Kcour=s.get("Kcour", 0.2), Krho=s.get("Krho", 0.06), gamma=s["gamma"],
                                    muiConst=s["mui"], minDt=self.minDt, minDt_m1=self.minDt_m1)

    def _allreduce(self, tensor, op):
        if self.world > 1:
            self.dist.all_reduce(tensor, op=op, group=self.group)
        return tensor

    def _update_tree(self, tree, keys, bucket, reduce_counts):
        """computeOctree on the first call, one updateOctree step afterwards; global variant all-reduces the counts"""
        t, ops = self.torch, self.ops
        if tree is None:
            leaves = t.tensor([0, -(1 << 63)], dtype=t.int64, device=self.dev)  # {0, 2^63}
            n = t.tensor([keys.numel()], dtype=t.int32, device=self.dev)
            if reduce_counts:
                self._allreduce(n, self.dist.ReduceOp.SUM)
            counts, conv = n, False
            while not conv:
                leaves, counts, conv = ops.updateOctreeGpu(self.ctx, keys, bucket, leaves, counts)
                if reduce_counts:
                    self._allreduce(counts, self.dist.ReduceOp.SUM)
            return leaves, counts
        leaves, counts, _ = ops.updateOctreeGpu(self.ctx, keys, bucket, tree[0], tree[1])
        if reduce_counts:
            self._allreduce(counts, self.dist.ReduceOp.SUM)
        return leaves, counts

    # ---- Domain::sync
    def _st(self, name):
        """optional fine-grained timers of sync() (SPHB200_SYNC_PROFILE=1), synchronising"""
        if not self._sync_profile:
            return
        self.torch.cuda.synchronize()
        now = time.perf_counter()
        if name is not None:
            self.sync_times[name] = self.sync_times.get(name, 0.0) + (now - self._st0)
        self._st0 = now

    def sync(self, lazy=False):
        """cstone::Domain::sync (domain/domain.hpp:196-244).  lazy: leave the fields that nothing needs before EOS
        un-reordered until _finish_reorder() (computeForces does that)"""
        self._sync_impl()
        if not lazy:
            self._finish_reorder()

    def _sync_impl(self):
        t, ops, dist, f = self.torch, self.ops, self.dist, self.f
        P, me = self.world, self.rank
        self._st(None)
        # 1. global box for open axes (makeGlobalBox, box_mpi.hpp:84-119)
        if any(b != 1 for b in self.bc):
            mm = ops.minMaxGpu(self.ctx, f["x"], f["y"], f["z"])
            v = t.tensor([mm[0], -mm[1], mm[2], -mm[3], mm[4], -mm[5]], dtype=t.float64, device=self.dev)
            self._allreduce(v, dist.ReduceOp.MIN)
            v = v.tolist()
            for d in range(3):
                if self.bc[d] != 1:
                    self.lim[2 * d], self.lim[2 * d + 1] = v[2 * d], -v[2 * d + 1]
            self.box = self.pkg.make_box(self.lim, self.bc)

        self._st("box")
        # 2. keys, sort, reorder
        keys = ops.computeSfcKeysGpu(self.ctx, f["x"], f["y"], f["z"], self.box)
        perm = ops.sortByKeyGpu(self.ctx, keys)
        # the reorder of the fields is deferred until the halo layout is known, so that it can write straight into the
        # [halos | owned | halos] arrays; only h is needed before that (halo discovery)
        pending = dict(f) if P > 1 else None
        self._late = None
        if P > 1:
            self._wait_upload("h")
            f["h"] = ops.gatherGpu(self.ctx, perm, [f["h"]])[0]
        else:
            out = ops.gatherGpu(self.ctx, perm, [f[k] for k in self.CONSERVED])
            for k, a in zip(self.CONSERVED, out):
                f[k] = a

        self._st("keys+sort+gather")
        # 3. replicated global tree + assignment
        self.gtree = self._update_tree(self.gtree, keys, self.bucket, True)
        self._st("global tree")
        if P > 1:
            g_leaves = self.gtree[0].cpu().numpy().view(np.uint64)
            g_counts = self.gtree[1].cpu().numpy().view(np.uint32)
            bounds, _ = make_sfc_assignment(P, g_counts, g_leaves)
            self.assignment = bounds
            # 4. migration (exchangeParticles, domaindecomp_mpi.hpp:85-173)
            bd = t.from_numpy(bounds.view(np.int64).copy()).to(self.dev)
            # keys are < 2^63 so signed compare is fine; the terminal boundary 2^63 wraps negative -> use n
            splits = t.searchsorted(keys, bd[:-1].contiguous()).tolist() + [keys.numel()]
            send = [splits[q + 1] - splits[q] for q in range(P)]
            # every rank sees the whole send matrix: the migration collectives below must be entered by all ranks or none
            send_t = t.tensor(send, dtype=t.int64, device=self.dev)
            mat = t.empty((P, P), dtype=t.int64, device=self.dev)
            dist.all_gather_into_tensor(mat, send_t, group=self.group)
            mat = mat.tolist()
            recv = [mat[q][me] for q in range(P)]
            moved = sum(mat[q][r] for q in range(P) for r in range(P) if q != r)
            self._st("assignment")
            if moved:
                self._wait_upload("m")
                self._wait_upload("late")
                out = ops.gatherGpu(self.ctx, perm, [pending[k] for k in self.CONSERVED])
                for k, a in zip(self.CONSERVED, out):
                    f[k] = a
                pending = None
                def a2a(recvbuf, sendbuf, out_splits, in_splits):
                    dist.all_to_all_single(recvbuf, sendbuf, output_split_sizes=out_splits,
                                           input_split_sizes=in_splits, group=self.group)

                f.update(migrate_fields(t, {k: f[k] for k in self.CONSERVED}, splits, send, recv, me, a2a))
                keys = ops.computeSfcKeysGpu(self.ctx, f["x"], f["y"], f["z"], self.box)
                perm = ops.sortByKeyGpu(self.ctx, keys)
                out = ops.gatherGpu(self.ctx, perm, [f[k] for k in self.CONSERVED])
                for k, a in zip(self.CONSERVED, out):
                    f[k] = a
        self._st("migration")
        self.own_keys = keys
        n_own = keys.numel()

        if P == 1:
            self.start, self.end, self.n_total = 0, n_own, n_own
            self.ftree = self._update_tree(self.ftree, keys, self.bucket_focus, False)
            self.tree = ops.Octree(self.ctx, self.ftree[0], self.box, self.ftree[1])
            self.all_keys = keys
            return

        # 5. local tree over owned keys, halo discovery
        self.ftree = self._update_tree(self.ftree, keys, self.bucket_focus, False)
        ftree = ops.Octree(self.ctx, self.ftree[0], self.box, self.ftree[1])
        self._st("focus tree")
        L, ni = ftree.numLeaves, ftree.numInternal
        hmax = t.empty(L, dtype=t.float64, device=self.dev)
        self.B.check(self.B.lib.sphb200_segment_max(self.ctx.h, self.B.ptr(f["h"]), self.dtype, self.B.ptr(ftree.layout),
                                                    L, self.B.ptr(hmax)))
        l2i = ftree.leafToInternal[ni:].long()
        leaf_c = ftree.centers[l2i].contiguous()
        leaf_s = ftree.sizes[l2i].contiguous()
        occupied = self.ftree[1] > 0
        # Domain::setHaloFactor (domain.hpp): block time steps widen the halo search so that the halos stay valid while
        # particles drift between two full syncs
        half = leaf_s + (2.0 * self.halo_factor * hmax * (1 + 1e-9)).unsqueeze(1)
        lo, hi = leaf_c - half, leaf_c + half
        # union of the halo boxes per coarse SFC block so that at most MAX_QUERY_BOXES boxes are published
        lk = self.ftree[0][:-1]
        level = getattr(self, "_qb_level", 7)  # start from last step's level: the tree changes slowly
        while True:
            gid = lk >> (3 * (21 - level))
            uniq, inv = t.unique_consecutive(gid, return_inverse=True)
            if uniq.numel() <= MAX_QUERY_BOXES or level == 0:
                break
            level -= 1
        self._qb_level = level
        G = uniq.numel()
        big = 1e300
        lo_m = t.where(occupied.unsqueeze(1), lo, t.full_like(lo, big))
        hi_m = t.where(occupied.unsqueeze(1), hi, t.full_like(hi, -big))
        glo = t.full((G, 3), big, dtype=t.float64, device=self.dev).scatter_reduce(0, inv.unsqueeze(1).expand(-1, 3),
                                                                                   lo_m, "amin")
        ghi = t.full((G, 3), -big, dtype=t.float64, device=self.dev).scatter_reduce(0, inv.unsqueeze(1).expand(-1, 3),
                                                                                    hi_m, "amax")
        valid = ghi[:, 0] >= glo[:, 0]
        glo, ghi = glo[valid], ghi[valid]
        G = glo.shape[0]
        qb = t.zeros((MAX_QUERY_BOXES, 6), dtype=t.float64, device=self.dev)
        qb[:G, :3] = 0.5 * (glo + ghi)
        qb[:G, 3:] = 0.5 * (ghi - glo) * (1 + 1e-12)
        self._st("query boxes")
        all_qb = t.empty((P, MAX_QUERY_BOXES, 6), dtype=t.float64, device=self.dev)
        all_cnt = t.empty(P, dtype=t.int32, device=self.dev)
        dist.all_gather_into_tensor(all_qb, qb, group=self.group)
        dist.all_gather_into_tensor(all_cnt, t.tensor([G], dtype=t.int32, device=self.dev), group=self.group)
        mask = t.empty(L, dtype=t.int32, device=self.dev)
        self.B.check(self.B.lib.sphb200_mark_halo_leaves(self.ctx.h, self.B.ptr(leaf_c), self.B.ptr(leaf_s), L,
                                                         self.B.ptr(all_qb), self.B.ptr(all_cnt), P, MAX_QUERY_BOXES, me,
                                                         C.byref(self.box), self.B.ptr(mask)))
        self._st("gather boxes + mark")
        layout = ftree.layout
        counts = self.ftree[1]
        # send lists of all peers at once: (peer, leaf) pairs in peer-major order, one range per marked leaf
        qs = t.arange(P, device=self.dev, dtype=t.int32)
        bits = (((mask.unsqueeze(0) >> qs.unsqueeze(1)) & 1) > 0) & occupied.unsqueeze(0)
        bits[me] = False
        pl = t.nonzero(bits)
        leaf = pl[:, 1]
        lens = counts[leaf].long()
        self.send_scan = (t.cumsum(lens, 0) - lens).int().contiguous()
        send_begins = layout[:-1][leaf]
        per_peer = t.zeros(P, dtype=t.int64, device=self.dev).index_add_(0, pl[:, 0], lens)
        recv_t = t.empty_like(per_peer)
        dist.all_to_all_single(recv_t, per_peer, group=self.group)
        both = t.stack([per_peer, recv_t]).tolist()
        self.halo_send, self.halo_recv = both[0], both[1]
        self.start, self.end, self.n_total, self.halo_off = halo_layout(me, n_own, self.halo_recv)
        # send ranges were found in the owned-only arrays: shift them to the [halos | owned | halos] layout
        self.send_begins = (send_begins + self.start).int().contiguous()
        # receive side: one contiguous range per peer
        rc = t.tensor(self.halo_recv, dtype=t.int64)
        self.recv_scan = (t.cumsum(rc, 0) - rc).int().to(self.dev)
        self.recv_begins = t.tensor([max(0, o) for o in self.halo_off], dtype=t.int32, device=self.dev)
        self._st("send lists")
        # 6. grow every conserved field to [halos | owned | halos]; exchange x,y,z,h,m (setupHalos, domain.hpp:512-529)
        if pending is not None:
            # only what the halo discovery, the search and XMass need is reordered now; the other fields follow right
            # before their first use (_finish_reorder), by which time a host-resident caller has them uploaded
            self._wait_upload("m")
            srcs = [pending[k] for k in self.EARLY]
            grown = [t.empty(self.n_total, dtype=a.dtype, device=self.dev) for a in srcs]
            ops.gatherGpu(self.ctx, perm, srcs, [g[self.start:self.end] for g in grown])
            for k, g in zip(self.EARLY, grown):
                f[k] = g
            self._late = (perm, pending, self.start, self.end, self.n_total)
        else:
            for k in self.CONSERVED:
                a = t.empty(self.n_total, dtype=f[k].dtype, device=self.dev)
                a[self.start:self.end] = f[k]
                f[k] = a
        self._st("grow arrays")
        self.exchange_halos(["x", "y", "z", "h", "m"])
        self._st("halo xyzhm")
        # 7. search tree over owned + halo keys (already SFC-ordered: peers' key ranges are ordered by rank)
        a, b, n = self.start, self.end, self.n_total
        self.all_keys = t.empty(n, dtype=keys.dtype, device=self.dev)
        self.all_keys[a:b] = keys
        if a > 0:
            self.all_keys[:a] = ops.computeSfcKeysGpu(self.ctx, f["x"][:a], f["y"][:a], f["z"][:a], self.box)
        if n > b:
            self.all_keys[b:] = ops.computeSfcKeysGpu(self.ctx, f["x"][b:], f["y"][b:], f["z"][b:], self.box)
        self.stree = self._update_tree(self.stree, self.all_keys, self.bucket_focus, False)
        self.tree = ops.Octree(self.ctx, self.stree[0], self.box, self.stree[1])
        self._st("search tree")

    EARLY = ["x", "y", "z", "h", "m"]

    def _wait_upload(self, name):
        """step_host: make the compute stream wait for the upload group `name` (no-op otherwise)"""
        ev = getattr(self, "_upload_events", {}).pop(name, None)
        if ev is not None:
            self.torch.cuda.current_stream().wait_event(ev)

    def _finish_reorder(self):
        """reorder of the conserved fields that sync() deferred, into the [halos | owned | halos] arrays"""
        if getattr(self, "_late", None) is None:
            return
        t, ops, f = self.torch, self.ops, self.f
        perm, pending, a, b, n = self._late
        self._late = None
        self._wait_upload("late")
        keys = [k for k in self.CONSERVED if k not in self.EARLY]
        srcs = [pending[k] for k in keys]
        grown = [t.empty(n, dtype=x.dtype, device=self.dev) for x in srcs]
        ops.gatherGpu(self.ctx, perm, srcs, [g[a:b] for g in grown])
        for k, g in zip(keys, grown):
            f[k] = g

    def exchange_halos(self, names):
        """haloExchangeGpu (halos/exchange_halos_gpu.cuh:50-143): pack ranges per peer with one kernel, one NCCL
        all-to-all on the packed bytes, unpack into the contiguous halo slots"""
        if self.world == 1:
            return
        t, ops, dist, f = self.torch, self.ops, self.dist, self.f
        P, me = self.world, self.rank
        arrs = [f[k] for k in names]
        sb = [ops.packedBytes(self.halo_send[q], arrs) for q in range(P)]
        rb = [ops.packedBytes(self.halo_recv[q], arrs) for q in range(P)]
        sendbuf = t.empty(max(1, sum(sb)), dtype=t.uint8, device=self.dev)
        recvbuf = t.empty(max(1, sum(rb)), dtype=t.uint8, device=self.dev)
        ops.packRangesPeers(self.ctx, self.send_scan, self.send_begins, self.halo_send, arrs, sendbuf)
        dist.all_to_all_single(recvbuf[:sum(rb)], sendbuf[:sum(sb)], output_split_sizes=rb, input_split_sizes=sb,
                               group=self.group)
        ops.unpackRangesPeers(self.ctx, self.recv_scan, self.recv_begins, self.halo_recv, arrs, recvbuf)
        self.halo_bytes += sum(sb)

    # ---- HydroVeProp::computeForces / integrate
    def computeForces(self):
        t, ops, f = self.torch, self.ops, self.f
        T = self._tick
        T(None)
        self.sync(lazy=True)
        T("domain::sync")
        n, a, b = self.n_total, self.start, self.end
        for k in self.DEPENDENT:
            f[k] = t.zeros(n, dtype=self.real_t, device=self.dev)
        f["du"] = t.zeros(n, dtype=t.float64, device=self.dev)
        f["nc"] = t.zeros(n, dtype=t.int32, device=self.dev)
        hc = self._hc()
        self.nb, nc, self.fails = ops.findNeighbors(self.ctx, self.tree, f["x"], f["y"], f["z"], f["h"], a, b, self.box,
                                                    self.ng0, self.ngmax, True)
        f["nc"][a:b] = nc
        T("FindNeighbors")
        ops.computeXMass(self.ctx, a, b, f, self.box, hc, self.nb)
        T("XMass")
        self.exchange_halos(["xm"])
        T("halo")
        ops.computeVeDefGradh(self.ctx, a, b, f, self.box, hc, self.nb)
        T("Normalization & Gradh")
        self._finish_reorder()
        ops.computeEOS(self.ctx, a, b, f, hc)
        T("EquationOfState")
        self.exchange_halos(["vx", "vy", "vz", "prho", "c", "kx"])
        T("halo")
        ops.computeIadDivvCurlv(self.ctx, a, b, f, self.box, hc, self.nb)
        T("IadVelocityDivCurl")
        self.exchange_halos(["c11", "c12", "c13", "c22", "c23", "c33", "divv"])
        T("halo")
        ops.computeAVswitches(self.ctx, a, b, f, self.box, hc, self.nb)
        T("AVswitches")
        self.exchange_halos(["alpha"])
        T("halo")
        self.minDtCourant = None
        f_struct = ops.fields_struct(f)
        self.B.check(self.B.lib.sphb200_momentum_energy(self.ctx.h, self.dtype, C.c_uint32(a), C.c_uint32(b),
                                                        C.byref(self.box), C.byref(hc), self.B.ptr(self.nb),
                                                        C.byref(f_struct), None))
        if self.turb is not None:
            # TurbVeProp::computeForces (turb_ve.hpp:99-119): every rank advances the same OU sequence (same seed) and
            # stirs its own particles with the time step of the previous iteration; no communication
            self.turb.drive(self.ctx, a, b, f, self.minDt)
        T("MomentumAndEnergy")

    def attachTurbulence(self, turb):
        """turbulence-ve on P ranks: `turb` (turbulence.Turbulence) must be built from the same constants and seed on
        every rank, as the reference's TurbulenceData is"""
        self.turb = turb

    def integrate(self):
        t, ops, f, dist = self.torch, self.ops, self.f, self.dist
        hc = self._hc()
        a, b = self.start, self.end
        f_struct = ops.fields_struct(f)
        # local min{courant, rho, 1.1 dt} stays on the device and is all-reduced in-stream (ts_global.hpp:104-106)
        self.B.check(self.B.lib.sphb200_timestep(self.ctx.h, self.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(hc),
                                                 C.byref(f_struct), None))
        dptr = self.B.lib.sphb200_timestep_device_ptr(self.ctx.h)
        dt_dev = self._dt_view(dptr)
        if self.world > 1:
            dist.all_reduce(dt_dev[2:3], op=dist.ReduceOp.MIN, group=self.group)
        dt = float(dt_dev[2].item())
        self.ttot += dt
        self.minDt_m1, self.minDt = self.minDt, dt
        ops.computePositionsAndH(self.ctx, a, b, f, self.box, hc, self.minDt, self.minDt_m1)
        # drop halos: the owned range is the state carried to the next step
        for k in self.CONSERVED:
            f[k] = f[k][a:b]
        self._tick("Timestep+UpdateQuantities")

    def _dt_view(self, dptr):
        """torch view of the library's 3 device doubles (no copy), via the CUDA array interface"""
        t = self.torch
        if getattr(self, "_dt_cache", None) is None or self._dt_cache[0] != dptr:

            class _Holder:
                pass

            h = _Holder()
            h.__cuda_array_interface__ = {"shape": (3,), "typestr": "<f8", "data": (int(dptr), False), "version": 3}
            self._dt_cache = (dptr, t.as_tensor(h, device=self.dev))
        return self._dt_cache[1]

    def step(self):
        self.computeForces()
        self.integrate()

    def step_host(self, host, n_in, ranges=16):
        """one step on host-resident state: `host[k]` are pinned tensors holding this rank's n_in owned particles; on
        return they hold the (possibly different number of) owned particles after the step.  Returns that number.

        Same schedule as sphb200_sim_step_host (api.cu) with the halo exchanges in between: uploads in order of first
        use on a copy stream, m / next-step h / alpha sent back as soon as they are final, the Courant step evaluated
        with AVswitches and all-reduced before momentumEnergy, which then runs range by range with each range
        integrated and sent back while the next one computes.  Results are bit-identical to upload + step + download."""
        t, ops, f, dist, B = self.torch, self.ops, self.f, self.dist, self.B
        cur = t.cuda.current_stream()
        if not hasattr(self, "_copy_in"):
            self._copy_in, self._copy_out = t.cuda.Stream(), t.cuda.Stream()
        cin, cout = self._copy_in, self._copy_out
        late = [k for k in self.CONSERVED if k not in self.EARLY]
        groups = [("xyz", ["x", "y", "z"]), ("h", ["h"]), ("m", ["m"]), ("late", late)]
        cin.wait_stream(cur)
        cout.wait_stream(cur)
        events = {}
        with t.cuda.stream(cin):
            for name, keys in groups:
                for k in keys:
                    f[k] = host[k][:n_in].to(self.dev, non_blocking=True)
                    f[k].record_stream(cur)
                events[name] = cin.record_event()
        cout.wait_event(events["late"])  # the host buffers are reused for the results
        cur.wait_event(events.pop("xyz"))
        self._upload_events = events

        def send(k, src, lo, hi):
            """host[k][lo:hi] <- src, once everything issued so far on the compute stream is done"""
            ev = cur.record_event()
            with t.cuda.stream(cout):
                cout.wait_event(ev)
                host[k][lo:hi].copy_(src, non_blocking=True)
                src.record_stream(cout)

        self.sync(lazy=True)
        n, a, b = self.n_total, self.start, self.end
        n_own = b - a
        if any(host[k].numel() < n_own for k in self.CONSERVED):
            raise RuntimeError("step_host: host buffers too small for the particles this rank owns now")
        if self.world == 1:
            self._finish_reorder()
        send("m", f["m"][a:b], 0, n_own)
        for k in self.DEPENDENT:
            f[k] = t.zeros(n, dtype=self.real_t, device=self.dev)
        f["du"] = t.zeros(n, dtype=t.float64, device=self.dev)
        f["nc"] = t.zeros(n, dtype=t.int32, device=self.dev)
        hc = self._hc()
        self.nb, nc, self.fails = ops.findNeighbors(self.ctx, self.tree, f["x"], f["y"], f["z"], f["h"], a, b, self.box,
                                                    self.ng0, self.ngmax, True)
        f["nc"][a:b] = nc
        h_new = t.empty_like(f["h"])
        ops.updateSmoothingLengthInto(self.ctx, a, b, self.ng0, f["nc"], f["h"], h_new)
        send("h", h_new[a:b], 0, n_own)
        ops.computeXMass(self.ctx, a, b, f, self.box, hc, self.nb)
        self.exchange_halos(["xm"])
        ops.computeVeDefGradh(self.ctx, a, b, f, self.box, hc, self.nb)
        self._finish_reorder()
        ops.computeEOS(self.ctx, a, b, f, hc)
        self.exchange_halos(["vx", "vy", "vz", "prho", "c", "kx"])
        ops.computeIadDivvCurlv(self.ctx, a, b, f, self.box, hc, self.nb)
        self.exchange_halos(["c11", "c12", "c13", "c22", "c23", "c33", "divv"])
        ops.computeAVswitchesCourant(self.ctx, a, b, f, self.box, hc, self.nb)
        # computeTimestep (ts_global.hpp:96-112) before the last pair kernel
        f_struct = ops.fields_struct(f)
        B.check(B.lib.sphb200_timestep(self.ctx.h, self.dtype, C.c_uint32(a), C.c_uint32(b), C.byref(hc),
                                       C.byref(f_struct), None))
        dt_dev = self._dt_view(B.lib.sphb200_timestep_device_ptr(self.ctx.h))
        if self.world > 1:
            dist.all_reduce(dt_dev[2:3], op=dist.ReduceOp.MIN, group=self.group)
        dt = float(dt_dev[2].item())
        self.ttot += dt
        dt_stir = self.minDt
        self.minDt_m1, self.minDt = self.minDt, dt
        if self.turb is not None:
            self.turb.advance(self.ctx, dt_stir)
        send("alpha", f["alpha"][a:b], 0, n_own)
        self.exchange_halos(["alpha"])
        rest = [k for k in self.CONSERVED if k not in ("m", "h", "alpha")]
        per = ((n_own + ranges - 1) // ranges + 127) // 128 * 128
        for r in range(ranges):
            rb, re = min(b, a + r * per), min(b, a + (r + 1) * per)
            if re <= rb:
                break
            ops.computeMomentumEnergyRange(self.ctx, a, b, rb, re, r == 0, f, self.box, hc, self.nb)
            if self.turb is not None:
                self.turb.stir(self.ctx, rb, re, f)
            ops.computePositionsAndH(self.ctx, rb, re, f, self.box, hc, self.minDt, self.minDt_m1)
            ev = cur.record_event()
            with t.cuda.stream(cout):
                cout.wait_event(ev)
                for k in rest:
                    host[k][rb - a:re - a].copy_(f[k][rb:re], non_blocking=True)
        for k in self.CONSERVED:
            f[k].record_stream(cout)
            f[k] = f[k][a:b]
        cout.synchronize()
        self._upload_events = {}
        return n_own

    def conserved(self):
        """computeConservedQuantities (conserved_quantities.hpp:113-180): local fp64 reduction on the device, then one
        in-stream all-reduce of the 10 sums (the reference reduces to rank 0 with MPI_Reduce)"""
        t, ops = self.torch, self.ops
        f = dict(self.f)
        f.setdefault("nc", None)
        n_own = self.end - self.start
        a, b = (0, n_own) if f["x"].numel() == n_own else (self.start, self.end)
        if f.get("nc") is not None and f["nc"].numel() != f["x"].numel():
            f["nc"] = None
        fs = ops.fields_struct(f)
        self.B.check(self.B.lib.sphb200_conserved_quantities(self.ctx.h, self.dtype, C.c_uint32(a), C.c_uint32(b),
                                                             C.byref(self._hc()), C.byref(fs), None))
        dptr = self.B.lib.sphb200_conserved_device_ptr(self.ctx.h)

        class _Holder:
            pass

        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (10,), "typestr": "<f8", "data": (int(dptr), False), "version": 3}
        v = t.as_tensor(h, device=self.dev).clone()
        if self.world > 1:
            self.dist.all_reduce(v, op=self.dist.ReduceOp.SUM, group=self.group)
        return ops.conserved_dict(v.tolist())

    def _tick(self, name):
        if not getattr(self, "profile", False):
            return
        self.torch.cuda.synchronize()
        now = time.perf_counter()
        if name is not None:
            self.times[name] = self.times.get(name, 0.0) + (now - self._t0)
        self._t0 = now

    def download_owned(self, names=None):
        names = names or self.CONSERVED
        out = {}
        n_own = self.end - self.start
        for k in names:
            a = self.f[k]
            if a.numel() != n_own:
                a = a[self.start:self.end]
            out[k] = a.cpu().numpy()
        out["keys"] = self.own_keys.cpu().numpy().view(np.uint64)
        return out


    # ---- restart files (restart.py, csrc/checkpoint.cu)
    def save_restart(self, path, iteration):
        """every rank dumps its owned particles to `<path>.<rank>` (sphexa.cpp:232-246 writes the same content through
        parallel HDF5); collective only in the sense that all ranks should call it at the same step"""
        from . import restart
        B = self.B
        own = self.download_owned()
        n = len(own["x"])
        info = B.RestartInfo()
        info.iteration, info.numParticlesGlobal, info.numParticles = iteration, self.n_global, n
        info.dtype = self.dtype
        info.time, info.minDt, info.minDt_m1 = self.ttot, self.minDt, self.minDt_m1
        info.box = self.box
        fs = B.Fields()
        for k in B.FIELD_NAMES:
            setattr(fs, k, None)
        keep = []
        for k in DistHydroVeProp.CONSERVED:
            a = np.ascontiguousarray(own[k])
            keep.append(a)
            setattr(fs, k, a.ctypes.data)
        p = restart.rank_path(path, self.rank, self.world)
        hc = self._hc()
        bdt = self._restart_bdt(own)  # block-time-step propagators: ts:: attributes and the rung field
        B.check(B.lib.sphb200_restart_write_bdt(p.encode(), C.byref(info), C.byref(hc),
                                                self.turb.h if self.turb is not None else None, self.rank, self.world,
                                                C.byref(fs), C.byref(bdt) if bdt is not None else None))
        return p

    def _restart_bdt(self, own):
        return None

    def _restart_extra(self, files, pieces, arrays):
        """load_restart hook: pieces = [(file, first, count, offset)] of this rank's slice"""

    def load_restart(self, path):
        """this rank's share of a dump written by any number of ranks (file_init.hpp:93-140); the next sync
        redistributes.  Returns the iteration stored in the dump."""
        from . import restart
        B = self.B
        files = [restart.RestartFile(p) for p in restart.find_files(path)]
        try:
            counts = [int(f.info.numParticles) for f in files]
            total = sum(counts)
            if total != int(files[0].info.numParticlesGlobal):
                raise B.SphError("restart: the files of the dump are incomplete")
            if len(files) == self.world:
                first, count = sum(counts[:self.rank]), counts[self.rank]
            else:
                first, count = restart.slice_for_rank(total, self.rank, self.world)
            real = np.float64 if self.dtype == B.F64 else np.float32
            arrays = {k: np.empty(count, np.float64 if k in B.ALWAYS_F64 else real)
                      for k in DistHydroVeProp.CONSERVED}
            filled, start, pieces = 0, 0, []
            for f, c in zip(files, counts):
                lo, hi = max(first, start), min(first + count, start + c)
                if hi > lo:
                    f.read_fields(lo - start, hi - lo, self.dtype, arrays, offset=filled)
                    pieces.append((f, lo - start, hi - lo, filled))
                    filled += hi - lo
                start += c
            self._restart_extra(files, pieces, arrays)
            if self.turb is not None:
                # TurbulenceData::loadOrStore (turbulence_data.hpp:90-125): OU phases and the random engine
                phases = files[0].attribute("turbulence::phases")
                rng = files[0].attribute("rngEngineState", np.uint8)
                if phases is not None and rng is not None:
                    self.turb.setState(phases=phases, rng_state=rng.tobytes().rstrip(b"\0").decode())
            info = files[0].info
            self.n_global = total
            self.minDt, self.minDt_m1, self.ttot = info.minDt, info.minDt_m1, info.time
            self.lim, self.bc = list(info.box.lim), list(info.box.bc)
            self.box = self.pkg.make_box(self.lim, self.bc)
            for name, key in (("Kcour", "Kcour"), ("Krho", "Krho"), ("gamma", "gamma"), ("muiConst", "mui")):
                v = files[0].attribute(name)
                if v is not None:
                    self.s[key] = float(v[0])
            self.upload(arrays)
            return int(info.iteration)
        finally:
            for f in files:
                f.close()


# ---------------------------------------------------------------------------------------------- bench entry (N > 1)
def bench_main(args, pkg, METRIC, UNIT, ALGO_BYTES, load_peaks, ClockSampler):
    import torch
    import torch.distributed as dist
    from . import binding as B

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} needs torchrun with {args.gpus} ranks (WORLD_SIZE={world})")
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dtype = B.F64 if args.dtype == "f64" else B.MIXED

    # weak scaling: args.side^3 particles per GPU
    side = int(round(args.side * world ** (1.0 / 3.0)))
    n_global = side ** 3
    first, last = partition_range(n_global, rank, world)
    case = getattr(args, "case", "sedov")
    if case == "turbulence":
        # BASELINE.json configs[3]: stirred subsonic turbulence, weak scaling
        f, box, s = pkg.ic.turbulence_range(side, first, last, args.dtype)
    elif case == "noh":
        # BASELINE.json configs[2]: the lattice is cut to the sphere, numParticlesGlobal is the sum over the ranks
        xyz = pkg.ic.noh_range_coords(side, first, last)
        cnt = torch.tensor([xyz[0].size], dtype=torch.int64, device="cuda")
        dist.all_reduce(cnt)
        n_global = int(cnt.item())
        f, box, s = pkg.ic.noh_range_fields(*xyz, n_global, args.dtype)
    elif case == "sedov":
        f, box, s = pkg.ic.sedov_range(side, first, last, args.dtype)
    else:
        raise SystemExit("--case on more than one GPU: sedov, noh or turbulence")
    ctx = pkg.Context(local_rank)
    sim = DistHydroVeProp(pkg, ctx, box, s, dtype, rank, world, n_global)
    if case == "turbulence":
        from .turbulence import Turbulence
        sim.attachTurbulence(Turbulence())
    sim.upload(f)
    del f

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        sim.step()
    barrier()
    sim.sync_times = {}
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches
    sim.halo_bytes = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        sim.step()
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    launches = ctx.launches - launches0
    halo_bytes = sim.halo_bytes / args.steps
    clocks = sampler.stop() if rank == 0 else None
    own = torch.tensor([sim.end - sim.start, sim.n_total], dtype=torch.int64, device="cuda")
    gathered = [torch.zeros_like(own) for _ in range(world)]
    dist.all_gather(gathered, own)

    # per-phase wall times with synchronisation (separate, untimed-for-the-metric steps)
    sim.profile, sim.times = True, {}
    reps = 2
    for _ in range(reps):
        sim.step()
    phase_ms = {k: v / reps * 1e3 for k, v in sim.times.items()}
    sim.profile = False

    # end to end: every step uploads the owned conserved fields from pinned host memory and reads the new state back
    e2e = None
    if not getattr(args, "no_e2e", False):
        e2e_steps = max(2, min(args.steps, 5))
        cap = int((sim.end - sim.start) * 1.25) + 1024
        host = {k: torch.empty(cap, dtype=sim.f[k].dtype).pin_memory() for k in sim.CONSERVED}
        n_h = sim.end - sim.start
        for k in sim.CONSERVED:
            host[k][:n_h].copy_(sim.f[k][:n_h])
        torch.cuda.synchronize()
        h2d = d2h = 0
        for it in range(1 + e2e_steps):
            if it == 1:
                barrier()
                h2d = d2h = 0
                t0 = time.perf_counter()
            esz = sum(host[k].element_size() for k in sim.CONSERVED)
            h2d += n_h * esz
            n_h = sim.step_host(host, n_h)
            d2h += n_h * esz
        barrier()
        dt_e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(dt_e, op=dist.ReduceOp.MAX)
        dt_e = float(dt_e.item())
        e2e = {"value": n_global * e2e_steps / dt_e, "unit": UNIT, "h2d_bytes_per_step": h2d // e2e_steps,
               "d2h_bytes_per_step": d2h // e2e_steps, "ms_per_step": dt_e / e2e_steps * 1e3, "steps": e2e_steps,
               "note": "bytes are rank 0's; every rank moves its own share"}

    if rank == 0:
        peak, peak_src = load_peaks()
        value = n_global * args.steps / (total_ms * 1e-3)
        kernels = {k: v for k, v in phase_ms.items() if k not in ("domain::sync", "halo")}
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
                                   f"({args.side}^3 per GPU), NCCL halo exchange, full step",
                       "l2": "inputs larger than L2", "parallelism": f"sfc{world}"},
            "find_neighbors_ms": phase_ms.get("FindNeighbors"), "phase_ms": phase_ms,
            "particles_per_rank": [[int(g[0]), int(g[1])] for g in gathered],
            "halo_bytes_per_step_rank0": halo_bytes, "clocks": clocks, "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "kernel_ms": phase_ms[dom]},
            "e2e": e2e, "cpu_baseline": None,
        }
        print(json.dumps(line), flush=True)
        if sim._sync_profile:
            calls = args.steps + reps + (0 if e2e is None else 1 + e2e["steps"])
            print("sync profile (ms per call): " + json.dumps({k: round(v / calls * 1e3, 3)
                                                                 for k, v in sim.sync_times.items()}),
                  file=sys.stderr, flush=True)
    barrier()
    dist.destroy_process_group()
    return 0
