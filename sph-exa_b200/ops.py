"""Host-side mirror of the reference's GPU operator interface for the hot path.

Names and argument meaning follow the reference's `extern` GPU functions
(cstone/*_gpu.h, sph/sph_gpu.hpp); every function forwards to the C ABI of
libsphexa_b200.so with raw device pointers.  torch is only the device allocator.
"""
import ctypes as C
import torch

from . import binding as B
from .binding import lib, check, ptr, vp


def _i32(n):
    return C.c_int(int(n))


def real_dtype(dtype):
    return torch.float64 if dtype == B.F64 else torch.float32


# ---------------------------------------------------------------- cstone: sfc / primitives
def computeSfcKeysGpu(ctx, x, y, z, box, keys=None):
    """cstone::computeSfcKeysGpu (sfc/sfc_gpu.h:39)"""
    n = x.numel()
    if keys is None:
        keys = torch.empty(n, dtype=torch.int64, device=x.device)
    check(lib.sphb200_sfc_keys(ctx.h, ptr(x), ptr(y), ptr(z), C.c_size_t(n), C.byref(box), ptr(keys)))
    return keys


def assembleCuboid(ctx, key_start, key_end, box, multiplicity, x_block, y_block, z_block):
    """sphexa::assembleCuboid (main/src/init/grid.hpp:238-264): the part of the global box, tiled by `multiplicity`
    copies of a template block in [0,1)^3, that lies in the SFC range [key_start, key_end); reference order"""
    m = (C.c_int * 3)(*[int(v) for v in multiplicity])
    nb = x_block.numel()
    count = C.c_uint64(0)
    # first call sizes the output (nothing is written when the capacity is too small)
    rc = lib.sphb200_assemble_cuboid(ctx.h, C.c_uint64(int(key_start)), C.c_uint64(int(key_end)), C.byref(box), m,
                                     ptr(x_block), ptr(y_block), ptr(z_block), C.c_size_t(nb), None, None, None,
                                     C.c_size_t(0), C.byref(count))
    if rc not in (0, B.ERR_CAPACITY):
        check(rc)
    n = int(count.value)
    x, y, z = (torch.empty(n, dtype=torch.float64, device=x_block.device) for _ in range(3))
    if n:
        check(lib.sphb200_assemble_cuboid(ctx.h, C.c_uint64(int(key_start)), C.c_uint64(int(key_end)), C.byref(box), m,
                                          ptr(x_block), ptr(y_block), ptr(z_block), C.c_size_t(nb), ptr(x), ptr(y),
                                          ptr(z), C.c_size_t(n), C.byref(count)))
    return x, y, z


def cutSphere(ctx, radius, x, y, z):
    """sphexa::cutSphere (grid.hpp:267-300): the particles with |r| <= radius, order preserved (new tensors)"""
    x, y, z = x.clone(), y.clone(), z.clone()
    count = C.c_uint64(0)
    check(lib.sphb200_cut_sphere(ctx.h, C.c_double(radius), ptr(x), ptr(y), ptr(z), C.c_size_t(x.numel()),
                                 C.byref(count)))
    n = int(count.value)
    return x[:n].clone(), y[:n].clone(), z[:n].clone()


def minMaxGpu(ctx, x, y, z):
    """MinMaxGpu per axis (primitives_gpu.h:52) -> [xmin,xmax,ymin,ymax,zmin,zmax]"""
    out = (C.c_double * 6)()
    check(lib.sphb200_minmax3(ctx.h, ptr(x), ptr(y), ptr(z), C.c_size_t(x.numel()), out))
    return list(out)


def sortByKeyGpu(ctx, keys):
    """sequenceGpu + sortByKeyGpu (primitives_gpu.h:87-111): keys sorted in place, returns the ordering"""
    n = keys.numel()
    perm = torch.empty(n, dtype=torch.int32, device=keys.device)
    check(lib.sphb200_sort_by_key(ctx.h, ptr(keys), ptr(perm), C.c_size_t(n)))
    return perm


def sortPairsGpu(ctx, keys, values):
    """sortByKeyGpu with caller-provided values (primitives_gpu.h:96): keys and values sorted in place by key, stable"""
    lib.sphb200_sort_pairs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
    check(lib.sphb200_sort_pairs(ctx.h, ptr(keys), ptr(values), C.c_size_t(keys.numel())))


def gatherGpu(ctx, perm, srcs, dsts=None):
    """gatherGpu for a list of fields in one launch (primitives_gpu.h:62); dsts: optional preallocated outputs
    (contiguous views of perm.numel() elements)"""
    n = perm.numel()
    if dsts is None:
        dsts = [torch.empty_like(s) for s in srcs]
    k = len(srcs)
    sp = (vp * k)(*[s.data_ptr() for s in srcs])
    dp = (vp * k)(*[d.data_ptr() for d in dsts])
    eb = (C.c_int * k)(*[s.element_size() for s in srcs])
    check(lib.sphb200_gather(ctx.h, ptr(perm), C.c_size_t(n), k, sp, dp, eb))
    return dsts


# ---------------------------------------------------------------- cstone: tree
def computeNodeCountsGpu(ctx, leaves, keys):
    """computeNodeCountsGpu (tree/csarray_gpu.h:56)"""
    nl = leaves.numel() - 1
    counts = torch.empty(nl, dtype=torch.int32, device=keys.device)
    check(lib.sphb200_node_counts(ctx.h, ptr(leaves), nl, ptr(keys), C.c_size_t(keys.numel()), ptr(counts)))
    return counts


def rebalanceDecisionGpu(ctx, leaves, counts, bucket):
    nl = counts.numel()
    ops = torch.empty(nl, dtype=torch.int32, device=leaves.device)
    conv = C.c_int(0)
    check(lib.sphb200_rebalance_decision(ctx.h, ptr(leaves), ptr(counts), nl, C.c_uint(bucket), ptr(ops),
                                         C.byref(conv)))
    return ops, bool(conv.value)


def computeOctreeGpu(ctx, keys, bucket, capacity=None):
    """computeOctree (tree/csarray.hpp:454-468) on the device: returns (leaves, counts, iterations)"""
    n = keys.numel()
    cap = capacity or max(4096, 16 * (n // max(1, bucket)) + 4096)
    leaves = torch.empty(cap + 2, dtype=torch.int64, device=keys.device)
    counts = torch.empty(cap + 2, dtype=torch.int32, device=keys.device)
    nl, it = C.c_int(0), C.c_int(0)
    check(lib.sphb200_compute_octree(ctx.h, ptr(keys), C.c_size_t(n), C.c_uint(bucket), ptr(leaves), ptr(counts), cap,
                                     C.byref(nl), C.byref(it)))
    return leaves[:nl.value + 1].clone(), counts[:nl.value].clone(), it.value


def updateOctreeGpu(ctx, keys, bucket, leaves, counts):
    """updateOctreeGpu (tree/update_gpu.cuh:58-82): one rebalance + recount; returns (leaves, counts, converged)"""
    nl0 = counts.numel()
    cap = 8 * nl0 + 4096  # usual growth; a leaf may split into up to 4096 (csarray.hpp:305) -> retry below
    while True:
        lv = torch.empty(cap + 2, dtype=torch.int64, device=keys.device)
        ct = torch.empty(cap + 2, dtype=torch.int32, device=keys.device)
        lv[:nl0 + 1] = leaves
        ct[:nl0] = counts
        nl, conv = C.c_int(nl0), C.c_int(0)
        rc = lib.sphb200_update_octree(ctx.h, ptr(keys), C.c_size_t(keys.numel()), C.c_uint(bucket), ptr(lv), ptr(ct),
                                       cap, C.byref(nl), C.byref(conv))
        if rc == B.ERR_CAPACITY and nl.value > cap:
            cap = nl.value + 64
            continue
        check(rc)
        return lv[:nl.value + 1].clone(), ct[:nl.value].clone(), bool(conv.value)


class Octree:
    """device twin of cstone::OctreeData (tree/octree.hpp:318-386) + geometry + layout"""

    def __init__(self, ctx, leaves, box=None, counts=None):
        dev = leaves.device
        self.leaves = leaves
        self.numLeaves = nl = leaves.numel() - 1
        self.numInternal = ni = (nl - 1) // 7
        self.numNodes = nn = nl + ni
        i32 = dict(dtype=torch.int32, device=dev)
        self.prefixes = torch.zeros(nn, dtype=torch.int64, device=dev)
        self.childOffsets = torch.zeros(nn + 1, **i32)
        self.parents = torch.zeros(max(1, (nn - 1) // 8), **i32)
        self.levelRange = torch.zeros(23, **i32)
        self.internalToLeaf = torch.zeros(nn, **i32)
        self.leafToInternal = torch.zeros(nn, **i32)
        check(lib.sphb200_build_octree(ctx.h, ptr(leaves), nl, ptr(self.prefixes), ptr(self.childOffsets),
                                       ptr(self.parents), ptr(self.levelRange), ptr(self.internalToLeaf),
                                       ptr(self.leafToInternal)))
        self.centers = self.sizes = self.layout = None
        if box is not None:
            self.centers = torch.empty((nn, 3), dtype=torch.float64, device=dev)
            self.sizes = torch.empty((nn, 3), dtype=torch.float64, device=dev)
            check(lib.sphb200_node_centers(ctx.h, ptr(self.prefixes), nn, C.byref(box), ptr(self.centers),
                                           ptr(self.sizes)))
        if counts is not None:
            self.layout = torch.empty(nl + 1, **i32)
            check(lib.sphb200_layout(ctx.h, ptr(counts), nl, ptr(self.layout)))

    def view(self):
        t = B.TreeView()
        t.numLeafNodes, t.numNodes = self.numLeaves, self.numNodes
        t.prefixes, t.childOffsets = ptr(self.prefixes), ptr(self.childOffsets)
        t.internalToLeaf, t.levelRange = ptr(self.internalToLeaf), ptr(self.levelRange)
        t.leaves, t.layout = ptr(self.leaves), ptr(self.layout)
        t.centers, t.sizes = ptr(self.centers), ptr(self.sizes)
        return t


# ---------------------------------------------------------------- neighbors
def neighborsSize(num_targets, ngmax):
    return lib.sphb200_neighbors_size(num_targets, ngmax)


def findNeighbors(ctx, tree, x, y, z, h, first, last, box, ng0, ngmax, iterate_h=True):
    """neighbor search + coupled h iteration (sph/find_neighbors.hpp:10-45); h is updated in place.

    returns (neighbor list buffer, nc[last-first] incl. self, numFails)"""
    dev = x.device
    nt = last - first
    nb = torch.empty(max(1, neighborsSize(nt, ngmax)), dtype=torch.int32, device=dev)
    nc = torch.zeros(max(1, nt), dtype=torch.int32, device=dev)
    fails = C.c_uint(0)
    dtype = B.F64 if h.dtype == torch.float64 else B.MIXED
    tv = tree.view()
    check(lib.sphb200_find_neighbors(ctx.h, C.byref(tv), ptr(x), ptr(y), ptr(z), ptr(h), dtype, C.c_uint32(first),
                                     C.c_uint32(last), C.byref(box), C.c_uint(ng0), C.c_uint(ngmax), int(iterate_h),
                                     ptr(nb), ptr(nc), C.byref(fails)))
    return nb, nc[:nt], fails.value


def neighborsRowMajor(ctx, nb, nc, ngmax):
    nt = nc.numel()
    out = torch.zeros((max(nt, 1), ngmax), dtype=torch.int32, device=nb.device)
    check(lib.sphb200_neighbors_to_rowmajor(ctx.h, ptr(nb), ptr(nc), C.c_uint32(nt), C.c_uint(ngmax), ptr(out)))
    return out[:nt]


def neighborsFromRowMajor(ctx, rowmajor, nc, ngmax):
    """reference CPU layout [nt, ngmax] (+ nc incl. self) -> list buffer read by the pair operators"""
    nt = nc.numel()
    nb = torch.zeros(max(1, neighborsSize(nt, ngmax)), dtype=torch.int32, device=rowmajor.device)
    check(lib.sphb200_neighbors_from_rowmajor(ctx.h, ptr(rowmajor), ptr(nc), C.c_uint32(nt), C.c_uint(ngmax), ptr(nb)))
    return nb


# ---------------------------------------------------------------- SPH-VE operators
def fields_struct(d):
    """dict name -> torch tensor  =>  sphb200_fields"""
    f = B.Fields()
    for n in B.FIELD_NAMES:
        t = d.get(n)
        setattr(f, n, t.data_ptr() if t is not None else None)
    f.n = d["x"].numel()
    return f


def hydro_const(K, wh, whd, ng0=100, ngmax=150, Kcour=0.2, Krho=0.06, gamma=5.0 / 3.0, muiConst=10.0, alphamin=0.05,
                alphamax=1.0, decay_constant=0.2, Atmin=0.1, Atmax=0.2, ramp=None, minDt=1e-6, minDt_m1=1e-6,
                avClean=False):
    hc = B.HydroConst()
    hc.K, hc.ng0, hc.ngmax, hc.Kcour, hc.Krho, hc.gamma, hc.muiConst = K, ng0, ngmax, Kcour, Krho, gamma, muiConst
    hc.alphamin, hc.alphamax, hc.decay_constant = alphamin, alphamax, decay_constant
    hc.Atmin, hc.Atmax, hc.ramp = Atmin, Atmax, (1.0 / (Atmax - Atmin) if ramp is None else ramp)
    hc.minDt, hc.minDt_m1 = minDt, minDt_m1
    hc.wh = wh.data_ptr() if hasattr(wh, "data_ptr") else wh
    hc.whd = whd.data_ptr() if hasattr(whd, "data_ptr") else whd
    hc.avClean = int(avClean)
    return hc


def _dt(d):
    return B.F64 if d["h"].dtype == torch.float64 else B.MIXED


def computeXMass(ctx, first, last, d, box, hc, nb):
    f = fields_struct(d)
    check(lib.sphb200_xmass(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box), C.byref(hc), ptr(nb),
                            C.byref(f)))


def computeVeDefGradh(ctx, first, last, d, box, hc, nb):
    f = fields_struct(d)
    check(lib.sphb200_ve_def_gradh(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box), C.byref(hc),
                                   ptr(nb), C.byref(f)))


def computeEOS(ctx, first, last, d, hc):
    f = fields_struct(d)
    check(lib.sphb200_eos(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(hc), C.byref(f)))


def computeIadDivvCurlv(ctx, first, last, d, box, hc, nb):
    f = fields_struct(d)
    check(lib.sphb200_iad_divv_curlv(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box), C.byref(hc),
                                     ptr(nb), C.byref(f)))


def computeAVswitches(ctx, first, last, d, box, hc, nb):
    f = fields_struct(d)
    check(lib.sphb200_av_switches(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box), C.byref(hc),
                                  ptr(nb), C.byref(f)))


def computeMomentumEnergy(ctx, first, last, d, box, hc, nb):
    """returns minDtCourant"""
    f = fields_struct(d)
    out = C.c_double(0)
    check(lib.sphb200_momentum_energy(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box), C.byref(hc),
                                      ptr(nb), C.byref(f), C.byref(out)))
    return out.value


def computeTimestep(ctx, first, last, d, hc):
    """(minDtCourant, minDtRho, local min incl. 1.1*minDt) — ts_global.hpp:71-112 before the global MIN"""
    f = fields_struct(d)
    out = (C.c_double * 3)()
    check(lib.sphb200_timestep(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(hc), C.byref(f), out))
    return tuple(out)


def computePositionsAndH(ctx, first, last, d, box, hc, dt, dt_m1):
    f = fields_struct(d)
    check(lib.sphb200_positions_update_h(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box),
                                         C.byref(hc), C.c_double(dt), C.c_double(dt_m1), C.byref(f)))


# ---------------------------------------------------------------- halo pack / unpack
def packRanges(ctx, range_scan, range_begin, total, srcs, packed):
    k = len(srcs)
    sp = (vp * k)(*[s.data_ptr() for s in srcs])
    eb = (C.c_int * k)(*[s.element_size() for s in srcs])
    check(lib.sphb200_pack_ranges(ctx.h, ptr(range_scan), ptr(range_begin), range_begin.numel(), C.c_uint32(total), k,
                                  sp, eb, ptr(packed)))


def unpackRanges(ctx, range_scan, range_begin, total, dsts, packed):
    k = len(dsts)
    dp = (vp * k)(*[s.data_ptr() for s in dsts])
    eb = (C.c_int * k)(*[s.element_size() for s in dsts])
    check(lib.sphb200_unpack_ranges(ctx.h, ptr(range_scan), ptr(range_begin), range_begin.numel(), C.c_uint32(total),
                                    k, dp, eb, ptr(packed)))


def packRangesPeers(ctx, range_scan, range_begin, peer_counts, srcs, packed):
    """all peers of one halo exchange in one launch (sphb200_pack_ranges_peers); peer_counts: host list"""
    k, P = len(srcs), len(peer_counts)
    sp = (vp * k)(*[s.data_ptr() for s in srcs])
    eb = (C.c_int * k)(*[s.element_size() for s in srcs])
    pc = (C.c_uint32 * P)(*peer_counts)
    check(lib.sphb200_pack_ranges_peers(ctx.h, ptr(range_scan), ptr(range_begin), range_begin.numel(), P, pc, k, sp, eb,
                                        ptr(packed), None))


def unpackRangesPeers(ctx, range_scan, range_begin, peer_counts, dsts, packed):
    k, P = len(dsts), len(peer_counts)
    dp = (vp * k)(*[s.data_ptr() for s in dsts])
    eb = (C.c_int * k)(*[s.element_size() for s in dsts])
    pc = (C.c_uint32 * P)(*peer_counts)
    check(lib.sphb200_unpack_ranges_peers(ctx.h, ptr(range_scan), ptr(range_begin), range_begin.numel(), P, pc, k, dp,
                                          eb, ptr(packed), None))


def packedBytes(total, srcs):
    return sum((total * s.element_size() + 7) // 8 * 8 for s in srcs)


# ---------------------------------------------------------------- observables
CONSERVED_NAMES = ["ecin", "eint", "egrav", "linmom_x", "linmom_y", "linmom_z", "angmom_x", "angmom_y", "angmom_z", "ncsum"]


def computeConservedQuantities(ctx, first, last, d, hc):
    """local part of sphexa::computeConservedQuantities (conserved_quantities.hpp:50-180): dict of the 10 sums over
    [first, last) plus etot, |linmom|, |angmom| as the reference derives them (lines 165-175)"""
    f = fields_struct(d)
    out = (C.c_double * 10)()
    check(lib.sphb200_conserved_quantities(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(hc), C.byref(f),
                                           out))
    return conserved_dict(list(out))


def conserved_dict(v):
    r = dict(zip(CONSERVED_NAMES, v))
    r["etot"] = r["ecin"] + r["eint"] + r["egrav"]
    r["linmom"] = (r["linmom_x"] ** 2 + r["linmom_y"] ** 2 + r["linmom_z"] ** 2) ** 0.5
    r["angmom"] = (r["angmom_x"] ** 2 + r["angmom_y"] ** 2 + r["angmom_z"] ** 2) ** 0.5
    return r


# ---------------------------------------------------------------- groups / block time steps (csrc/groups.cu)
class Groups:
    """device twin of sph::GroupData (sph/include/sph/groups.hpp): delimiters or separate start/end arrays"""

    def __init__(self, first, last, start, end):
        self.firstBody, self.lastBody = int(first), int(last)
        self.start, self.end = start, end
        self.numGroups = start.numel()

    def view(self, lo=0, hi=None):
        """makeSlicedView (groups.hpp:50-58): groups [lo, hi)"""
        hi = self.numGroups if hi is None else hi
        v = B.GroupView()
        v.firstBody, v.lastBody, v.numGroups = self.firstBody, self.lastBody, hi - lo
        v.groupStart = self.start.data_ptr() + 4 * lo
        v.groupEnd = self.end.data_ptr() + 4 * lo
        return v


def computeGroups(ctx, first, last, x, y, z, tree, box, tolFactor=2.0):
    """computeSpatialGroups (sph/include/sph/groups.cu:30-47); tree: ops.Octree with layout"""
    n = last - first
    cap = n // 32 + 2
    while True:
        delim = torch.empty(cap + 1, dtype=torch.int32, device=x.device)
        ng = C.c_uint32(0)
        rc = lib.sphb200_compute_groups(ctx.h, first, last, ptr(x), ptr(y), ptr(z), ptr(tree.leaves), tree.numLeaves,
                                        ptr(tree.layout), C.byref(box), C.c_float(tolFactor), ptr(delim), cap,
                                        C.byref(ng))
        if rc == B.ERR_CAPACITY:
            cap = ng.value + 1
            continue
        check(rc)
        break
    delim = delim[:ng.value + 1]
    return Groups(first, last, delim[:-1], delim[1:])


def groupDivvTimestep(ctx, Krho, grp_view, divv, groupDt):
    check(lib.sphb200_group_divv_timestep(ctx.h, B.F64 if divv.dtype == torch.float64 else B.MIXED, C.c_float(Krho),
                                          C.byref(grp_view), ptr(divv), ptr(groupDt)))


def groupAccTimestep(ctx, etaAcc, grp_view, ax, ay, az, groupDt):
    check(lib.sphb200_group_acc_timestep(ctx.h, B.F64 if ax.dtype == torch.float64 else B.MIXED, C.c_float(etaAcc),
                                         C.byref(grp_view), ptr(ax), ptr(ay), ptr(az), ptr(groupDt)))


def groupMinTimestep(ctx, grp_view, dtParticle, groupDt):
    check(lib.sphb200_group_min_timestep(ctx.h, B.F64 if dtParticle.dtype == torch.float64 else B.MIXED,
                                         C.byref(grp_view), ptr(dtParticle), ptr(groupDt)))


def storeRung(ctx, grp_view, rung, particleRungs):
    check(lib.sphb200_store_rung(ctx.h, C.byref(grp_view), C.c_uint8(rung), ptr(particleRungs)))


def extractGroups(ctx, grp, indices, first, last):
    """extractGroupGpu (groups.hpp:32-47): Groups made of grp[indices[first:last]]"""
    n = last - first
    out = torch.empty(2 * max(n, 1), dtype=torch.int32, device=indices.device)
    v = grp.view()
    check(lib.sphb200_extract_groups(ctx.h, C.byref(v), ptr(indices), first, last, ptr(out), ptr(out[max(n, 1):])))
    return Groups(grp.firstBody, grp.lastBody, out[:n], out[max(n, 1):max(n, 1) + n])


def targetsFromGroups(ctx, grp_view):
    """ascending index list of the particles in the groups of grp_view"""
    n = grp_view.lastBody - grp_view.firstBody
    dev = torch.device("cuda", torch.cuda.current_device())
    t = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    nt = C.c_uint32(0)
    check(lib.sphb200_targets_from_groups(ctx.h, C.byref(grp_view), ptr(t), C.byref(nt)))
    return t[:nt.value]


def markGroups(ctx, grp_view, n):
    """uint8 mask [n]: 1 for the members of the groups"""
    dev = torch.device("cuda", torch.cuda.current_device())
    mask = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
    check(lib.sphb200_mark_groups(ctx.h, C.byref(grp_view), C.c_uint32(n), ptr(mask)))
    return mask[:n]


class scope:
    """with ops.scope(ctx, active=mask, searchExtFactor=f, dtParticle=t): the GroupView argument of the reference's GPU
    operators (sph_gpu.hpp:13-82) and OctreeNsView::searchExtFactor as context state (sphb200_set_scope)"""

    def __init__(self, ctx, active=None, searchExtFactor=1.0, dtParticle=None, gatherMask=None):
        self.ctx = ctx
        self.s = B.OpScope()
        self.s.active = active.data_ptr() if active is not None else None
        self.s.searchExtFactor = searchExtFactor
        self.s.dtParticle = dtParticle.data_ptr() if dtParticle is not None else None
        self.s.gatherMask = gatherMask.data_ptr() if gatherMask is not None else None
        self.keep = (active, dtParticle, gatherMask)

    def __enter__(self):
        check(lib.sphb200_set_scope(self.ctx.h, C.byref(self.s)))
        return self

    def __exit__(self, *exc):
        check(lib.sphb200_set_scope(self.ctx.h, None))
        return False


def sortGroupDt(ctx, groupDt, numGroups=None):
    """sortGroupDt (sph/include/sph/ts_rungs.hpp:55-66): sorts groupDt[:numGroups] in place, returns the ordering"""
    n = groupDt.numel() if numGroups is None else numGroups
    idx = torch.empty(max(n, 1), dtype=torch.int32, device=groupDt.device)
    check(lib.sphb200_sort_group_dt(ctx.h, ptr(groupDt), ptr(idx), C.c_uint32(n)))
    return idx[:n]


def rungRanges(ctx, minDt, groupDt, numGroups, numRungs, fastFraction=0.4):
    """findRungRanges + timestepRangeGpu (ts_rungs.hpp:69-75,115-129): ([5 ranges], (dt[0], dt[fast]))"""
    rr = (C.c_uint32 * 5)()
    mdt = (C.c_float * 2)()
    check(lib.sphb200_rung_ranges(ctx.h, C.c_float(minDt), ptr(groupDt), C.c_uint32(numGroups), int(numRungs),
                                  C.c_float(fastFraction), rr, mdt))
    return list(rr), (mdt[0], mdt[1])


def _dt4(dt_m1):
    return (C.c_float * 4)(*[float(v) for v in dt_m1])


def driftPositions(ctx, dtype, grp_view, dt, dt_back, dt_m1, rung, hc, f):
    """driftPositionsGpu (sph/include/sph/positions_gpu.cu:47-111)"""
    check(lib.sphb200_drift_positions(ctx.h, dtype, C.byref(grp_view), C.c_float(dt), C.c_float(dt_back), _dt4(dt_m1),
                                      ptr(rung), C.byref(hc), C.byref(f)))


def computePositionsGroups(ctx, dtype, grp_view, dt, dt_m1, rung, box, hc, f):
    """computePositionsGpu (positions_gpu.cu:118-180)"""
    check(lib.sphb200_compute_positions_groups(ctx.h, dtype, C.byref(grp_view), C.c_float(dt), _dt4(dt_m1), ptr(rung),
                                               C.byref(box), C.byref(hc), C.byref(f)))


def updateSmoothingLengthGroups(ctx, grp_view, ng0, nc, h):
    """updateSmoothingLengthGpu (sph/include/sph/update_h_gpu.cu:39-60)"""
    check(lib.sphb200_update_h_groups(ctx.h, B.F64 if h.dtype == torch.float64 else B.MIXED, C.byref(grp_view),
                                      C.c_uint(ng0), ptr(nc), ptr(h)))


# ---------------------------------------------------------------- early Courant step / ranged last stage
def computeAVswitchesCourant(ctx, first, last, d, box, hc, nb):
    """AVswitches that also leaves momentumEnergy's Courant step on the device (read it with computeTimestep)"""
    f = fields_struct(d)
    check(lib.sphb200_av_switches_courant(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(box),
                                          C.byref(hc), ptr(nb), C.byref(f)))


def computeMomentumEnergyRange(ctx, first, last, begin, end, pack, d, box, hc, nb):
    """momentumEnergy for the targets [begin,end) of [first,last); pack=False reuses the previous call's records"""
    f = fields_struct(d)
    check(lib.sphb200_momentum_energy_range(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.c_uint32(begin),
                                            C.c_uint32(end), int(bool(pack)), C.byref(box), C.byref(hc), ptr(nb),
                                            C.byref(f)))


def updateSmoothingLengthInto(ctx, first, last, ng0, nc, h, h_new):
    """h_new[i] = updateH(ng0, nc[i], h[i]) for [first,last), h untouched"""
    check(lib.sphb200_update_h_into(ctx.h, B.F64 if h.dtype == torch.float64 else B.MIXED, C.c_uint32(first),
                                    C.c_uint32(last), C.c_uint(ng0), ptr(nc), ptr(h), ptr(h_new)))


def machSquareSum(ctx, first, last, d):
    """localMachSquareSum (main/src/observables/turbulence_mach_rms.hpp:38-56) over [first,last)"""
    f = fields_struct(d)
    out = C.c_double(0)
    check(lib.sphb200_mach_square_sum(ctx.h, _dt(d), C.c_uint32(first), C.c_uint32(last), C.byref(f), C.byref(out)))
    return out.value
