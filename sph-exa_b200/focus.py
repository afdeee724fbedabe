"""Python mirror of the focus-tree part of the C ABI (csrc/focus.cu).

`FocusedOctree` has the interface of cstone::FocusedOctreeSingleNode (focus/octree_focus.hpp:219-288): `update(box,
keys, focusStart, focusEnd, mandatoryKeys)`, `treeLeaves()`, `leafCounts()`; the staged calls follow
cstone::FocusedOctree (focus/octree_focus_mpi.hpp: updateMinMac, updateTree, updateCounts).  The operator-level
functions take device tensors of a linked octree like the reference's *Gpu functions.  Everything here only marshals
arguments: the work is in libsphexa_b200.so.
"""
import ctypes as C

import numpy as np

from . import binding as B
from .binding import lib, check, vp, ptr

u64 = C.c_uint64


class FocusView(C.Structure):
    _fields_ = [("numLeaves", C.c_int), ("numNodes", C.c_int), ("lastStatus", C.c_int), ("leaves", vp),
                ("leafCounts", vp), ("prefixes", vp), ("childOffsets", vp), ("parents", vp), ("levelRange", vp),
                ("internalToLeaf", vp), ("leafToInternal", vp), ("counts", vp), ("macs", vp), ("centers", vp)]


lib.sphb200_min_mac_r2.argtypes = [vp, vp, C.c_int, C.c_float, vp, vp]
lib.sphb200_mark_macs.argtypes = [vp, vp, vp, vp, C.c_int, vp, u64, u64, vp]
lib.sphb200_rebalance_decision_essential.argtypes = [vp, vp, vp, vp, vp, vp, u64, u64, C.c_uint, C.c_int, vp]
lib.sphb200_enforce_keys.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, vp]
lib.sphb200_protect_ancestors.argtypes = [vp, vp, vp, C.c_int, vp, vp]
lib.sphb200_find_peers_mac.argtypes = [C.c_int, C.c_int, vp, vp, C.c_int, vp, C.c_float, vp]
lib.sphb200_focus_create.argtypes = [vp, C.c_uint, C.c_float, vp]
lib.sphb200_focus_destroy.argtypes = [vp]
lib.sphb200_focus_destroy.restype = None
lib.sphb200_focus_update.argtypes = [vp, vp, vp, C.c_size_t, u64, u64, vp, C.c_int, vp]
lib.sphb200_focus_update_tree.argtypes = [vp, u64, u64, vp, C.c_int, vp]
lib.sphb200_focus_update_min_mac.argtypes = [vp, vp, u64, u64]
lib.sphb200_focus_update_counts.argtypes = [vp, vp, C.c_size_t]
lib.sphb200_focus_set_leaf_counts.argtypes = [vp, vp]
lib.sphb200_focus_get_view.argtypes = [vp, vp]

#: ResolutionStatus (focus/rebalance.hpp:178-187)
CONVERGED, CANCEL_MERGE, REBALANCE, FAILED = 0, 1, 2, 3


def invThetaMinMac(theta):
    """invThetaMinMac (traversal/macs.hpp:44): 1/theta + 0.5 in float arithmetic"""
    return float(np.float32(1.0) / np.float32(theta) + np.float32(0.5))


def findPeersMac(myRank, bounds, leaves, box, invThetaEff):
    """findPeersMac (traversal/peers.hpp:63-120); bounds: P + 1 keys, leaves: the global tree; host arrays"""
    bounds = np.ascontiguousarray(bounds, np.uint64)
    leaves = np.ascontiguousarray(leaves, np.uint64)
    P = bounds.size - 1
    peers = np.zeros(P, np.int32)
    check(lib.sphb200_find_peers_mac(myRank, P, bounds.ctypes.data, leaves.ctypes.data, leaves.size - 1, C.byref(box),
                                     C.c_float(invThetaEff), peers.ctypes.data))
    return np.nonzero(peers)[0].tolist()


def _keys_h(keys):
    a = np.ascontiguousarray(np.asarray(keys, np.uint64))
    return a, (a.ctypes.data if a.size else None)


def minMacR2(ctx, prefixes, invThetaEff, box, centers4):
    check(lib.sphb200_min_mac_r2(ctx.h, ptr(prefixes), prefixes.numel(), C.c_float(invThetaEff), C.byref(box),
                                 ptr(centers4)))


def markMacs(ctx, prefixes, parents, centers4, box, focusStart, focusEnd, markings):
    check(lib.sphb200_mark_macs(ctx.h, ptr(prefixes), ptr(parents), ptr(centers4), prefixes.numel(), C.byref(box),
                                int(focusStart), int(focusEnd), ptr(markings)))


def rebalanceDecisionEssential(ctx, prefixes, childOffsets, parents, counts, macs, focusStart, focusEnd, bucketSize,
                               nodeOps):
    check(lib.sphb200_rebalance_decision_essential(ctx.h, ptr(prefixes), ptr(childOffsets), ptr(parents), ptr(counts),
                                                   ptr(macs), int(focusStart), int(focusEnd), bucketSize,
                                                   prefixes.numel(), ptr(nodeOps)))


def enforceKeys(ctx, keys, prefixes, childOffsets, parents, nodeOps):
    status = C.c_int(0)
    check(lib.sphb200_enforce_keys(ctx.h, ptr(keys), keys.numel(), ptr(prefixes), ptr(childOffsets), ptr(parents),
                                   ptr(nodeOps), C.byref(status)))
    return status.value


def protectAncestors(ctx, prefixes, parents, nodeOps):
    conv = C.c_int(0)
    check(lib.sphb200_protect_ancestors(ctx.h, ptr(prefixes), ptr(parents), prefixes.numel(), ptr(nodeOps),
                                        C.byref(conv)))
    return bool(conv.value)


class FocusedOctree:
    """cstone::FocusedOctreeSingleNode on the device"""

    def __init__(self, ctx, bucketSize, theta):
        self.ctx = ctx
        self.h = vp()
        check(lib.sphb200_focus_create(ctx.h, bucketSize, C.c_float(theta), C.byref(self.h)))

    def close(self):
        if self.h:
            lib.sphb200_focus_destroy(self.h)
            self.h = vp()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def update(self, box, keys, focusStart, focusEnd, mandatoryKeys=()):
        """keys: sorted device tensor (int64 view of the uint64 keys)"""
        mk, mkp = _keys_h(mandatoryKeys)
        conv = C.c_int(0)
        check(lib.sphb200_focus_update(self.h, C.byref(box), ptr(keys), keys.numel(), int(focusStart), int(focusEnd),
                                       mkp, mk.size, C.byref(conv)))
        return bool(conv.value)

    def updateTree(self, focusStart, focusEnd, mandatoryKeys=()):
        mk, mkp = _keys_h(mandatoryKeys)
        conv = C.c_int(0)
        check(lib.sphb200_focus_update_tree(self.h, int(focusStart), int(focusEnd), mkp, mk.size, C.byref(conv)))
        return bool(conv.value)

    def updateMinMac(self, box, focusStart, focusEnd):
        check(lib.sphb200_focus_update_min_mac(self.h, C.byref(box), int(focusStart), int(focusEnd)))

    def updateCounts(self, keys):
        check(lib.sphb200_focus_update_counts(self.h, ptr(keys), keys.numel()))

    def setLeafCounts(self, leafCounts):
        assert leafCounts.numel() == self.view().numLeaves
        check(lib.sphb200_focus_set_leaf_counts(self.h, ptr(leafCounts)))

    def view(self):
        v = FocusView()
        check(lib.sphb200_focus_get_view(self.h, C.byref(v)))
        return v

    _rt = None

    def _download(self, p, count, dtype):
        out = np.empty(count, dtype)
        if count:
            if FocusedOctree._rt is None:
                FocusedOctree._rt = C.CDLL("libcudart.so.12")
            self.ctx.sync()
            rc = FocusedOctree._rt.cudaMemcpy(out.ctypes.data_as(vp), vp(p), C.c_size_t(out.nbytes), 2)  # D2H
            if rc != 0:
                raise B.SphError(f"cudaMemcpy D2H failed: {rc}")
        return out

    def treeLeaves(self):
        v = self.view()
        return self._download(v.leaves, v.numLeaves + 1, np.uint64)

    def leafCounts(self):
        v = self.view()
        return self._download(v.leafCounts, v.numLeaves, np.uint32)

    def nodeArrays(self):
        """host copies of the linked octree with its counts, MACs and centers"""
        v = self.view()
        nn = v.numNodes
        out = dict(numLeaves=v.numLeaves, numNodes=nn, prefixes=self._download(v.prefixes, nn, np.uint64),
                   childOffsets=self._download(v.childOffsets, nn, np.int32),
                   parents=self._download(v.parents, max(1, (nn - 1) // 8), np.int32),
                   leafToInternal=self._download(v.leafToInternal, nn, np.int32),
                   internalToLeaf=self._download(v.internalToLeaf, nn, np.int32))
        if v.counts:
            out["counts"] = self._download(v.counts, nn, np.uint32)
        if v.macs:
            out["macs"] = self._download(v.macs, nn, np.int8)
            out["centers"] = self._download(v.centers, 4 * nn, np.float64).reshape(nn, 4)
        return out
