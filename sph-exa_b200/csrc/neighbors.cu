// Neighbor search: ONE traversal per step that writes compact, ascending-j neighbor lists.
//
// The reference GPU path re-runs its warp traversal inside every pair kernel (5x per step) with a
// ring-buffer queue in global memory (domain/include/cstone/traversal/find_neighbors.cuh:201-341).
// Here one warp owns one slab of 32 SFC-consecutive targets, walks the linked octree depth-first in
// ascending SFC order with its stack in shared memory, stages each overlapping leaf's particles in
// shared memory with coalesced loads and tests them against all 32 targets via broadcast reads.
// Membership arithmetic follows cstone::findNeighbors / distanceSq (findneighbors.hpp:51-165) in the
// same operation order; this file is compiled with --fmad=false so that d^2 < 4h^2 is bit-exact
// against the reference CPU build.
#include "common.cuh"

namespace sphb200
{

constexpr int kNbWarps     = 4;   // warps per CTA
constexpr int kStackSize   = 160; // 7 * 21 levels + 8, singleTraversal uses 128 (traversal.hpp:80)
constexpr int kMaxHIter    = 10;  // sph/find_neighbors.hpp:20

struct DevTree
{
    const TreeNodeIndex* childOffsets;
    const TreeNodeIndex* internalToLeaf;
    const LocalIndex* layout;
    const double* centers;
    const double* sizes;
    const double* leafBounds; // [numLeaves][6] tight particle extents: xlo ylo zlo xhi yhi zhi
};

//! @brief tight bounding box of the particles in every leaf (one warp per leaf); empty leaf: lo=+inf, hi=-inf
__global__ void __launch_bounds__(128) leafBoundsKernel(const LocalIndex* __restrict__ layout, int numLeaves,
                                                        const double* __restrict__ x, const double* __restrict__ y,
                                                        const double* __restrict__ z, double* __restrict__ bounds)
{
    int leaf = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    int lane = threadIdx.x & 31;
    if (leaf >= numLeaves) return;
    LocalIndex jb = layout[leaf], je = layout[leaf + 1];
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (LocalIndex j = jb + lane; j < je; j += 32)
    {
        double p[3] = {x[j], y[j], z[j]};
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            lo[d] = fmin(lo[d], p[d]);
            hi[d] = fmax(hi[d], p[d]);
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            lo[d] = fmin(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
            hi[d] = fmax(hi[d], __shfl_xor_sync(0xffffffffu, hi[d], o));
        }
    }
    if (lane < 3) bounds[6 * size_t(leaf) + lane] = lo[lane];
    else if (lane < 6) bounds[6 * size_t(leaf) + lane] = hi[lane - 3];
}

//! @brief sph::updateH (sph/include/sph/kernels.hpp:26-32)
template<class T>
__device__ __forceinline__ T updateH(unsigned ng0, unsigned nc, T h)
{
    const T c0  = T(1023.0);
    const T exp = T(1.0 / 10.0);
    return h * T(0.5) * pow(T(1) + c0 * T(ng0) / T(nc), exp);
}

template<bool UsePbc>
__device__ __forceinline__ double distSq(double xj, double yj, double zj, double xi, double yi, double zi,
                                         const sphb200_box& box)
{
    double dx = xj - xi;
    double dy = yj - yi;
    double dz = zj - zi;
    if (UsePbc)
    {
        // fold into [-l/2, l/2]; factor pbc*l is l or 0 exactly (findneighbors.hpp:58-64)
        if (box.bc[0] == SPHB200_PERIODIC) dx -= box.len[0] * rint(dx * box.inv[0]);
        if (box.bc[1] == SPHB200_PERIODIC) dy -= box.len[1] * rint(dy * box.inv[1]);
        if (box.bc[2] == SPHB200_PERIODIC) dz -= box.len[2] * rint(dz * box.inv[2]);
    }
    return dx * dx + dy * dy + dz * dz;
}

template<bool UsePbc>
__device__ __forceinline__ void testLeaf(const double* __restrict__ x, const double* __restrict__ y,
                                         const double* __restrict__ z, LocalIndex jb, LocalIndex je, double* cand,
                                         int lane, bool active, LocalIndex i, double xi, double yi, double zi,
                                         double radiusSq, const sphb200_box& box, unsigned ngmax, uint32_t* myList,
                                         unsigned& count)
{
    for (LocalIndex base = jb; base < je; base += 32)
    {
        LocalIndex j = base + lane;
        if (j < je)
        {
            cand[lane]      = x[j];
            cand[32 + lane] = y[j];
            cand[64 + lane] = z[j];
        }
        __syncwarp();
        int cnt = min(32u, je - base);
        if (active)
        {
#pragma unroll 4
            for (int k = 0; k < cnt; ++k)
            {
                double d2     = distSq<UsePbc>(cand[k], cand[32 + k], cand[64 + k], xi, yi, zi, box);
                LocalIndex jj = base + k;
                if (d2 < radiusSq && jj != i)
                {
                    if (count < ngmax) myList[size_t(count) * 32] = jj;
                    count++;
                }
            }
        }
        __syncwarp();
    }
}

constexpr int kFrontier = 256; // nodes per level of the breadth-first walk; larger walks fall back to the DFS

//! @brief extents of the warp's (active) target positions and its largest squared search radius
struct TargetGroup
{
    double lo[3], hi[3];
    double rmaxSq;
};

//! @brief inclusive warp prefix sum
__device__ __forceinline__ int warpInclusiveScan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

/*! @brief level-wise tree walk + pipelined, prefiltered leaf tests (the main search path)
 *
 * 1. The octree is walked breadth-first with one lane per frontier node (8 independent child tests per lane instead
 *    of one dependent node visit per step); children replace their parent in place, so the frontier stays in
 *    ascending SFC order and so do the neighbor lists.
 * 2. One lane per leaf: particle range + cull of the leaf's tight extents against the box of the target positions.
 * 3. The surviving ranges are streamed in 32-particle chunks; the loads of chunk c+1 are in flight while chunk c is
 *    tested.  Each lane first checks ITS candidate against the target box: a candidate farther than the largest
 *    search radius from that box is nobody's neighbor (about 70 % of the staged particles: the leaves cover a box,
 *    the neighbors a rounded box).  Survivors are broadcast and tested against all 32 targets with the reference
 *    arithmetic (distSq).
 * Both culls are conservative in floating point: for lo <= x_i <= hi, |fl(x_j - x_i)| >= fl(max(lo - x_j, x_j - hi, 0))
 * since rounding is monotone, and squaring and the left-associated sum of distSq preserve the ordering, so a computed
 * box distance never exceeds the computed distance to any target inside the box.
 * With periodic folding (UsePbc) the node tests fold like the DFS and the leaf cull of step 2 takes the nearest of the
 * leaf's periodic images per axis.
 * @return false if the frontier overflowed (nothing usable was written; the caller runs the DFS instead)
 */
template<bool UsePbc>
__device__ __forceinline__ bool searchBfs(const DevTree& tree, const double* __restrict__ x,
                                          const double* __restrict__ y, const double* __restrict__ z, int* bufA,
                                          int* bufB, int* bufC, double* cand, int lane, bool active, LocalIndex i,
                                           double xi, double yi, double zi, double radiusSq, const double* tc,
                                           const double* ts, const TargetGroup& tg, const sphb200_box& box,
                                           unsigned ngmax, uint32_t* myList, unsigned& count)
{
    // ---- 1. level-wise walk
    int* cur = bufA;
    int* nxt = bufB;
    int n    = 1;
    if (lane == 0) cur[0] = 0;
    __syncwarp();
    while (true)
    {
        int nOut         = 0;
        bool anyInternal = false;
        for (int b = 0; b < n; b += 32)
        {
            int e      = b + lane;
            bool have  = e < n;
            int node   = have ? cur[e] : 0;
            int child0 = have ? tree.childOffsets[node] : 0;
            unsigned childMask = 0;
            if (have && child0 != 0)
            {
#pragma unroll
                for (int c = 0; c < 8; ++c)
                {
                    const double* ctr = tree.centers + 3 * size_t(child0 + c);
                    const double* sz  = tree.sizes + 3 * size_t(child0 + c);
                    bool ok = true;
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                    {
                        double dx = ctr[d] - tc[d];
                        if (UsePbc && box.bc[d] == SPHB200_PERIODIC) dx -= box.len[d] * rint(dx * box.inv[d]);
                        ok = ok && (fabs(dx) - sz[d] - ts[d] <= 0.0);
                    }
                    childMask |= unsigned(ok) << c;
                }
            }
            int emit    = have ? (child0 == 0 ? 1 : __popc(childMask)) : 0;
            anyInternal = anyInternal || __any_sync(0xffffffffu, have && child0 != 0);
            int incl    = warpInclusiveScan(emit, lane);
            int total   = __shfl_sync(0xffffffffu, incl, 31);
            if (nOut + total > kFrontier) return false;
            int pos = nOut + incl - emit;
            if (have)
            {
                if (child0 == 0) nxt[pos] = node;
                else
                {
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        if (childMask >> c & 1u) nxt[pos++] = child0 + c;
                }
            }
            nOut += total;
        }
        __syncwarp();
        int* t = cur;
        cur    = nxt;
        nxt    = t;
        n      = nOut;
        if (!anyInternal) break;
    }

    // ---- 2. leaves -> particle ranges, culled against the target box; leaves that follow each other in the particle
    // array are merged into one range (SFC neighbors are mostly kept together), so that the 32-particle chunks of
    // step 3 are full instead of one partly filled chunk per leaf
    int* rb = nxt;
    int* re = bufC;
    int nR  = 0;
    int prevEnd = -1; // end of the last range written so far
    for (int b = 0; b < n; b += 32)
    {
        int e     = b + lane;
        bool keep = false;
        LocalIndex jb = 0, je = 0;
        if (e < n)
        {
            TreeNodeIndex leafIdx = tree.internalToLeaf[cur[e]];
            jb = tree.layout[leafIdx];
            je = tree.layout[leafIdx + 1];
            const double* lb = tree.leafBounds + 6 * size_t(leafIdx);
            double ex = fmax(fmax(lb[0] - tg.hi[0], tg.lo[0] - lb[3]), 0.0);
            double ey = fmax(fmax(lb[1] - tg.hi[1], tg.lo[1] - lb[4]), 0.0);
            double ez = fmax(fmax(lb[2] - tg.hi[2], tg.lo[2] - lb[5]), 0.0);
            double rCull = tg.rmaxSq;
            if (UsePbc)
            {
                // periodic axes: the nearest of the leaf's three images, per axis (a lower bound of the folded distance
                // of every particle in the leaf to every target in the box); margin as in the candidate prefilter
                double e[3] = {ex, ey, ez};
#pragma unroll
                for (int d = 0; d < 3; ++d)
                    if (box.bc[d] == SPHB200_PERIODIC)
                    {
                        double l  = box.len[d];
                        double em = fmax(fmax(lb[d] - l - tg.hi[d], tg.lo[d] - (lb[3 + d] - l)), 0.0);
                        double ep = fmax(fmax(lb[d] + l - tg.hi[d], tg.lo[d] - (lb[3 + d] + l)), 0.0);
                        e[d]      = fmin(e[d], fmin(em, ep));
                    }
                ex = e[0], ey = e[1], ez = e[2];
                rCull = tg.rmaxSq * (1.0 + 1e-9);
            }
            keep = je > jb && ex * ex + ey * ey + ez * ez < rCull;
        }
        unsigned m      = __ballot_sync(0xffffffffu, keep);
        unsigned below  = m & ((1u << lane) - 1);
        int prevLane    = below ? 31 - __clz(below) : -1;
        int prevJe      = __shfl_sync(0xffffffffu, int(je), prevLane < 0 ? 0 : prevLane);
        if (prevLane < 0) prevJe = prevEnd;
        bool start      = keep && prevJe != int(jb);
        unsigned starts = __ballot_sync(0xffffffffu, start);
        if (keep)
        {
            int idx = nR + __popc(starts & (0xffffffffu >> (31 - lane))) - 1; // the range this leaf belongs to
            if (start) rb[idx] = int(jb);
            unsigned above = lane == 31 ? 0u : (m & (0xffffffffu << (lane + 1)));
            bool closes    = above == 0u || ((starts >> (__ffs(above) - 1)) & 1u);
            if (closes) re[idx] = int(je);
        }
        nR += __popc(starts);
        if (m) prevEnd = __shfl_sync(0xffffffffu, int(je), 31 - __clz(m));
    }
    __syncwarp();
    if (nR == 0) return true;

    // ---- 3. stream the ranges, one chunk ahead
    // periodic case: candidates are moved to the image nearest to the centre of the target box for the prefilter only
    // (the caller guarantees box half-size + search radius < len/4, so that image is the one every target would
    // pick); the 1e-9 margin covers the different rounding order, membership is decided by distSq<true> as always
    const double tgc[3]  = {0.5 * (tg.lo[0] + tg.hi[0]), 0.5 * (tg.lo[1] + tg.hi[1]), 0.5 * (tg.lo[2] + tg.hi[2])};
    const double rFilter = UsePbc ? tg.rmaxSq * (1.0 + 1e-9) : tg.rmaxSq;
    int* candIdx         = cur; // the frontier is no longer needed
    int li = 0, nBuf = 0;
    LocalIndex base = LocalIndex(rb[0]), end = LocalIndex(re[0]);
    double X = 0, Y = 0, Z = 0;
    bool have = base + lane < end;
    if (have)
    {
        X = x[base + lane];
        Y = y[base + lane];
        Z = z[base + lane];
    }
    while (li < nR)
    {
        // prefilter the chunk whose loads were issued one iteration ago and compact the survivors in shared memory
        bool pass = false;
        if (have)
        {
            double Xs = X, Ys = Y, Zs = Z;
            if (UsePbc)
            {
                if (box.bc[0] == SPHB200_PERIODIC) Xs -= box.len[0] * rint((X - tgc[0]) * box.inv[0]);
                if (box.bc[1] == SPHB200_PERIODIC) Ys -= box.len[1] * rint((Y - tgc[1]) * box.inv[1]);
                if (box.bc[2] == SPHB200_PERIODIC) Zs -= box.len[2] * rint((Z - tgc[2]) * box.inv[2]);
            }
            double ex = fmax(fmax(tg.lo[0] - Xs, Xs - tg.hi[0]), 0.0);
            double ey = fmax(fmax(tg.lo[1] - Ys, Ys - tg.hi[1]), 0.0);
            double ez = fmax(fmax(tg.lo[2] - Zs, Zs - tg.hi[2]), 0.0);
            pass      = ex * ex + ey * ey + ez * ez < rFilter;
        }
        unsigned survivors = __ballot_sync(0xffffffffu, pass);
        if (pass)
        {
            int pos         = nBuf + __popc(survivors & ((1u << lane) - 1));
            cand[pos]       = X;
            cand[64 + pos]  = Y;
            cand[128 + pos] = Z;
            candIdx[pos]    = int(base + lane);
        }
        nBuf += __popc(survivors);

        // issue the next chunk's loads
        base += 32;
        if (base >= end)
        {
            ++li;
            if (li < nR)
            {
                base = LocalIndex(rb[li]);
                end  = LocalIndex(re[li]);
            }
        }
        have = li < nR && base + lane < end;
        if (have)
        {
            X = x[base + lane];
            Y = y[base + lane];
            Z = z[base + lane];
        }
        __syncwarp();

        // survivors are tested in full batches of 32 (about a third of a chunk survives the prefilter)
        if (nBuf >= 32)
        {
            if (active)
            {
#pragma unroll 8
                for (int k = 0; k < 32; ++k)
                {
                    double d2     = distSq<UsePbc>(cand[k], cand[64 + k], cand[128 + k], xi, yi, zi, box);
                    LocalIndex jj = LocalIndex(candIdx[k]);
                    bool in       = d2 < radiusSq && jj != i;
                    if (in && count < ngmax) myList[size_t(count) * 32] = jj;
                    count += in;
                }
            }
            __syncwarp();
            int rem = nBuf - 32;
            double tx = 0, ty = 0, tz = 0;
            int ti = 0;
            if (lane < rem)
            {
                tx = cand[32 + lane];
                ty = cand[96 + lane];
                tz = cand[160 + lane];
                ti = candIdx[32 + lane];
            }
            __syncwarp();
            if (lane < rem)
            {
                cand[lane]       = tx;
                cand[64 + lane]  = ty;
                cand[128 + lane] = tz;
                candIdx[lane]    = ti;
            }
            nBuf = rem;
            __syncwarp();
        }
    }
    if (active)
    {
        for (int k = 0; k < nBuf; ++k)
        {
            double d2     = distSq<UsePbc>(cand[k], cand[64 + k], cand[128 + k], xi, yi, zi, box);
            LocalIndex jj = LocalIndex(candIdx[k]);
            bool in       = d2 < radiusSq && jj != i;
            if (in && count < ngmax) myList[size_t(count) * 32] = jj;
            count += in;
        }
    }
    __syncwarp();
    return true;
}

template<class Th>
__global__ void __launch_bounds__(kNbWarps * 32)
    findNeighborsKernel(DevTree tree, const double* __restrict__ x, const double* __restrict__ y,
                        const double* __restrict__ z, Th* __restrict__ h, LocalIndex first, LocalIndex last,
                        sphb200_box box, unsigned ng0, unsigned ngmax, int iterateH, uint32_t* __restrict__ nidx,
                        uint32_t* __restrict__ nc, unsigned* __restrict__ numFails,
                        const uint8_t* __restrict__ activeMask, double extFactor, uint8_t* __restrict__ gatherMask)
{
    __shared__ TreeNodeIndex stackMem[kNbWarps][kStackSize];
    __shared__ double candMem[kNbWarps][192];
    __shared__ int frontierMem[kNbWarps][3][kFrontier];

    const int lane  = threadIdx.x & 31;
    const int warp  = threadIdx.x >> 5;
    const size_t group = size_t(blockIdx.x) * kNbWarps + warp;
    const size_t gBase = size_t(first) + group * 32;
    if (gBase >= last) return; // whole warp exits together

    TreeNodeIndex* stack = stackMem[warp];
    double* cand         = candMem[warp];

    // GroupView scope (block time steps): only the marked targets are searched, a slab without any is skipped
    const bool inRange = gBase + lane < last;
    const bool valid   = inRange && (!activeMask || activeMask[gBase + lane]);
    if (!__any_sync(0xffffffffu, valid)) return;
    const LocalIndex i = inRange ? LocalIndex(gBase + lane) : last - 1;
    const double xi = x[i], yi = y[i], zi = z[i];
    Th hi           = h[i];
    uint32_t* myList = nidx + nbOffset(group * 32 + lane, 0, ngmax);

    const bool anyPbc =
        box.bc[0] == SPHB200_PERIODIC || box.bc[1] == SPHB200_PERIODIC || box.bc[2] == SPHB200_PERIODIC;
    const unsigned ngmin = ng0 / 4;

    bool active   = valid;
    int iteration = 0;
    unsigned count = 0;

    while (__any_sync(0xffffffffu, active))
    {
        // membership radius in the type of h (findneighbors.hpp:111)
        Th radiusSqT    = Th(4.0) * hi * hi;
        double radiusSq = double(radiusSqT);
        double r        = 2.0 * double(hi);

        // bounding box of the active targets' search spheres, and of their positions (candidate prefilter)
        double lo[3] = {xi - r, yi - r, zi - r};
        double hi3[3] = {xi + r, yi + r, zi + r};
        TargetGroup tg;
        tg.lo[0] = tg.hi[0] = xi;
        tg.lo[1] = tg.hi[1] = yi;
        tg.lo[2] = tg.hi[2] = zi;
        tg.rmaxSq = radiusSq;
        if (!active)
        {
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                lo[d] = tg.lo[d] = INFINITY;
                hi3[d] = tg.hi[d] = -INFINITY;
            }
            tg.rmaxSq = 0.0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                lo[d]    = fmin(lo[d], __shfl_xor_sync(0xffffffffu, lo[d], o));
                hi3[d]   = fmax(hi3[d], __shfl_xor_sync(0xffffffffu, hi3[d], o));
                tg.lo[d] = fmin(tg.lo[d], __shfl_xor_sync(0xffffffffu, tg.lo[d], o));
                tg.hi[d] = fmax(tg.hi[d], __shfl_xor_sync(0xffffffffu, tg.hi[d], o));
            }
            tg.rmaxSq = fmax(tg.rmaxSq, __shfl_xor_sync(0xffffffffu, tg.rmaxSq, o));
        }
        double tc[3], ts[3];
        bool inside = true;
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            tc[d] = 0.5 * (hi3[d] + lo[d]);
            // conservative: pruning only, membership is decided by the exact distance test
            // extFactor: OctreeNsView::searchExtFactor (find_neighbors.cuh:462), > 1 while particles have drifted
            // out of the cells of a tree that was not rebuilt (HydroVeBdtProp::partialSync)
            ts[d]  = 0.5 * (hi3[d] - lo[d]) * (1.0 + 1e-9) * extFactor + 1e-13 * box.len[d];
            inside = inside && (lo[d] >= box.lim[2 * d]) && (hi3[d] <= box.lim[2 * d + 1]);
        }
        // same switch as findneighbors.hpp:116-118 / find_neighbors.cuh:470-472: no fold needed if all search
        // spheres lie inside the box (then rint(dx/l) == 0 for every possible neighbor)
        const bool usePbc = anyPbc && !inside;

        if (active) count = 0;
        bool done = false;
        int *fa = frontierMem[warp][0], *fb = frontierMem[warp][1], *fc = frontierMem[warp][2];
        if (!usePbc)
        {
            done = searchBfs<false>(tree, x, y, z, fa, fb, fc, cand, lane, active, i, xi, yi, zi, radiusSq, tc, ts, tg,
                                    box, ngmax, myList, count);
        }
        else
        {
            // nearest-image prefilter needs the target box plus the search radius well inside half a period
            bool small = true;
#pragma unroll
            for (int d = 0; d < 3; ++d)
                small = small && (box.bc[d] != SPHB200_PERIODIC || ts[d] < 0.25 * box.len[d]);
            if (small)
                done = searchBfs<true>(tree, x, y, z, fa, fb, fc, cand, lane, active, i, xi, yi, zi, radiusSq, tc, ts,
                                       tg, box, ngmax, myList, count);
        }
        if (!done && active) count = 0;
        int sp = 0;
        if (lane == 0) stack[0] = 0;
        sp = done ? 0 : 1;
        __syncwarp();
        while (sp > 0)
        {
            TreeNodeIndex node   = stack[--sp];
            TreeNodeIndex child0 = tree.childOffsets[node];
            __syncwarp();
            if (child0 == 0)
            {
                TreeNodeIndex leafIdx = tree.internalToLeaf[node];
                LocalIndex jb = tree.layout[leafIdx], je = tree.layout[leafIdx + 1];
                if (usePbc) testLeaf<true>(x, y, z, jb, je, cand, lane, active, i, xi, yi, zi, radiusSq, box, ngmax, myList, count);
                else
                {
                    // per-target cull against the tight leaf extents.  Every particle of the leaf has
                    // lo <= x_j <= hi, rounding is monotone and the sum below associates like distSq, so the computed
                    // box distance never exceeds the computed particle distance: a leaf skipped here holds no neighbor.
                    const double* lb = tree.leafBounds + 6 * size_t(leafIdx);
                    double ex = fmax(fmax(lb[0] - xi, xi - lb[3]), 0.0);
                    double ey = fmax(fmax(lb[1] - yi, yi - lb[4]), 0.0);
                    double ez = fmax(fmax(lb[2] - zi, zi - lb[5]), 0.0);
                    bool hit  = active && (ex * ex + ey * ey + ez * ez < radiusSq);
                    if (__any_sync(0xffffffffu, hit))
                        testLeaf<false>(x, y, z, jb, je, cand, lane, hit, i, xi, yi, zi, radiusSq, box, ngmax, myList, count);
                }
            }
            else
            {
                bool ok = false;
                if (lane < 8)
                {
                    TreeNodeIndex c = child0 + lane;
                    ok              = true;
#pragma unroll
                    for (int d = 0; d < 3; ++d)
                    {
                        double dx = tree.centers[3 * size_t(c) + d] - tc[d];
                        if (box.bc[d] == SPHB200_PERIODIC) dx -= box.len[d] * rint(dx * box.inv[d]);
                        ok = ok && (fabs(dx) - tree.sizes[3 * size_t(c) + d] - ts[d] <= 0.0);
                    }
                }
                unsigned m = __ballot_sync(0xffffffffu, ok);
                // push in descending octant order so that children pop in ascending SFC order
                if (ok)
                {
                    int above     = __popc(m >> (lane + 1)); // overlapping children with higher octant
                    stack[sp + above] = child0 + lane;
                }
                sp += __popc(m);
                __syncwarp();
            }
        }

        // coupled h / neighbor-count iteration, per target (sph/find_neighbors.hpp:24-35)
        if (active)
        {
            unsigned ncSph = count + 1;
            bool need      = iterateH && (ngmin > ncSph || (ncSph - 1) > ngmax);
            if (need && iteration++ < kMaxHIter) { hi = updateH(ng0, ncSph, hi); }
            else { active = false; }
        }
    }

    if (valid)
    {
        if (gatherMask)
        {
            // what the pair operators will gather for this target: itself and its list
            gatherMask[i] = 1;
            unsigned stored = min(count, ngmax);
            for (unsigned k = 0; k < stored; ++k)
                gatherMask[myList[size_t(k) * 32]] = 1;
        }
        nc[i - first] = count + 1;
        if (iterateH)
        {
            if (iteration > 0) h[i] = hi;
            if (iteration >= kMaxHIter) atomicAdd(numFails, 1u);
        }
    }
}

//! @brief list buffer <-> row-major [target][ngmax]
template<bool ToRowMajor>
__global__ void convertListsKernel(uint32_t* __restrict__ lists, const uint32_t* __restrict__ nc, uint32_t numTargets,
                                   unsigned ngmax, uint32_t* __restrict__ rowMajor)
{
    size_t t = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (t >= numTargets) return;
    unsigned cnt = min(nc[t] - 1, ngmax);
    for (unsigned k = 0; k < cnt; ++k)
    {
        if (ToRowMajor) rowMajor[t * ngmax + k] = lists[nbOffset(t, k, ngmax)];
        else lists[nbOffset(t, k, ngmax)] = rowMajor[t * ngmax + k];
    }
}

} // namespace sphb200

using namespace sphb200;

extern "C" size_t sphb200_neighbors_size(size_t numTargets, unsigned ngmax)
{
    return ((numTargets + 31) / 32) * size_t(ngmax) * 32;
}

extern "C" int sphb200_find_neighbors(sphb200_ctx* ctx, const sphb200_tree_view* t, const double* x, const double* y,
                                      const double* z, void* h, int dtype, uint32_t first, uint32_t last,
                                      const sphb200_box* box, unsigned ng0, unsigned ngmax, int iterateH,
                                      uint32_t* neighbors, uint32_t* nc, unsigned* numFails_h)
{
    try
    {
        if (numFails_h) *numFails_h = 0;
        if (last <= first) return 0;
        auto* bounds = static_cast<double*>(ctx->nbTmp.get(size_t(t->numLeafNodes) * 6 * sizeof(double)));
        SPHB200_LAUNCH(ctx, leafBoundsKernel, gridFor(t->numLeafNodes, 4), 128, 0, t->layout, t->numLeafNodes, x, y, z,
                       bounds);
        DevTree tree{t->childOffsets, t->internalToLeaf, t->layout, t->centers, t->sizes, bounds};
        unsigned* dFails = reinterpret_cast<unsigned*>(ctx->devScalars + 4);
        SPHB200_CUDA(cudaMemsetAsync(dFails, 0, sizeof(unsigned), ctx->stream));
        size_t numGroups = (size_t(last - first) + 31) / 32;
        int grid         = int((numGroups + kNbWarps - 1) / kNbWarps);
        if (dtype == SPHB200_F64)
            SPHB200_LAUNCH(ctx, findNeighborsKernel<double>, grid, kNbWarps * 32, 0, tree, x, y, z,
                           static_cast<double*>(h), first, last, *box, ng0, ngmax, iterateH, neighbors, nc, dFails,
                           ctx->scopeActive, double(ctx->searchExtFactor), ctx->scopeGather);
        else
            SPHB200_LAUNCH(ctx, findNeighborsKernel<float>, grid, kNbWarps * 32, 0, tree, x, y, z,
                           static_cast<float*>(h), first, last, *box, ng0, ngmax, iterateH, neighbors, nc, dFails,
                           ctx->scopeActive, double(ctx->searchExtFactor), ctx->scopeGather);
        if (numFails_h)
        {
            SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars + 4, dFails, sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
            SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
            *numFails_h = unsigned(ctx->hostScalars[4]);
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_neighbors_to_rowmajor(sphb200_ctx* ctx, const uint32_t* slab, const uint32_t* nc,
                                             uint32_t numTargets, unsigned ngmax, uint32_t* rowmajor)
{
    try
    {
        if (numTargets)
            SPHB200_LAUNCH(ctx, convertListsKernel<true>, gridFor(numTargets, 128), 128, 0,
                           const_cast<uint32_t*>(slab), nc, numTargets, ngmax, rowmajor);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_neighbors_from_rowmajor(sphb200_ctx* ctx, const uint32_t* rowmajor, const uint32_t* nc,
                                               uint32_t numTargets, unsigned ngmax, uint32_t* neighbors)
{
    try
    {
        if (numTargets)
            SPHB200_LAUNCH(ctx, convertListsKernel<false>, gridFor(numTargets, 128), 128, 0, neighbors, nc, numTargets,
                           ngmax, const_cast<uint32_t*>(rowmajor));
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}
