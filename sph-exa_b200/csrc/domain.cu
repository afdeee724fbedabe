// Multi-rank cstone::Domain + HydroVeProp behind the C ABI (sphb200_domain_*): SFC-partitioned run, one sphb200_sim per
// rank, every collective issued from C++ on the compute stream through the transport of comm.cuh (NCCL, or the
// single-GPU loopback over P virtual ranks).
//
// Sequence of a sync (cstone: domain/domain.hpp:196-244, assignment.hpp:84-168, halos/halos.hpp:138-228):
//   1 global box (open axes): local min/max -> all-reduce                                     box_mpi.hpp:84-119
//   2 keys + sort of the owned particles
//   3 replicated global tree: rebalance, local counts, all-reduce SUM, max(local, reduced)     tree/update_mpi.hpp:81-104
//   4 assignment: uniformBins + limitBoundaryShifts (host), send ranges = lower_bound          domaindecomp.hpp:49-231
//   5 migration of the leaving key ranges (only when some rank has any), second sort           domaindecomp_mpi_gpu.cuh
//   6 focus tree (bucketSizeFocus) over the owned keys, max h per leaf, halo boxes             halos.hpp:138-195
//   7 halo discovery.  cstone: requester-side (findHalos on the focus tree, then key ranges are sent to the owners).
//     Here owner-side with the SAME criterion: every rank publishes the integer halo boxes (makeHaloBox) of those of
//     its leaves whose box is not contained in its own key range (containedIn), and each owner walks its own linked
//     octree (findCollisions) and marks the leaves a peer's box overlaps.  A leaf is a halo of rank A iff it overlaps
//     the halo box of one of A's leaves -- the reference's definition, evaluated against the owner's exact
//     bucket-64 leaves instead of the requester's MAC-coarsened image of them.
//   8 layout [halos of lower ranks | owned | halos of higher ranks], final reorder straight into it
//   9 halo exchange of x, y, z, h, m; keys of the halos; bucket-64 tree over all local keys for the neighbor search
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <numeric>
#include "sim.cuh"
#include "comm.cuh"
#include "hilbert.cuh"
#include "primitives.cuh"

using namespace sphb200;

namespace sphb200
{

constexpr int kMaxPeersDomain = 32; // one mask bit per peer

// layout of DistState::dSmall / pinned (uint32 units; the device and the pinned host copy use the same offsets)
constexpr int OFF_META    = 0;    // {firstLeaf, lastLeaf, numBoxes}
constexpr int OFF_BOXCNT  = 16;   // halo boxes per rank [P]
constexpr int OFF_BOUNDS  = 64;   // uint64 [P + 1]
constexpr int OFF_SPLITS  = 136;  // [P + 1] padded to 34 | send counts [32] | matrix [P * P]: contiguous
constexpr int OFF_SEND    = OFF_SPLITS + kMaxPeersDomain + 2;
constexpr int OFF_MATRIX  = OFF_SEND + kMaxPeersDomain;
constexpr int OFF_HALOCNT = 1232; // elements sent to each peer [P], then the number of ranges
constexpr int OFF_HALOMAT = 1272; // [P * (P + 1)]
constexpr int OFF_TOTALS  = 2400; // {numRanges, numElements}
constexpr int OFF_RECVSCAN = 2408, OFF_RECVBEGIN = 2448;
constexpr int OFF_DBL     = 2496; // doubles: bounding box [6]
static_assert(OFF_HALOMAT + kMaxPeersDomain * (kMaxPeersDomain + 1) <= OFF_TOTALS, "small-buffer layout");
static_assert(OFF_DBL % 2 == 0 && OFF_BOUNDS % 2 == 0, "8-byte alignment");

// ---------------------------------------------------------------------------------------------- host: assignment
//! @brief uniformBins (domaindecomp.hpp:49-72)
static void uniformBins(const uint32_t* counts, int numLeaves, int numBins, std::vector<int>& bins,
                        std::vector<uint32_t>& binCounts)
{
    std::vector<uint64_t> scan(size_t(numLeaves) + 1, 0);
    for (int i = 0; i < numLeaves; ++i)
        scan[i + 1] = scan[i] + counts[i];
    double binCount = double(scan.back()) / numBins;
    bins.assign(numBins + 1, 0);
    binCounts.assign(numBins, 0);
    bins[numBins] = numLeaves;
    for (int i = 1; i < numBins; ++i)
    {
        uint64_t target = uint64_t(i * binCount); // in double, then truncated (domaindecomp.hpp:63)
        bins[i]         = int(std::lower_bound(scan.begin(), scan.end(), target) - scan.begin());
    }
    for (int i = 1; i < numBins; ++i)
        binCounts[i - 1] = uint32_t(scan[bins[i]] - scan[bins[i - 1]]);
    binCounts[numBins - 1] = uint32_t(scan.back() - scan[bins[numBins - 1]]);
}

/*! @brief makeSfcAssignment + limitBoundaryShifts (domaindecomp.hpp:115-181)
 *
 * @param old  boundaries of the previous sync (numRanks + 1) or nullptr on the first call
 */
static void sfcAssignment(int numRanks, const uint32_t* counts, const uint64_t* leaves, int numLeaves,
                          const uint64_t* old, uint64_t* bounds, uint32_t* rankCounts)
{
    std::vector<int> bins;
    std::vector<uint32_t> binCounts;
    uniformBins(counts, numLeaves, numRanks, bins, binCounts);
    for (int r = 0; r <= numRanks; ++r)
        bounds[r] = leaves[bins[r]];
    for (int r = 0; r < numRanks; ++r)
        rankCounts[r] = binCounts[r];
    if (!old) return;
    bool recount = false;
    for (int r = 1; r < numRanks; ++r)
    {
        uint64_t nb = std::min(std::max(bounds[r], old[r - 1]), old[r + 1]);
        if (nb != bounds[r])
        {
            recount   = true;
            bounds[r] = nb;
        }
    }
    if (!recount) return;
    for (int r = 0; r < numRanks; ++r)
    {
        // findNodeAbove: first leaf key >= boundary
        auto a = std::lower_bound(leaves, leaves + numLeaves + 1, bounds[r]) - leaves;
        auto b = std::lower_bound(leaves, leaves + numLeaves + 1, bounds[r + 1]) - leaves;
        uint64_t c = 0;
        for (auto i = a; i < b; ++i)
            c += counts[i];
        rankCounts[r] = uint32_t(c);
    }
}

//! @brief the initial global tree: computeSpanningTree(initialDomainSplits(P, log8ceil(100 P))) (assignment.hpp:62-69)
static std::vector<uint64_t> initialGlobalTree(int numRanks)
{
    auto log8ceil = [](uint64_t n) -> unsigned
    {
        if (n == 0) return 0;
        unsigned lz = clz64(n - 1);
        return kMaxTreeLevel - (lz - 1) / 3;
    };
    unsigned level = log8ceil(uint64_t(100) * numRanks);
    std::vector<uint64_t> splits(numRanks + 1);
    uint64_t delta = nodeRange(0) / numRanks;
    splits[0]      = 0;
    for (int i = 1; i < numRanks; ++i)
        splits[i] = (uint64_t(i) * delta) & ~(nodeRange(level) - 1); // enclosingBoxCode
    splits[numRanks] = nodeRange(0);
    std::vector<uint64_t> tree;
    std::vector<uint64_t> cells(256);
    for (int i = 0; i < numRanks; ++i)
    {
        if (splits[i] == splits[i + 1]) continue;
        int n = sphb200_span_sfc_range(splits[i], splits[i + 1], cells.data(), cells.size());
        if (n < 0) throw std::runtime_error("initialGlobalTree: span_sfc_range failed");
        tree.insert(tree.end(), cells.begin(), cells.begin() + n);
    }
    tree.push_back(nodeRange(0));
    return tree;
}

// ---------------------------------------------------------------------------------------------- device kernels
__global__ void maxCountsKernel(const uint32_t* __restrict__ local, uint32_t* __restrict__ reduced, int n)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) reduced[i] = max(local[i], reduced[i]); // update_mpi.hpp:99-101
}

//! @brief createSendRangesGpu (domaindecomp_gpu.cuh:55-70): splits[r] = lower_bound(keys, bounds[r]), splits[P] = n
__global__ void sendSplitsKernel(const KeyType* __restrict__ keys, uint32_t n, const KeyType* __restrict__ bounds,
                                 int numRanks, uint32_t* __restrict__ splits, uint32_t* __restrict__ sendCounts)
{
    __shared__ uint32_t sh[kMaxPeersDomain + 1];
    int r = threadIdx.x;
    if (r < numRanks) sh[r] = uint32_t(lowerBound(keys, size_t(n), bounds[r]));
    if (r == numRanks) sh[r] = n;
    __syncthreads();
    if (r <= numRanks) splits[r] = sh[r];
    if (r < numRanks) sendCounts[r] = sh[r + 1] - sh[r];
}

struct IBox
{
    int lo[3], hi[3];
};

//! @brief hilbertIBox (hilbert.hpp:264-281)
__device__ __forceinline__ IBox nodeIBox(KeyType start, unsigned level)
{
    unsigned cube = (1u << kMaxTreeLevel) >> level;
    unsigned mask = ~(cube - 1);
    unsigned ix, iy, iz;
    hilbertDecode(start, ix, iy, iz);
    IBox b;
    b.lo[0] = int(ix & mask);
    b.lo[1] = int(iy & mask);
    b.lo[2] = int(iz & mask);
    for (int d = 0; d < 3; ++d)
        b.hi[d] = b.lo[d] + int(cube);
    return b;
}

//! @brief toNBitIntCeil<uint64_t> (sfc/common.hpp:80-90)
__device__ __forceinline__ int toNBitIntCeil(double x)
{
    constexpr unsigned top = (1u << kMaxTreeLevel) - 1u;
    unsigned r             = unsigned(ceil(__dmul_rn(x, double(1u << kMaxTreeLevel))));
    return int(min(r, top));
}

//! @brief overlapRange<2^21> (boxoverlap.hpp:56-71)
__device__ __forceinline__ bool overlapRangePbc(int a, int b, int c, int d)
{
    constexpr int R = 1 << kMaxTreeLevel;
    auto two        = [](int a, int b, int c, int d) { return b > c && d > a; };
    return two(a, b, c, d) || two(a + R, b + R, c, d) || two(a, b, c + R, d + R);
}

__device__ __forceinline__ bool overlapIBox(const IBox& a, const IBox& b)
{
    return overlapRangePbc(a.lo[0], a.hi[0], b.lo[0], b.hi[0]) && overlapRangePbc(a.lo[1], a.hi[1], b.lo[1], b.hi[1]) &&
           overlapRangePbc(a.lo[2], a.hi[2], b.lo[2], b.hi[2]);
}

//! @brief containedIn(codeStart, codeEnd, box) (boxoverlap.hpp:94-117)
__device__ __forceinline__ bool boxContainedIn(KeyType codeStart, KeyType codeEnd, const IBox& b)
{
    constexpr int pbcRange = 1 << kMaxTreeLevel;
    if (min(min(b.lo[0], b.lo[1]), b.lo[2]) < 0 || max(max(b.hi[0], b.hi[1]), b.hi[2]) > pbcRange)
        return codeStart == 0 && codeEnd == nodeRange(0);
    KeyType lowCode  = hilbertKey(unsigned(b.lo[0]), unsigned(b.lo[1]), unsigned(b.lo[2]));
    KeyType highCode = hilbertKey(unsigned(b.hi[0] - 1), unsigned(b.hi[1] - 1), unsigned(b.hi[2] - 1));
    // smallestCommonBox (sfc/common.hpp:283-289)
    unsigned commonLevel = unsigned(commonPrefix(lowCode, highCode)) / 3;
    KeyType nodeStart    = lowCode & ~(nodeRange(commonLevel) - 1);
    KeyType nodeEnd      = nodeStart + nodeRange(commonLevel);
    return nodeStart >= codeStart && nodeEnd <= codeEnd;
}

/*! @brief halo radius per leaf as Halos::discover computes it on the host (halos.hpp:180-188):
 *         float(max h of the leaf * 2 * searchExt), 0 for empty leaves
 */
template<class T>
__global__ void haloRadiiKernel(const T* __restrict__ h, const uint32_t* __restrict__ layout, int firstLeaf, int lastLeaf,
                                float searchExt, float* __restrict__ radii)
{
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    int leaf = firstLeaf + warp;
    if (leaf >= lastLeaf) return;
    uint32_t a = layout[leaf], b = layout[leaf + 1];
    T m        = 0;
    for (uint32_t i = a + lane; i < b; i += 32)
        m = max(m, h[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    // `*max_element(h) * 2 * searchExtFact` is evaluated in the type of h x float, then stored as float
    if (lane == 0) radii[leaf] = (b > a) ? float(m * 2 * searchExt) : 0.0f;
}

/*! @brief halo boxes of the leaves [firstLeaf, lastLeaf) that reach outside the key range they span
 *         (findHalos, traversal/collisions.hpp:91-103: makeHaloBox, then the containedIn early exit)
 *
 * boxes[k] = 6 ints (xmin, xmax, ymin, ymax, zmin, zmax as in IBox) + the leaf index; *numBoxes counts them
 */
__global__ void haloBoxesKernel(const KeyType* __restrict__ leaves, const float* __restrict__ radii, int firstLeaf,
                                int lastLeaf, sphb200_box box, int* __restrict__ boxes, int capacity,
                                int* __restrict__ numBoxes, const uint32_t* __restrict__ counts, int* __restrict__ extent)
{
    int leaf   = firstLeaf + blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = leaf < lastLeaf;
    IBox nb;
    if (valid)
    {
        KeyType a = leaves[leaf], b = leaves[leaf + 1];
        nb = nodeIBox(a, treeLevel(b - a));
    }
    if (extent)
    {
        // integer bounding box of the non-empty leaves of the range (extent = {xlo, xhi, ylo, yhi, zlo, zhi}, preset to
        // {INT_MAX, INT_MIN, ...}): a halo box of a peer that misses it needs no walk through this rank's tree
        bool mine = valid && counts[leaf] > 0;
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            int lo = __reduce_min_sync(0xffffffffu, mine ? nb.lo[d] : INT_MAX);
            int hi = __reduce_max_sync(0xffffffffu, mine ? nb.hi[d] : INT_MIN);
            if ((threadIdx.x & 31) == 0 && lo <= hi)
            {
                atomicMin(extent + 2 * d, lo);
                atomicMax(extent + 2 * d + 1, hi);
            }
        }
    }
    if (!valid) return;
    float radius = radii[leaf];
    IBox hb;
    constexpr int maxCoord = 1 << kMaxTreeLevel;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        int delta = toNBitIntCeil(__dmul_rn(double(radius), box.inv[d]));
        bool pbc  = box.bc[d] == SPHB200_PERIODIC;
        int lo = nb.lo[d] - delta, hi = nb.hi[d] + delta;
        hb.lo[d] = pbc ? lo : min(max(0, lo), maxCoord);
        hb.hi[d] = pbc ? hi : min(max(0, hi), maxCoord);
    }
    if (boxContainedIn(leaves[firstLeaf], leaves[lastLeaf], hb)) return;
    int k = atomicAdd(numBoxes, 1);
    if (k >= capacity) return;
    int* o = boxes + size_t(k) * 8;
    o[0] = hb.lo[0]; o[1] = hb.hi[0]; o[2] = hb.lo[1]; o[3] = hb.hi[1]; o[4] = hb.lo[2]; o[5] = hb.hi[2];
    o[6] = leaf; o[7] = 0;
}

/*! @brief findCollisions (traversal/collisions.hpp:39-57) of every box of one sender against a linked octree
 *
 * One thread per box, depth-first with a local stack.  A node is entered when its integer box overlaps the halo box
 * (periodic-aware) and it is not contained in [exLo, exHi); only nodes that intersect [visitLo, visitHi) are walked
 * (the owner's key range: everything else in its tree is empty).  Leaves that are reached get `bit` or-ed into their
 * mask (flags[leaf] = 1 when mask is NULL); counts != NULL: empty leaves are not marked.
 */
constexpr int kMarkWarps = 4;    // boxes per CTA
constexpr int kMarkStack = 1280; // node stack per warp (shared memory)
constexpr int kMarkWide  = 1024; // below this fill four nodes are expanded per step, above it one (plain depth-first)

/*! @brief one warp walks one halo box through the linked octree: up to four nodes are popped per step and their 32
 *         children tested by the 32 lanes.  A thread per box with a private stack walked a box that overlaps many
 *         leaves one dependent load after the other, and a launch lasts as long as its longest walk.
 */
__device__ __forceinline__ void markBox(const KeyType* __restrict__ prefixes,
                                        const TreeNodeIndex* __restrict__ childOffsets,
                                        const TreeNodeIndex* __restrict__ internalToLeaf,
                                        const uint32_t* __restrict__ counts, const int* __restrict__ bx, uint32_t bit,
                                        KeyType exLo, KeyType exHi, KeyType visitLo, KeyType visitHi,
                                        uint32_t* __restrict__ mask, int* __restrict__ flags,
                                        TreeNodeIndex* __restrict__ stack)
{
    const int lane = threadIdx.x & 31;
    IBox target;
    target.lo[0] = bx[0]; target.hi[0] = bx[1]; target.lo[1] = bx[2]; target.hi[1] = bx[3];
    target.lo[2] = bx[4]; target.hi[2] = bx[5];

    auto enters = [&](TreeNodeIndex idx)
    {
        KeyType prefix  = prefixes[idx];
        KeyType nodeKey = decodePlaceholderBit(prefix);
        unsigned level  = decodePrefixLength(prefix) / 3;
        KeyType nodeEnd = nodeKey + nodeRange(level);
        if (nodeEnd <= visitLo || nodeKey >= visitHi) return false;
        if (!(nodeKey < exLo || nodeEnd > exHi)) return false; // containedIn(node, exclude range)
        return overlapIBox(nodeIBox(nodeKey, level), target);
    };

    if (!enters(0)) return; // same answer in every lane
    if (lane == 0) stack[0] = 0;
    int top = 1;
    __syncwarp();
    while (top > 0)
    {
        int take = top > kMarkWide ? 1 : min(top, 4);
        int slot = lane >> 3;
        TreeNodeIndex node = slot < take ? stack[top - 1 - slot] : -1;
        __syncwarp();
        top -= take;
        bool push       = false;
        TreeNodeIndex c = 0;
        if (node >= 0)
        {
            TreeNodeIndex child = childOffsets[node];
            if (child == 0)
            {
                if ((lane & 7) == 0)
                {
                    TreeNodeIndex leaf = internalToLeaf[node];
                    if (!(counts && counts[leaf] == 0))
                    {
                        if (mask) atomicOr(mask + leaf, bit);
                        else flags[leaf] = 1;
                    }
                }
            }
            else
            {
                c    = child + (lane & 7);
                push = enters(c);
            }
        }
        unsigned m = __ballot_sync(0xffffffffu, push);
        if (push) stack[top + __popc(m & ((1u << lane) - 1))] = c;
        top += __popc(m);
        __syncwarp();
    }
}

__global__ void __launch_bounds__(kMarkWarps * 32)
    markCollisionsKernel(const KeyType* __restrict__ prefixes, const TreeNodeIndex* __restrict__ childOffsets,
                         const TreeNodeIndex* __restrict__ internalToLeaf, const uint32_t* __restrict__ counts,
                         const int* __restrict__ boxes, int numBoxes, uint32_t bit, KeyType exLo, KeyType exHi,
                         KeyType visitLo, KeyType visitHi, uint32_t* __restrict__ mask, int* __restrict__ flags)
{
    __shared__ TreeNodeIndex stacks[kMarkWarps][kMarkStack];
    int k = blockIdx.x * kMarkWarps + (threadIdx.x >> 5);
    if (k >= numBoxes) return;
    markBox(prefixes, childOffsets, internalToLeaf, counts, boxes + size_t(k) * 8, bit, exLo, exHi, visitLo, visitHi,
            mask, flags, stacks[threadIdx.x >> 5]);
}

/*! @brief the boxes of all peers in one launch: block q of the gathered array (stride boxes, boxCount[q] of them valid)
 *         marks bit q (one launch per sender cost P - 1 times the longest walk)
 */
__global__ void __launch_bounds__(kMarkWarps * 32)
    markCollisionsAllKernel(const KeyType* __restrict__ prefixes, const TreeNodeIndex* __restrict__ childOffsets,
                            const TreeNodeIndex* __restrict__ internalToLeaf, const uint32_t* __restrict__ counts,
                            const int* __restrict__ boxes, int stride, int numRanks, int myRank,
                            const uint32_t* __restrict__ boxCount, KeyType visitLo, KeyType visitHi,
                            uint32_t* __restrict__ mask, const int* __restrict__ extent)
{
    __shared__ TreeNodeIndex stacks[kMarkWarps][kMarkStack];
    size_t k = size_t(blockIdx.x) * kMarkWarps + (threadIdx.x >> 5);
    if (k >= size_t(stride) * numRanks) return;
    int q = int(k / stride), b = int(k % stride);
    if (q == myRank || b >= int(boxCount[q])) return;
    {
        // most boxes of most peers are nowhere near this rank: test against the extent of its non-empty leaves first
        const int* bx = boxes + k * 8;
        IBox target, mine;
        for (int d = 0; d < 3; ++d)
        {
            target.lo[d] = bx[2 * d];
            target.hi[d] = bx[2 * d + 1];
            mine.lo[d]   = extent[2 * d];
            mine.hi[d]   = extent[2 * d + 1];
        }
        if (mine.lo[0] > mine.hi[0] || !overlapIBox(mine, target)) return;
    }
    markBox(prefixes, childOffsets, internalToLeaf, counts, boxes + k * 8, 1u << q, KeyType(0), KeyType(0), visitLo,
            visitHi, mask, nullptr, stacks[threadIdx.x >> 5]);
}

//! @brief per (peer, leaf): does the leaf go to that peer, and with how many particles
__global__ void sendFlagsKernel(const uint32_t* __restrict__ mask, const uint32_t* __restrict__ counts, int numLeaves,
                                int numRanks, int myRank, uint32_t* __restrict__ flags, uint32_t* __restrict__ lens)
{
    size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= size_t(numLeaves) * numRanks) return;
    int q    = int(idx / numLeaves);
    int leaf = int(idx % numLeaves);
    bool on  = q != myRank && ((mask[leaf] >> q) & 1u) && counts[leaf] > 0;
    flags[idx] = on;
    lens[idx]  = on ? counts[leaf] : 0u;
}

__global__ void sendRangesKernel(const uint32_t* __restrict__ flagsScan, const uint32_t* __restrict__ lensScan,
                                 const uint32_t* __restrict__ mask, const uint32_t* __restrict__ counts,
                                 const uint32_t* __restrict__ layout, int numLeaves, int numRanks, int myRank,
                                 uint32_t* __restrict__ rangeScan, uint32_t* __restrict__ rangeBegin)
{
    size_t idx = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= size_t(numLeaves) * numRanks) return;
    int q    = int(idx / numLeaves);
    int leaf = int(idx % numLeaves);
    bool on  = q != myRank && ((mask[leaf] >> q) & 1u) && counts[leaf] > 0;
    if (!on) return;
    uint32_t r    = flagsScan[idx];
    rangeScan[r]  = lensScan[idx];
    rangeBegin[r] = layout[leaf]; // relative to the first owned particle; shifted by addOffsetKernel
}

//! @brief out[q] = elements sent to peer q, out[P] = number of ranges
__global__ void peerTotalsKernel(const uint32_t* __restrict__ flagsScan, const uint32_t* __restrict__ lensScan,
                                 const uint32_t* __restrict__ totals /* {numRanges, numElements} */, int numLeaves,
                                 int numRanks, uint32_t* __restrict__ out)
{
    int q = threadIdx.x;
    if (q < numRanks)
    {
        uint32_t a = lensScan[size_t(q) * numLeaves];
        uint32_t b = (q + 1 < numRanks) ? lensScan[size_t(q + 1) * numLeaves] : totals[1];
        out[q]     = b - a;
    }
    if (q == numRanks) out[q] = totals[0];
}

__global__ void addOffsetKernel(uint32_t* __restrict__ a, int n, uint32_t off)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] += off;
}

} // namespace sphb200

// ---------------------------------------------------------------------------------------------- host helpers
namespace
{

struct Group
{
    std::vector<sphb200_sim*> sims;
    std::vector<RankStream> rs;
    sphb200_comm* comm;
    int P;
    int numLocal() const { return int(sims.size()); }
};

Group makeGroup(sphb200_sim* const* sims, int numLocal)
{
    if (numLocal < 1 || !sims) throw std::runtime_error("domain: no simulation handles");
    Group g;
    for (int l = 0; l < numLocal; ++l)
    {
        sphb200_sim* s = sims[l];
        if (!s || !s->dist) throw std::runtime_error("domain: handle was not created by sphb200_domain_create");
        g.sims.push_back(s);
        g.rs.push_back({s->dist->rank, s->ctx->stream, s->ctx});
    }
    g.comm = g.sims[0]->dist->comm;
    g.P    = g.comm->numRanks;
    for (auto* s : g.sims)
        if (s->dist->comm != g.comm) throw std::runtime_error("domain: handles of different communicators");
    if (g.comm->loopback)
    {
        if (numLocal != g.P) throw std::runtime_error("domain: the loopback transport drives all ranks at once");
        for (int l = 0; l < numLocal; ++l)
        {
            if (g.sims[l]->dist->rank != l) throw std::runtime_error("domain: pass the handles in rank order");
            // scratch arenas, gather records and the time-step scalars live in the context
            for (int k = 0; k < l; ++k)
                if (g.sims[k]->ctx == g.sims[l]->ctx)
                    throw std::runtime_error("domain: every virtual rank needs its own sphb200_ctx (on the same stream)");
        }
    }
    else if (numLocal != 1) throw std::runtime_error("domain: one handle per process with the NCCL transport");
    return g;
}

void syncStream(sphb200_sim* s) { SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream)); }

void updateTreeOnce(sphb200_sim* s, Tree& t, unsigned bucket, const uint64_t* keys, size_t n)
{
    if (t.first)
    {
        int it = 0;
        CHECK_RC(sphb200_compute_octree(s->ctx, keys, n, bucket, t.leaves, t.counts, t.capacity, &t.numLeaves, &it));
        t.first = false;
    }
    else
    {
        int conv = 0;
        CHECK_RC(sphb200_update_octree(s->ctx, keys, n, bucket, t.leaves, t.counts, t.capacity, &t.numLeaves, &conv));
    }
}

//! @brief reorder conserved fields: alt[dstOff + i] = field[srcOff + perm[i]], i < n
void gatherInto(sphb200_sim* s, const int* ids, int count, size_t n, size_t srcOff, size_t dstOff, bool swapAfter)
{
    const void* src[kNumConserved];
    void* dst[kNumConserved];
    int eb[kNumConserved];
    for (int k = 0; k < count; ++k)
    {
        int id = ids[k];
        eb[k]  = s->elemBytes(id);
        src[k] = static_cast<char*>(s->field[id]) + srcOff * eb[k];
        dst[k] = static_cast<char*>(s->fieldAlt[id]) + dstOff * eb[k];
    }
    CHECK_RC(sphb200_gather(s->ctx, s->perm, n, count, src, dst, eb));
    if (swapAfter)
        for (int k = 0; k < count; ++k)
            std::swap(s->field[ids[k]], s->fieldAlt[ids[k]]);
}

const int kAllConserved[kNumConserved] = {F_x, F_y, F_z, F_h, F_m, F_vx, F_vy, F_vz, F_x_m1, F_y_m1, F_z_m1,
                                          F_temp, F_du_m1, F_alpha};

// ---- 1. global box
void syncBox(Group& g)
{
    sphb200_sim* s0 = g.sims[0];
    bool anyOpen    = false;
    for (int d = 0; d < 3; ++d)
        anyOpen = anyOpen || s0->box.bc[d] != SPHB200_PERIODIC;
    if (!anyOpen) return;
    std::vector<void*> ptr;
    for (auto* s : g.sims)
    {
        size_t n = s->end - s->start;
        double mm[6] = {1e300, -1e300, 1e300, -1e300, 1e300, -1e300};
        if (n)
            CHECK_RC(sphb200_minmax3(s->ctx, static_cast<double*>(s->field[F_x]) + s->start,
                                     static_cast<double*>(s->field[F_y]) + s->start,
                                     static_cast<double*>(s->field[F_z]) + s->start, n, mm));
        auto* hp = reinterpret_cast<double*>(s->dist->pinned + OFF_DBL);
        for (int d = 0; d < 3; ++d)
        {
            hp[2 * d]     = mm[2 * d];
            hp[2 * d + 1] = -mm[2 * d + 1]; // one MIN all-reduce for both ends
        }
        SPHB200_CUDA(cudaMemcpyAsync(s->dist->dSmall + OFF_DBL, hp, 6 * sizeof(double), cudaMemcpyHostToDevice,
                                     s->ctx->stream));
        ptr.push_back(s->dist->dSmall + OFF_DBL);
    }
    commAllReduce(g.comm, g.rs, ptr.data(), 6, COMM_F64, COMM_MIN);
    for (auto* s : g.sims)
    {
        auto* hp = reinterpret_cast<double*>(s->dist->pinned + OFF_DBL);
        SPHB200_CUDA(cudaMemcpyAsync(hp, s->dist->dSmall + OFF_DBL, 6 * sizeof(double), cudaMemcpyDeviceToHost,
                                     s->ctx->stream));
    }
    for (auto* s : g.sims)
    {
        syncStream(s);
        auto* hp = reinterpret_cast<double*>(s->dist->pinned + OFF_DBL);
        double lim[6];
        for (int d = 0; d < 3; ++d)
        {
            bool pbc       = s->box.bc[d] == SPHB200_PERIODIC;
            lim[2 * d]     = pbc ? s->box.lim[2 * d] : hp[2 * d];
            lim[2 * d + 1] = pbc ? s->box.lim[2 * d + 1] : -hp[2 * d + 1];
        }
        sphb200_make_box(lim, s->box.bc, &s->box);
    }
}

// ---- 2. keys + sort of the owned particles at [srcOff, srcOff + n) into dist->keysTmp[0, n), ordering in s->perm
void keysAndSort(sphb200_sim* s, size_t srcOff, size_t n)
{
    auto* x = static_cast<double*>(s->field[F_x]) + srcOff;
    auto* y = static_cast<double*>(s->field[F_y]) + srcOff;
    auto* z = static_cast<double*>(s->field[F_z]) + srcOff;
    CHECK_RC(sphb200_sfc_keys(s->ctx, x, y, z, n, &s->box, s->dist->keysTmp));
    CHECK_RC(sphb200_sort_by_key(s->ctx, s->dist->keysTmp, s->perm, n));
}

// ---- 3. updateOctreeGlobal (tree/update_mpi.hpp:81-104): rebalance with the reduced counts of the previous call,
//         local counts, all-reduce, max(local, reduced).  Returns the converged flag (identical on all ranks).
bool updateGlobalTree(Group& g)
{
    bool converged = true;
    std::vector<void*> ptr;
    for (auto* s : g.sims)
    {
        Tree& t  = s->global;
        int conv = 0;
        CHECK_RC(sphb200_update_octree(s->ctx, s->dist->keysTmp, s->dist->numOwn, s->cfg.bucketSize, t.leaves, t.counts,
                                       t.capacity, &t.numLeaves, &conv));
        converged = converged && conv;
        // keep the local counts for the overflow guard
        auto* local = static_cast<uint32_t*>(s->dist->scanArena.get(size_t(t.numLeaves) * sizeof(uint32_t)));
        SPHB200_CUDA(cudaMemcpyAsync(local, t.counts, size_t(t.numLeaves) * sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                                     s->ctx->stream));
        ptr.push_back(t.counts);
    }
    int L = g.sims[0]->global.numLeaves;
    for (auto* s : g.sims)
        if (s->global.numLeaves != L) throw std::runtime_error("domain: the replicated global trees diverged");
    commAllReduce(g.comm, g.rs, ptr.data(), size_t(L), COMM_U32, COMM_SUM);
    for (auto* s : g.sims)
        SPHB200_LAUNCH(s->ctx, maxCountsKernel, gridFor(L, 256), 256, 0,
                       static_cast<const uint32_t*>(s->dist->scanArena.ptr), s->global.counts, L);
    return converged;
}

void syncGlobalTree(Group& g)
{
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        if (!d->globalFirst) continue;
        // GlobalAssignment constructor (assignment.hpp:62-69)
        std::vector<uint64_t> init = initialGlobalTree(d->P);
        int L                      = int(init.size()) - 1;
        if (L > s->global.capacity) throw std::runtime_error("domain: global tree capacity");
        std::vector<uint32_t> cnt(L, s->cfg.bucketSize - 1);
        SPHB200_CUDA(cudaMemcpyAsync(s->global.leaves, init.data(), init.size() * sizeof(uint64_t), cudaMemcpyHostToDevice,
                                     s->ctx->stream));
        SPHB200_CUDA(cudaMemcpyAsync(s->global.counts, cnt.data(), cnt.size() * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                     s->ctx->stream));
        syncStream(s);
        s->global.numLeaves = L;
        s->global.first     = false;
    }
    bool first = g.sims[0]->dist->globalFirst;
    updateGlobalTree(g);
    if (first)
    {
        int guard = 0;
        while (!updateGlobalTree(g))
            if (++guard > 64) throw std::runtime_error("domain: global tree does not converge");
        for (auto* s : g.sims)
            s->dist->globalFirst = false;
    }
}

// ---- 4. assignment + send ranges; returns the P x P matrix of particle counts (row = sender) on the host
std::vector<uint32_t> assign(Group& g)
{
    const int P = g.P;
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        int L        = s->global.numLeaves;
        d->gLeavesH.resize(size_t(L) + 1);
        d->gCountsH.resize(L);
        SPHB200_CUDA(cudaMemcpyAsync(d->gLeavesH.data(), s->global.leaves, (size_t(L) + 1) * sizeof(uint64_t),
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
        SPHB200_CUDA(cudaMemcpyAsync(d->gCountsH.data(), s->global.counts, size_t(L) * sizeof(uint32_t),
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
    }
    std::vector<const void*> sendPtr;
    std::vector<void*> recvPtr;
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        syncStream(s);
        int L = s->global.numLeaves;
        std::vector<uint64_t> nb(P + 1);
        std::vector<uint32_t> rc(P);
        sfcAssignment(P, d->gCountsH.data(), d->gLeavesH.data(), L, d->bounds.empty() ? nullptr : d->bounds.data(),
                      nb.data(), rc.data());
        d->oldBounds  = d->bounds;
        d->bounds     = nb;
        d->rankCounts = rc;
        // device: boundaries -> splits of the sorted owned keys -> send counts
        auto* hb = reinterpret_cast<uint64_t*>(d->pinned + OFF_BOUNDS);
        for (int r = 0; r <= P; ++r)
            hb[r] = nb[r];
        auto* dBounds     = reinterpret_cast<uint64_t*>(d->dSmall + OFF_BOUNDS);
        uint32_t* dSplits = d->dSmall + OFF_SPLITS;
        uint32_t* dSend   = d->dSmall + OFF_SEND;
        uint32_t* dMatrix = d->dSmall + OFF_MATRIX;
        SPHB200_CUDA(cudaMemcpyAsync(dBounds, hb, (P + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, s->ctx->stream));
        SPHB200_LAUNCH(s->ctx, sendSplitsKernel, 1, kMaxPeersDomain + 1, 0, d->keysTmp, d->numOwn, dBounds, P, dSplits,
                       dSend);
        sendPtr.push_back(dSend);
        recvPtr.push_back(dMatrix);
    }
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), size_t(P) * sizeof(uint32_t));
    std::vector<uint32_t> matrix;
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        // splits | send | matrix are contiguous: one copy
        size_t ints = size_t(OFF_MATRIX - OFF_SPLITS) + size_t(P) * P;
        SPHB200_CUDA(cudaMemcpyAsync(d->pinned + OFF_SPLITS, d->dSmall + OFF_SPLITS, ints * sizeof(uint32_t),
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
    }
    for (auto* s : g.sims)
        syncStream(s);
    {
        DistState* d = g.sims[0]->dist;
        const uint32_t* m = d->pinned + OFF_MATRIX;
        matrix.assign(m, m + size_t(P) * P);
    }
    return matrix;
}

const uint32_t* hostSplits(const sphb200_sim* s) { return s->dist->pinned + OFF_SPLITS; }

// ---- 5. exchangeParticles (domaindecomp_mpi_gpu.cuh:85-186) on key-sorted arrays: contiguous index ranges per
//         destination, sent straight out of the field arrays and received straight into the spare ones
void migrate(Group& g, const std::vector<uint32_t>& matrix)
{
    const int P = g.P, NL = g.numLocal();
    // sort every field first (the leaving ranges are contiguous in key order)
    for (auto* s : g.sims)
    {
        gatherInto(s, kAllConserved, kNumConserved, s->dist->numOwn, s->start, 0, true);
        s->start = 0;
        s->end   = s->dist->numOwn;
    }
    std::vector<size_t> sendFirst(size_t(NL) * P, 0), sendCount(size_t(NL) * P, 0), recvFirst(size_t(NL) * P, 0),
        recvCount(size_t(NL) * P, 0);
    std::vector<uint32_t> newOwn(NL);
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        DistState* d   = s->dist;
        int me         = d->rank;
        const uint32_t* splits = hostSplits(s);
        uint32_t kept = splits[me + 1] - splits[me];
        size_t cursor = kept;
        d->numMigratedOut = d->numMigratedIn = 0;
        for (int q = 0; q < P; ++q)
        {
            if (q == me) continue;
            sendFirst[l * P + q] = splits[q];
            sendCount[l * P + q] = splits[q + 1] - splits[q];
            recvFirst[l * P + q] = cursor;
            recvCount[l * P + q] = matrix[size_t(q) * P + me];
            cursor += matrix[size_t(q) * P + me];
            d->numMigratedOut += uint32_t(sendCount[l * P + q]);
            d->numMigratedIn += uint32_t(recvCount[l * P + q]);
        }
        if (cursor > s->capacity)
            throw std::runtime_error("domain: particle capacity exceeded by migration (" + std::to_string(cursor) + " > " +
                                     std::to_string(s->capacity) + ")");
        newOwn[l] = uint32_t(cursor);
        // the kept range moves to the front of the spare arrays
        for (int id = 0; id < kNumConserved; ++id)
        {
            size_t eb = s->elemBytes(id);
            SPHB200_CUDA(cudaMemcpyAsync(s->fieldAlt[id], static_cast<char*>(s->field[id]) + size_t(splits[me]) * eb,
                                         size_t(kept) * eb, cudaMemcpyDeviceToDevice, s->ctx->stream));
        }
    }
    std::vector<const void*> sendBase(size_t(kNumConserved) * NL);
    std::vector<void*> recvBase(size_t(kNumConserved) * NL);
    int eb[kNumConserved];
    for (int id = 0; id < kNumConserved; ++id)
    {
        eb[id] = g.sims[0]->elemBytes(id);
        for (int l = 0; l < NL; ++l)
        {
            sendBase[size_t(id) * NL + l] = g.sims[l]->field[id];
            recvBase[size_t(id) * NL + l] = g.sims[l]->fieldAlt[id];
        }
    }
    commExchangeArrays(g.comm, g.rs, kNumConserved, eb, sendBase.data(), sendFirst.data(), sendCount.data(),
                       recvBase.data(), recvFirst.data(), recvCount.data());
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        for (int id = 0; id < kNumConserved; ++id)
            std::swap(s->field[id], s->fieldAlt[id]);
        s->dist->numOwn = newOwn[l];
        s->start        = 0;
        s->end          = newOwn[l];
        // second sort: merges the arrivals (assignment.hpp:157-167)
        keysAndSort(s, 0, newOwn[l]);
    }
}

// ---- SPHB200_SYNC_PROFILE=1: wall time of every stage of a sync (a stream synchronisation closes each one), averaged
// over the calls after a warm-up of 4 (communicator set-up, arena growth) and printed by rank 0 every 8 calls
struct SyncProfile
{
    bool on = std::getenv("SPHB200_SYNC_PROFILE") != nullptr;
    std::vector<std::pair<std::string, double>> acc;
    int calls = 0;
    double t0 = 0;
    static double now()
    {
        return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }
    void begin(Group& g)
    {
        if (!on) return;
        for (auto* s : g.sims)
            syncStream(s);
        t0 = now();
    }
    void tick(Group& g, const char* name)
    {
        if (!on) return;
        for (auto* s : g.sims)
            syncStream(s);
        double t = now();
        if (calls >= 4)
        {
            auto it = std::find_if(acc.begin(), acc.end(), [&](auto& e) { return e.first == name; });
            if (it == acc.end()) acc.emplace_back(name, t - t0);
            else it->second += t - t0;
        }
        t0 = t;
    }
    void end(Group& g)
    {
        if (!on) return;
        ++calls;
        if (g.sims[0]->dist->rank != 0 || calls <= 4 || (calls - 4) % 8 != 0) return;
        std::string line = "domain sync profile (ms per call over " + std::to_string(calls - 4) + " calls):";
        for (auto& e : acc)
        {
            char buf[96];
            std::snprintf(buf, sizeof(buf), " %s %.3f |", e.first.c_str(), e.second / (calls - 4));
            line += buf;
        }
        std::fprintf(stderr, "%s\n", line.c_str());
    }
};
SyncProfile gSyncProfile;

//! @brief arena pointer with half as much again when it has to grow: the sizes of the halo bookkeeping (boxes, send
//!        lists, exchange buffers) creep up from step to step, and every regrowth is a cudaFree + cudaMalloc in the sync
inline void* roomy(Arena& a, size_t bytes)
{
    if (bytes > a.capacity) a.get(bytes + bytes / 2);
    return a.ptr;
}

// ---- 6. focus tree over the owned keys + halo boxes of the surface leaves
void focusAndBoxes(sphb200_sim* s, size_t srcOff)
{
    DistState* d = s->dist;
    auto* ctx    = s->ctx;
    size_t n     = d->numOwn;
    // h in SFC order for the per-leaf maxima (domain.hpp:213-216); the other fields wait for the final layout
    const int hOnly[1] = {F_h};
    gatherInto(s, hOnly, 1, n, srcOff, 0, false);
    // focus tree: bucketSizeFocus inside my key range, the global tree's leaves outside, every global leaf resolved
    // (rank boundaries are global leaf keys, so no focus leaf ever straddles one)
    {
        Tree& t = s->focus;
        int conv = 0, forced = 0;
        if (t.first)
        {
            if (t.capacity < 1) throw std::runtime_error("domain: focus tree capacity");
            initRootTree(ctx, t.leaves, t.counts, uint32_t(n));
            t.numLeaves = 1;
            int guard   = 0;
            do
            {
                updateOctreeGuided(ctx, d->keysTmp, n, s->cfg.bucketSizeFocus, t.leaves, t.counts, t.capacity,
                                   &t.numLeaves, &conv, s->global.leaves, s->global.numLeaves, &forced);
                if (++guard > 64) throw std::runtime_error("domain: focus tree does not converge");
            } while (!conv);
            t.first = false;
        }
        else
        {
            int guard = 0;
            do
            {
                updateOctreeGuided(ctx, d->keysTmp, n, s->cfg.bucketSizeFocus, t.leaves, t.counts, t.capacity,
                                   &t.numLeaves, &conv, s->global.leaves, s->global.numLeaves, &forced);
                if (++guard > 32) throw std::runtime_error("domain: focus tree cannot resolve the global tree");
            } while (forced);
        }
    }
    int L = s->focus.numLeaves;
    CHECK_RC(sphb200_build_octree(ctx, s->focus.leaves, L, d->fPrefixes, d->fChild, d->fParents, d->fLevelRange, d->fI2L,
                                  d->fL2I));
    CHECK_RC(sphb200_layout(ctx, s->focus.counts, L, d->fLayout));
    // my leaves: those inside [bounds[me], bounds[me+1]); the tree is built from my keys alone, so every other leaf is
    // empty and gets radius 0 -> its halo box is its own box -> containedIn(...) or no particles to ask for
    float ext = d->haloFactor;
    if (s->cfg.dtype == SPHB200_F64)
        SPHB200_LAUNCH(ctx, haloRadiiKernel<double>, gridFor(size_t(L) * 32, 256), 256, 0,
                       static_cast<const double*>(s->fieldAlt[F_h]), d->fLayout, 0, L, ext, d->radii);
    else
        SPHB200_LAUNCH(ctx, haloRadiiKernel<float>, gridFor(size_t(L) * 32, 256), 256, 0,
                       static_cast<const float*>(s->fieldAlt[F_h]), d->fLayout, 0, L, ext, d->radii);
}

} // namespace

namespace sphb200
{
//! @brief leaf index range of the focus tree that covers [lo, hi): findNodeAbove(lo), findNodeBelow(hi)
__global__ void leafRangeKernel(const KeyType* __restrict__ leaves, int numLeaves, KeyType lo, KeyType hi,
                                int* __restrict__ out)
{
    out[0] = int(lowerBound(leaves, size_t(numLeaves) + 1, lo));
    int b  = int(upperBound(leaves, size_t(numLeaves) + 1, hi)) - 1;
    out[1] = max(b, out[0]);
    out[2] = 0; // box counter
    for (int d = 0; d < 3; ++d)
    {
        out[4 + 2 * d] = INT_MAX; // extent of the owned leaves, see haloBoxesKernel
        out[5 + 2 * d] = INT_MIN;
    }
}
} // namespace sphb200

namespace
{

void discoverHalos(Group& g)
{
    const int P = g.P;
    // own leaf range + halo boxes
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        auto* ctx    = s->ctx;
        int L        = s->focus.numLeaves;
        int* dMeta   = reinterpret_cast<int*>(d->dSmall + OFF_META); // {firstLeaf, lastLeaf, numBoxes}
        SPHB200_LAUNCH(ctx, leafRangeKernel, 1, 1, 0, s->focus.leaves, L, d->bounds[d->rank], d->bounds[d->rank + 1],
                       dMeta);
        SPHB200_CUDA(cudaMemcpyAsync(d->pinned + OFF_META, dMeta, 3 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    }
    std::vector<int> firstLeaf, lastLeaf;
    for (auto* s : g.sims)
    {
        syncStream(s);
        firstLeaf.push_back(int(s->dist->pinned[OFF_META]));
        lastLeaf.push_back(int(s->dist->pinned[OFF_META + 1]));
    }
    gSyncProfile.tick(g, "hd: leaf range");
    std::vector<const void*> sendPtr;
    std::vector<void*> recvPtr;
    for (int l = 0; l < g.numLocal(); ++l)
    {
        sphb200_sim* s = g.sims[l];
        DistState* d   = s->dist;
        auto* ctx      = s->ctx;
        int nl         = lastLeaf[l] - firstLeaf[l];
        d->numFocusLeavesOwned = nl;
        int* dMeta     = reinterpret_cast<int*>(d->dSmall + OFF_META);
        int* boxes     = static_cast<int*>(roomy(d->boxArena, size_t(std::max(nl, 1)) * 8 * sizeof(int)));
        if (nl > 0)
            SPHB200_LAUNCH(ctx, haloBoxesKernel, gridFor(nl, 128), 128, 0, s->focus.leaves, d->radii, firstLeaf[l],
                           lastLeaf[l], s->box, boxes, nl, dMeta + 2, s->focus.counts, dMeta + 4);
        sendPtr.push_back(dMeta + 2);
        recvPtr.push_back(d->dSmall + OFF_BOXCNT);
    }
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), sizeof(int));
    for (auto* s : g.sims)
        SPHB200_CUDA(cudaMemcpyAsync(s->dist->pinned + OFF_BOXCNT, s->dist->dSmall + OFF_BOXCNT, P * sizeof(int),
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
    for (auto* s : g.sims)
        syncStream(s);
    std::vector<int> boxCount(P);
    int maxBoxes = 0;
    for (int q = 0; q < P; ++q)
    {
        boxCount[q] = int(g.sims[0]->dist->pinned[OFF_BOXCNT + q]);
        maxBoxes    = std::max(maxBoxes, boxCount[q]);
    }
    gSyncProfile.tick(g, "hd: boxes + count gather");
    // all-gather of the boxes, padded to the largest count
    size_t blk = size_t(std::max(maxBoxes, 1)) * 8 * sizeof(int);
    sendPtr.clear();
    recvPtr.clear();
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        d->numSurfaceBoxes = uint32_t(boxCount[d->rank]);
        // the send block must be blk bytes long: grow (contents preserved by hand, the arena only grows when too small)
        if (d->boxArena.capacity < blk)
        {
            Arena bigger;
            void* p = bigger.get(blk + blk / 2);
            SPHB200_CUDA(cudaMemcpyAsync(p, d->boxArena.ptr, size_t(boxCount[d->rank]) * 8 * sizeof(int),
                                         cudaMemcpyDeviceToDevice, s->ctx->stream));
            syncStream(s);
            d->boxArena.release();
            d->boxArena = bigger;
        }
        sendPtr.push_back(d->boxArena.ptr);
        recvPtr.push_back(roomy(d->allBoxArena, blk * P));
    }
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), blk);

    gSyncProfile.tick(g, "hd: box gather");

    // owner-side marking (findCollisions on my own linked octree, one launch per sender) + send lists
    sendPtr.clear();
    recvPtr.clear();
    const int stride = std::max(maxBoxes, 1);
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        auto* ctx    = s->ctx;
        int L        = s->focus.numLeaves;
        int me       = d->rank;
        SPHB200_CUDA(cudaMemsetAsync(d->leafMask, 0, size_t(L) * sizeof(uint32_t), ctx->stream));
        SPHB200_LAUNCH(ctx, markCollisionsAllKernel, gridFor(size_t(stride) * P, kMarkWarps), kMarkWarps * 32, 0, d->fPrefixes, d->fChild,
                       d->fI2L, s->focus.counts, static_cast<const int*>(d->allBoxArena.ptr), stride, P, me,
                       d->dSmall + OFF_BOXCNT, d->bounds[me], d->bounds[me + 1], d->leafMask,
                       reinterpret_cast<const int*>(d->dSmall + OFF_META) + 4);
    }
    gSyncProfile.tick(g, "hd: marking");
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        auto* ctx    = s->ctx;
        int L        = s->focus.numLeaves;
        int me       = d->rank;
        // (peer, leaf) flags -> ranges in peer-major order
        size_t PL        = size_t(P) * L;
        auto* base       = static_cast<uint32_t*>(roomy(d->listArena, (4 * PL + 8) * sizeof(uint32_t)));
        uint32_t* flags  = base;
        uint32_t* lens   = base + PL;
        uint32_t* rScan  = base + 2 * PL; // ranges: element prefix
        uint32_t* rBegin = base + 3 * PL; // ranges: first particle
        uint32_t* totals = d->dSmall + OFF_TOTALS;
        SPHB200_LAUNCH(ctx, sendFlagsKernel, gridFor(PL, 256), 256, 0, d->leafMask, s->focus.counts, L, P, me, flags, lens);
        exclusiveScan(ctx, flags, flags, PL, totals);
        exclusiveScan(ctx, lens, lens, PL, totals + 1);
        SPHB200_LAUNCH(ctx, sendRangesKernel, gridFor(PL, 256), 256, 0, flags, lens, d->leafMask, s->focus.counts,
                       d->fLayout, L, P, me, rScan, rBegin);
        SPHB200_LAUNCH(ctx, peerTotalsKernel, 1, kMaxPeersDomain + 1, 0, flags, lens, totals, L, P,
                       d->dSmall + OFF_HALOCNT);
        d->dSendScan  = rScan;
        d->dSendBegin = rBegin;
        sendPtr.push_back(d->dSmall + OFF_HALOCNT);
        recvPtr.push_back(d->dSmall + OFF_HALOMAT);
    }
    gSyncProfile.tick(g, "hd: send lists");
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), size_t(P + 1) * sizeof(uint32_t));
    for (auto* s : g.sims)
        SPHB200_CUDA(cudaMemcpyAsync(s->dist->pinned + OFF_HALOMAT, s->dist->dSmall + OFF_HALOMAT,
                                     size_t(P) * (P + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->ctx->stream));
    for (auto* s : g.sims)
    {
        syncStream(s);
        DistState* d      = s->dist;
        int me            = d->rank;
        const uint32_t* m = d->pinned + OFF_HALOMAT; // row q = {elements q sends to each peer, numRanges of q}
        d->sendCount.assign(P, 0);
        d->recvCount.assign(P, 0);
        for (int q = 0; q < P; ++q)
        {
            d->sendCount[q] = m[size_t(me) * (P + 1) + q];
            d->recvCount[q] = m[size_t(q) * (P + 1) + me];
        }
        d->numSendRanges = int(m[size_t(me) * (P + 1) + P]);
    }
}

// ---- 8. layout + final reorder
extern const int kEarlyIds[6];
extern const int kLateIds[8];
void layoutAndReorder(sphb200_sim* s, size_t srcOff, bool earlyOnly = false)
{
    DistState* d = s->dist;
    auto* ctx    = s->ctx;
    const int P  = d->P;
    int me       = d->rank;
    uint32_t before = 0, after = 0;
    for (int q = 0; q < me; ++q)
        before += d->recvCount[q];
    for (int q = me + 1; q < P; ++q)
        after += d->recvCount[q];
    size_t nTotal = size_t(before) + d->numOwn + after;
    if (nTotal > s->capacity)
        throw std::runtime_error("domain: particle capacity exceeded (" + std::to_string(nTotal) + " owned + halo > " +
                                 std::to_string(s->capacity) + ")");
    uint32_t start = before, end = before + d->numOwn;
    // halos of peer q are one contiguous block, blocks in rank order: the array is SFC-sorted as a whole
    d->haloOff.assign(P, 0);
    uint32_t cur = 0;
    auto* hScan  = d->pinned + OFF_RECVSCAN;
    auto* hBegin = d->pinned + OFF_RECVBEGIN;
    uint32_t scan = 0;
    for (int q = 0; q < P; ++q)
    {
        if (q == me) cur = end;
        d->haloOff[q] = cur;
        hScan[q]      = scan;
        hBegin[q]     = cur;
        if (q != me)
        {
            cur += d->recvCount[q];
            scan += d->recvCount[q];
        }
    }
    hScan[P] = scan;
    SPHB200_CUDA(cudaMemcpyAsync(d->dSmall + OFF_RECVSCAN, hScan, (P + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                 ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(d->dSmall + OFF_RECVBEGIN, hBegin, P * sizeof(uint32_t), cudaMemcpyHostToDevice,
                                 ctx->stream));
    d->dRecvScan  = d->dSmall + OFF_RECVSCAN;
    d->dRecvBegin = d->dSmall + OFF_RECVBEGIN;
    // send ranges were found relative to the first owned particle
    if (d->numSendRanges > 0 && start > 0)
        SPHB200_LAUNCH(ctx, addOffsetKernel, gridFor(d->numSendRanges, 256), 256, 0, d->dSendBegin, d->numSendRanges, start);
    // every conserved field straight to its final place (domain.hpp:532-560 updateLayout)
    if (earlyOnly) gatherInto(s, kEarlyIds, 6, d->numOwn, srcOff, start, true);
    else gatherInto(s, kAllConserved, kNumConserved, d->numOwn, srcOff, start, true);
    SPHB200_CUDA(cudaMemcpyAsync(s->keys + start, d->keysTmp, size_t(d->numOwn) * sizeof(uint64_t),
                                 cudaMemcpyDeviceToDevice, ctx->stream));
    s->start = start;
    s->end   = end;
    s->n     = uint32_t(nTotal);
}

// ---- haloExchangeGpu (halos/exchange_halos_gpu.cuh:50-143): one pack kernel, one message per peer, one unpack kernel
//! @brief exchangeHalos bracketed by the events of timer slot `slot` (sphb200_domain_info_t::haloExchangeMs)
void exchangeHalos(Group& g, const int* ids, int count);
void exchangeHalosTimed(Group& g, const int* ids, int count, int slot)
{
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        for (int k = 2 * slot; k < 2 * slot + 2; ++k)
            if (!d->haloEv[k]) SPHB200_CUDA(cudaEventCreate(&d->haloEv[k]));
        SPHB200_CUDA(cudaEventRecord(d->haloEv[2 * slot], s->ctx->stream));
    }
    exchangeHalos(g, ids, count);
    for (auto* s : g.sims)
    {
        SPHB200_CUDA(cudaEventRecord(s->dist->haloEv[2 * slot + 1], s->ctx->stream));
        s->dist->haloEvValid[slot] = true;
    }
}

void exchangeHalos(Group& g, const int* ids, int count)
{
    const int P = g.P, NL = g.numLocal();
    if (P == 1 || count == 0) return;
    std::vector<size_t> sendOff(size_t(NL) * P, 0), sendBytes(size_t(NL) * P, 0), recvOff(size_t(NL) * P, 0),
        recvBytes(size_t(NL) * P, 0);
    std::vector<const void*> sendBase(NL);
    std::vector<void*> recvBase(NL);
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        DistState* d   = s->dist;
        const void* src[kNumFields];
        int eb[kNumFields];
        for (int k = 0; k < count; ++k)
        {
            if (!s->field[ids[k]]) throw std::runtime_error("exchange_halos: field is not allocated");
            src[k] = s->field[ids[k]];
            eb[k]  = s->elemBytes(ids[k]);
        }
        std::vector<size_t> sb(P), rb(P);
        CHECK_RC(sphb200_pack_ranges_peers(s->ctx, nullptr, nullptr, 0, P, d->sendCount.data(), count, src, eb, nullptr,
                                           sb.data()));
        CHECK_RC(sphb200_pack_ranges_peers(s->ctx, nullptr, nullptr, 0, P, d->recvCount.data(), count, src, eb, nullptr,
                                           rb.data()));
        size_t so = 0, ro = 0;
        for (int q = 0; q < P; ++q)
        {
            sendOff[l * P + q]   = so;
            sendBytes[l * P + q] = sb[q];
            recvOff[l * P + q]   = ro;
            recvBytes[l * P + q] = rb[q];
            so += sb[q];
            ro += rb[q];
        }
        void* sendBuf = roomy(d->sendArena, std::max<size_t>(so, 8));
        void* recvBuf = roomy(d->recvArena, std::max<size_t>(ro, 8));
        CHECK_RC(sphb200_pack_ranges_peers(s->ctx, d->dSendScan, d->dSendBegin, d->numSendRanges, P, d->sendCount.data(),
                                           count, src, eb, sendBuf, nullptr));
        sendBase[l] = sendBuf;
        recvBase[l] = recvBuf;
        d->haloBytes += so;
    }
    commExchange(g.comm, g.rs, sendBase.data(), sendOff.data(), sendBytes.data(), recvBase.data(), recvOff.data(),
                 recvBytes.data());
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        DistState* d   = s->dist;
        void* dst[kNumFields];
        int eb[kNumFields];
        for (int k = 0; k < count; ++k)
        {
            dst[k] = s->field[ids[k]];
            eb[k]  = s->elemBytes(ids[k]);
        }
        CHECK_RC(sphb200_unpack_ranges_peers(s->ctx, d->dRecvScan, d->dRecvBegin, P, P, d->recvCount.data(), count, dst,
                                             eb, recvBase[l], nullptr));
    }
}

// ---- 9. keys of the halos + the tree the neighbor search walks
void searchTree(sphb200_sim* s)
{
    DistState* d = s->dist;
    auto* ctx    = s->ctx;
    auto* x = static_cast<double*>(s->field[F_x]);
    auto* y = static_cast<double*>(s->field[F_y]);
    auto* z = static_cast<double*>(s->field[F_z]);
    if (s->start > 0) CHECK_RC(sphb200_sfc_keys(ctx, x, y, z, s->start, &s->box, s->keys));
    if (s->n > s->end)
        CHECK_RC(sphb200_sfc_keys(ctx, x + s->end, y + s->end, z + s->end, s->n - s->end, &s->box, s->keys + s->end));
    updateTreeOnce(s, d->search, s->cfg.bucketSizeFocus, s->keys, s->n);
    int L = d->search.numLeaves;
    CHECK_RC(sphb200_build_octree(ctx, d->search.leaves, L, s->prefixes, s->childOffsets, s->parents, s->levelRange,
                                  s->internalToLeaf, s->leafToInternal));
    CHECK_RC(sphb200_node_centers(ctx, s->prefixes, L + (L - 1) / 7, &s->box, s->centers, s->sizes));
    CHECK_RC(sphb200_layout(ctx, d->search.counts, L, s->layout));
}

/*! @brief hooks of the host-resident step (sphb200_domain_step_host): the sync waits for the upload groups right where it
 *         first needs them and leaves the fields nothing needs before IAD un-reordered (lateIds) until finishReorder */
struct SyncStaging
{
    bool active = false;
    std::vector<size_t> lateSrcOff; // per local rank: where the un-reordered late fields sit
    std::vector<bool> latePending;
};

extern const int kEarlyIds[6] = {F_x, F_y, F_z, F_h, F_m, F_temp};
extern const int kLateIds[8]  = {F_vx, F_vy, F_vz, F_x_m1, F_y_m1, F_z_m1, F_du_m1, F_alpha};

void waitUpload(sphb200_sim* s, int group)
{
    SPHB200_CUDA(cudaStreamWaitEvent(s->ctx->stream, s->evUp[group], 0));
}

void domainSync(Group& g, SyncStaging* st = nullptr)
{
    gSyncProfile.begin(g);
    auto tick = [&](const char* name) { gSyncProfile.tick(g, name); };
    for (auto* s : g.sims)
    {
        s->ctx->packValid = 0;
        s->mark(0);
        s->dist->numOwn = s->end - s->start;
    }
    syncBox(g);
    tick("box");
    for (auto* s : g.sims)
        keysAndSort(s, s->start, s->dist->numOwn);
    tick("keys+sort");
    syncGlobalTree(g);
    tick("global tree");
    std::vector<uint32_t> matrix = assign(g);
    tick("assignment");
    uint64_t moved = 0;
    for (int a = 0; a < g.P; ++a)
        for (int q = 0; q < g.P; ++q)
            if (a != q) moved += matrix[size_t(a) * g.P + q];
    if (moved)
    {
        if (st)
            for (auto* s : g.sims)
                for (int k = 1; k < 4; ++k)
                    waitUpload(s, k); // the leaving particles take all their fields along
        migrate(g, matrix);
    }
    else
        for (auto* s : g.sims)
            s->dist->numMigratedIn = s->dist->numMigratedOut = 0;
    tick("migration");
    for (auto* s : g.sims)
    {
        if (st) waitUpload(s, 1); // h
        focusAndBoxes(s, s->start);
    }
    tick("focus tree");
    discoverHalos(g);
    tick("halo discovery");
    for (size_t l = 0; l < g.sims.size(); ++l)
    {
        sphb200_sim* s = g.sims[l];
        bool late      = st && !moved;
        if (st)
        {
            waitUpload(s, 3); // m, temp
            st->lateSrcOff[l]  = s->start;
            st->latePending[l] = late;
        }
        layoutAndReorder(s, s->start, late);
    }
    tick("layout+reorder");
    const int early[5] = {F_x, F_y, F_z, F_h, F_m};
    exchangeHalosTimed(g, early, 5, 0);
    tick("halo x,y,z,h,m");
    for (auto* s : g.sims)
    {
        searchTree(s);
        s->mark(1);
    }
    tick("search tree");
    gSyncProfile.end(g);
}

void domainForces(Group& g)
{
    auto phase = [&](int p)
    {
        for (auto* s : g.sims)
            CHECK_RC(sphb200_sim_forces_phase(s, p));
    };
    // ve_hydro.hpp:149-192: the halo exchanges between the operators
    // h travels with xm: the coupled h / neighbor-count iteration rewrote it on the owned particles, and
    // momentumEnergy reads 1/h_j.  (cstone leaves the halos' h at its pre-iteration value; sending it keeps a P-rank run
    // identical to the one-rank run also when the iteration triggers.)
    const int ex1[2] = {F_xm, F_h};
    const int ex2[6] = {F_vx, F_vy, F_vz, F_prho, F_c, F_kx};
    const int ex3[7] = {F_c11, F_c12, F_c13, F_c22, F_c23, F_c33, F_divv};
    const int ex4[1] = {F_alpha};
    phase(0);
    exchangeHalosTimed(g, ex1, 2, 1);
    phase(1);
    exchangeHalosTimed(g, ex2, 6, 2);
    phase(2);
    exchangeHalosTimed(g, ex3, 7, 3);
    phase(3);
    exchangeHalosTimed(g, ex4, 1, 4);
    if (g.sims[0]->cfg.hc.avClean)
    {
        const int ex5[6] = {F_dV11, F_dV12, F_dV13, F_dV22, F_dV23, F_dV33};
        exchangeHalos(g, ex5, 6);
    }
    phase(4);
}

void domainIntegrate(Group& g)
{
    // computeTimestep (ts_global.hpp:96-112): the local minimum stays on the device, MIN all-reduce in-stream
    std::vector<void*> ptr;
    for (auto* s : g.sims)
    {
        sphb200_fields f       = s->table();
        sphb200_hydro_const hc = s->hydroConst();
        CHECK_RC(sphb200_timestep(s->ctx, s->cfg.dtype, s->start, s->end, &hc, &f, nullptr));
        ptr.push_back(sphb200_timestep_device_ptr(s->ctx) + 2);
    }
    commAllReduce(g.comm, g.rs, ptr.data(), 1, COMM_F64, COMM_MIN);
    for (auto* s : g.sims)
        SPHB200_CUDA(cudaMemcpyAsync(s->ctx->hostDt, sphb200_timestep_device_ptr(s->ctx), 3 * sizeof(double),
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
    for (auto* s : g.sims)
    {
        syncStream(s);
        for (int k = 0; k < 3; ++k)
            s->dtLocal[k] = s->ctx->hostDt[k];
        CHECK_RC(sphb200_sim_integrate(s, s->ctx->hostDt[2]));
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------- C ABI
void destroyDist(sphb200_sim* s)
{
    DistState* d = s->dist;
    if (!d) return;
    cudaFree(d->search.leaves); cudaFree(d->search.counts);
    cudaFree(d->fPrefixes); cudaFree(d->fChild); cudaFree(d->fParents); cudaFree(d->fLevelRange); cudaFree(d->fI2L);
    cudaFree(d->fL2I); cudaFree(d->fLayout); cudaFree(d->keysTmp); cudaFree(d->leafMask); cudaFree(d->radii);
    d->boxArena.release(); d->allBoxArena.release(); d->listArena.release(); d->sendArena.release();
    d->recvArena.release(); d->scanArena.release();
    if (d->macTree) sphb200_focus_destroy(d->macTree);
    d->macSend.release(); d->macRecv.release(); d->macCounts.release();
    cudaFreeHost(d->pinned);
    cudaFree(d->dSmall);
    for (auto& e : d->haloEv)
        if (e) cudaEventDestroy(e);
    delete d;
    s->dist = nullptr;
}

extern "C" int sphb200_domain_create(sphb200_ctx* ctx, const sphb200_sim_config* cfg, sphb200_comm* comm,
                                     sphb200_sim** out)
{
    SPHB200_TRY
    if (!comm) throw std::runtime_error("domain_create: no communicator");
    if (cfg->numRanks != comm->numRanks) throw std::runtime_error("domain_create: cfg.numRanks != communicator size");
    if (cfg->rank < 0 || cfg->rank >= cfg->numRanks) throw std::runtime_error("domain_create: bad rank");
    if (!comm->loopback && cfg->rank != comm->rank) throw std::runtime_error("domain_create: cfg.rank != communicator rank");
    if (cfg->numRanks > kMaxPeersDomain) throw std::runtime_error("domain_create: at most 32 ranks");
    if (cfg->bucketSize < cfg->bucketSizeFocus) // domain.hpp:108-112
        throw std::runtime_error("The bucket size of the global tree must not be smaller than the bucket size"
                                 " of the focused tree\n");
    sphb200_sim* s = nullptr;
    CHECK_RC(sphb200_sim_create(ctx, cfg, &s));
    auto* d  = new DistState;
    s->dist  = d;
    d->comm  = comm;
    d->rank  = cfg->rank;
    d->P     = cfg->numRanks;
    try
    {
        size_t cap = s->capacity;
        d->search.capacity = s->focus.capacity;
        d->search.leaves   = devAlloc<uint64_t>(size_t(d->search.capacity) + 2);
        d->search.counts   = devAlloc<uint32_t>(size_t(d->search.capacity) + 2);
        size_t maxNodes    = size_t(s->focus.capacity) + s->focus.capacity / 7 + 2;
        d->fPrefixes   = devAlloc<uint64_t>(maxNodes);
        d->fChild      = devAlloc<int32_t>(maxNodes + 1);
        d->fParents    = devAlloc<int32_t>(maxNodes / 8 + 2);
        d->fLevelRange = devAlloc<int32_t>(32);
        d->fI2L        = devAlloc<int32_t>(maxNodes);
        d->fL2I        = devAlloc<int32_t>(maxNodes);
        d->fLayout     = devAlloc<uint32_t>(size_t(s->focus.capacity) + 2);
        d->keysTmp     = devAlloc<uint64_t>(cap);
        d->leafMask    = devAlloc<uint32_t>(size_t(s->focus.capacity) + 2);
        d->radii       = devAlloc<float>(size_t(s->focus.capacity) + 2);
        d->dSmall      = devAlloc<uint32_t>(DistState::kPinnedInts);
        SPHB200_CUDA(cudaMallocHost(&d->pinned, DistState::kPinnedInts * sizeof(uint32_t)));
        SPHB200_CUDA(cudaMemset(d->dSmall, 0, DistState::kPinnedInts * sizeof(uint32_t)));
    }
    catch (...)
    {
        sphb200_sim_destroy(s);
        throw;
    }
    *out = s;
    SPHB200_CATCH
}

extern "C" int sphb200_domain_sync(sphb200_sim* const* sims, int numLocal)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    domainSync(g);
    SPHB200_CATCH
}

extern "C" int sphb200_domain_exchange_halos(sphb200_sim* const* sims, int numLocal, const int* fieldIds, int numFields)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    for (int k = 0; k < numFields; ++k)
        if (fieldIds[k] < 0 || fieldIds[k] >= kNumFields) throw std::runtime_error("exchange_halos: bad field id");
    if (numFields > 16) throw std::runtime_error("exchange_halos: at most 16 fields per call");
    exchangeHalos(g, fieldIds, numFields);
    SPHB200_CATCH
}

extern "C" int sphb200_domain_compute_forces(sphb200_sim* const* sims, int numLocal)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    domainSync(g);
    domainForces(g);
    SPHB200_CATCH
}

extern "C" int sphb200_domain_integrate(sphb200_sim* const* sims, int numLocal)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    domainIntegrate(g);
    SPHB200_CATCH
}

extern "C" int sphb200_domain_step(sphb200_sim* const* sims, int numLocal)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    domainSync(g);
    domainForces(g);
    domainIntegrate(g);
    SPHB200_CATCH
}

extern "C" int sphb200_domain_conserved(sphb200_sim* const* sims, int numLocal, double* out10_h)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    std::vector<void*> ptr;
    for (auto* s : g.sims)
    {
        CHECK_RC(sphb200_sim_conserved(s, nullptr));
        ptr.push_back(sphb200_conserved_device_ptr(s->ctx));
    }
    commAllReduce(g.comm, g.rs, ptr.data(), 10, COMM_F64, COMM_SUM);
    sphb200_sim* s0 = g.sims[0];
    SPHB200_CUDA(cudaMemcpyAsync(s0->ctx->hostDt + 8, ptr[0], 10 * sizeof(double), cudaMemcpyDeviceToHost,
                                 s0->ctx->stream));
    syncStream(s0);
    for (int k = 0; k < 10; ++k)
        out10_h[k] = s0->ctx->hostDt[8 + k];
    SPHB200_CATCH
}

// ---- a10: the locally essential tree of every rank, MAC-coarsened outside its key range -------------------------
namespace
{

//! @brief out[0 .. stride) = leaves[0 .. numKeys) followed by 2^63 (a block of equal length for the all-gather)
__global__ void padLeavesKernel(const uint64_t* __restrict__ leaves, int numKeys, int stride, uint64_t* __restrict__ out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < stride) out[i] = i < numKeys ? leaves[i] : nodeRange(0);
}

/*! @brief one round of FocusedOctree::converge (focus/octree_focus_mpi.hpp:498-516) as Domain::sync runs it
 *         (domain.hpp:216-227): updateMinMac, updateTree with the boundaries of the MAC peers enforced, updateCounts
 *
 * updateCounts in the reference: own leaves are counted locally, the leaves in a peer's range are counted by that
 * peer (treelet exchange, exchange_focus.hpp:113-198), everything else is looked up in the replicated global tree.
 * Here every rank's leaf array goes to every rank in ONE all-gather (a few MB over NVLink), each rank counts its own
 * particles in all of them with one launch, and one all-reduce sums the counts: every leaf of every rank's tree
 * gets its exact global count, also outside the peers' ranges.
 */
bool updateMacFocusOnce(Group& g, float theta)
{
    const int P       = g.P;
    float invThetaEff = 1.0f / theta + 0.5f; // invThetaMinMac (traversal/macs.hpp:44)
    std::vector<const void*> sendPtr;
    std::vector<void*> recvPtr;
    std::vector<int> converged(g.numLocal(), 0);
    for (int l = 0; l < g.numLocal(); ++l)
    {
        sphb200_sim* s = g.sims[l];
        DistState* d   = s->dist;
        if (int(d->bounds.size()) != P + 1 || d->gLeavesH.empty())
            throw std::runtime_error("domain_update_focus: call sphb200_domain_sync first");
        if (d->macTree && d->macTheta != theta)
        {
            sphb200_focus_destroy(d->macTree);
            d->macTree = nullptr;
        }
        if (!d->macTree)
        {
            CHECK_RC(sphb200_focus_create(s->ctx, s->cfg.bucketSizeFocus, theta, &d->macTree));
            d->macTheta = theta;
            d->macFirst = true;
        }
        uint64_t focusStart = d->bounds[d->rank], focusEnd = d->bounds[d->rank + 1];
        if (focusEnd <= focusStart) throw std::runtime_error("domain_update_focus: rank without a key range");
        // peers: data identical on all ranks, no communication (domain.hpp:216-217)
        d->macPeers.assign(P, 0);
        CHECK_RC(sphb200_find_peers_mac(d->rank, P, d->bounds.data(), d->gLeavesH.data(), int(d->gLeavesH.size()) - 1,
                                        &s->box, invThetaEff, d->macPeers.data()));
        std::vector<uint64_t> enforced;
        for (int q = 0; q < P; ++q)
            if (d->macPeers[q])
            {
                enforced.push_back(d->bounds[q]);
                enforced.push_back(d->bounds[q + 1]);
            }
        CHECK_RC(sphb200_focus_update_min_mac(d->macTree, &s->box, focusStart, focusEnd));
        CHECK_RC(sphb200_focus_update_tree(d->macTree, focusStart, focusEnd, enforced.data(), int(enforced.size()),
                                           &converged[l]));
        sphb200_focus_view v;
        CHECK_RC(sphb200_focus_get_view(d->macTree, &v));
        d->pinned[OFF_META + 8] = uint32_t(v.numLeaves);
        d->pinned[OFF_META + 9] = uint32_t(converged[l]);
        SPHB200_CUDA(cudaMemcpyAsync(d->dSmall + OFF_META + 8, d->pinned + OFF_META + 8, 2 * sizeof(uint32_t),
                                     cudaMemcpyHostToDevice, s->ctx->stream));
        sendPtr.push_back(d->dSmall + OFF_META + 8);
        recvPtr.push_back(d->dSmall + OFF_HALOMAT); // free between syncs: P x {numLeaves, converged}
    }
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), 2 * sizeof(uint32_t));
    for (auto* s : g.sims)
        SPHB200_CUDA(cudaMemcpyAsync(s->dist->pinned + OFF_HALOMAT, s->dist->dSmall + OFF_HALOMAT,
                                     size_t(2 * P) * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->ctx->stream));
    for (auto* s : g.sims)
        syncStream(s);
    int maxLeaves = 0, numConverged = 0;
    for (int q = 0; q < P; ++q)
    {
        maxLeaves = std::max(maxLeaves, int(g.sims[0]->dist->pinned[OFF_HALOMAT + 2 * q]));
        numConverged += int(g.sims[0]->dist->pinned[OFF_HALOMAT + 2 * q + 1]);
    }
    const int stride = maxLeaves + 1;

    sendPtr.clear();
    recvPtr.clear();
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        sphb200_focus_view v;
        CHECK_RC(sphb200_focus_get_view(d->macTree, &v));
        auto* send = static_cast<uint64_t*>(d->macSend.get(size_t(stride) * sizeof(uint64_t)));
        SPHB200_LAUNCH(s->ctx, padLeavesKernel, gridFor(stride, 256), 256, 0, v.leaves, v.numLeaves + 1, stride, send);
        sendPtr.push_back(send);
        recvPtr.push_back(d->macRecv.get(size_t(stride) * P * sizeof(uint64_t)));
    }
    commAllGather(g.comm, g.rs, sendPtr.data(), recvPtr.data(), size_t(stride) * sizeof(uint64_t));
    std::vector<void*> countPtr;
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        auto* counts = static_cast<uint32_t*>(d->macCounts.get(size_t(stride) * P * sizeof(uint32_t)));
        // the P padded leaf arrays as one array of P * stride - 1 key intervals: the interval between two blocks
        // (2^63 followed by 0) yields a meaningless number in a slot nobody reads
        CHECK_RC(sphb200_node_counts(s->ctx, static_cast<const uint64_t*>(d->macRecv.ptr), stride * P - 1,
                                     s->keys + s->start, size_t(s->end - s->start), counts));
        SPHB200_CUDA(cudaMemsetAsync(counts + size_t(stride) * P - 1, 0, sizeof(uint32_t), s->ctx->stream));
        countPtr.push_back(counts);
    }
    commAllReduce(g.comm, g.rs, countPtr.data(), size_t(stride) * P, COMM_U32, COMM_SUM);
    for (auto* s : g.sims)
    {
        DistState* d = s->dist;
        CHECK_RC(sphb200_focus_set_leaf_counts(d->macTree,
                                               static_cast<const uint32_t*>(d->macCounts.ptr) + size_t(stride) * d->rank));
    }
    return numConverged == P;
}

} // namespace

extern "C" int sphb200_domain_update_focus(sphb200_sim* const* sims, int numLocal, float theta, int* converged_h)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    if (!(theta > 0)) throw std::runtime_error("domain_update_focus: theta must be positive");
    bool first = false;
    for (auto* s : g.sims)
        first = first || !s->dist->macTree || s->dist->macTheta != theta || s->dist->macFirst;
    bool converged = updateMacFocusOnce(g, theta);
    if (first)
    {
        // firstCall_: FocusedOctree::converge (domain.hpp:219-223), every rank iterates until all have converged
        int guard = 0;
        while (!converged)
        {
            converged = updateMacFocusOnce(g, theta);
            if (++guard > 64) throw std::runtime_error("domain_update_focus: the focus trees do not converge");
        }
        for (auto* s : g.sims)
            s->dist->macFirst = false;
    }
    if (converged_h) *converged_h = converged ? 1 : 0;
    SPHB200_CATCH
}

extern "C" int sphb200_domain_focus_view(const sphb200_sim* s, sphb200_focus_view* out, int* peers_h)
{
    SPHB200_TRY
    if (!s->dist || !s->dist->macTree) throw std::runtime_error("domain_focus_view: call sphb200_domain_update_focus first");
    CHECK_RC(sphb200_focus_get_view(s->dist->macTree, out));
    if (peers_h)
        for (int q = 0; q < s->dist->P; ++q)
            peers_h[q] = s->dist->macPeers[q];
    SPHB200_CATCH
}

extern "C" void sphb200_domain_set_halo_factor(sphb200_sim* s, float factor)
{
    if (s && s->dist) s->dist->haloFactor = factor;
}

extern "C" int sphb200_domain_info(const sphb200_sim* s, sphb200_domain_info_t* out)
{
    SPHB200_TRY
    if (!s->dist) throw std::runtime_error("domain_info: not a multi-rank handle");
    const DistState* d = s->dist;
    std::memset(out, 0, sizeof(*out));
    out->rank            = d->rank;
    out->numRanks        = d->P;
    out->startIndex      = s->start;
    out->endIndex        = s->end;
    out->nWithHalos      = s->n;
    out->numGlobalLeaves = s->global.numLeaves;
    out->numFocusLeaves  = s->focus.numLeaves;
    out->numSearchLeaves = d->search.numLeaves;
    out->numSurfaceBoxes = d->numSurfaceBoxes;
    out->numSendRanges   = uint32_t(d->numSendRanges);
    out->migratedIn      = d->numMigratedIn;
    out->migratedOut     = d->numMigratedOut;
    out->haloBytes       = d->haloBytes;
    for (int q = 0; q < d->P && q < 32; ++q)
    {
        out->haloSend[q] = q < int(d->sendCount.size()) ? d->sendCount[q] : 0;
        out->haloRecv[q] = q < int(d->recvCount.size()) ? d->recvCount[q] : 0;
    }
    for (int q = 0; q <= d->P && q < 33; ++q)
        out->assignment[q] = q < int(d->bounds.size()) ? d->bounds[q] : 0;
    for (int k = 0; k < 5; ++k)
    {
        float ms = 0;
        if (d->haloEvValid[k] && cudaEventSynchronize(d->haloEv[2 * k + 1]) == cudaSuccess &&
            cudaEventElapsedTime(&ms, d->haloEv[2 * k], d->haloEv[2 * k + 1]) == cudaSuccess)
        {
            out->haloExchangeMs[k] = ms;
            out->haloExchangeMs[5] += ms;
        }
    }
    cudaGetLastError();
    SPHB200_CATCH
}

//! @brief makeSfcAssignment + limitBoundaryShifts on the host (no device needed)
extern "C" int sphb200_sfc_assignment(int numRanks, const uint32_t* counts_h, const uint64_t* leaves_h, int numLeaves,
                                      const uint64_t* oldBounds_h, uint64_t* bounds_h, uint32_t* rankCounts_h)
{
    SPHB200_TRY
    if (numRanks < 1 || numLeaves < 1) throw std::runtime_error("sfc_assignment: bad sizes");
    sfcAssignment(numRanks, counts_h, leaves_h, numLeaves, oldBounds_h, bounds_h, rankCounts_h);
    SPHB200_CATCH
}

//! @brief the tree a GlobalAssignment starts from (assignment.hpp:62-69), host only; returns the number of leaves
extern "C" int sphb200_initial_global_tree(int numRanks, uint64_t* leaves_h, int capacity)
{
    try
    {
        std::vector<uint64_t> t = initialGlobalTree(numRanks);
        if (int(t.size()) > capacity) return -int(t.size());
        std::copy(t.begin(), t.end(), leaves_h);
        return int(t.size()) - 1;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 0;
    }
}

/*! @brief findHalos (traversal/collisions.hpp:78-105) as the two kernels of the owner-side discovery, on ONE tree:
 *         halo boxes of the leaves [firstLeaf, lastLeaf) (makeHaloBox + containedIn), then findCollisions of those
 *         boxes against the whole linked octree, nodes inside [leaves[firstLeaf], leaves[lastLeaf]) excluded.
 *         radii: float[numLeaves] (2 * searchExt * max h of the leaf); flags: int[numLeaves], zeroed by the caller.
 */
extern "C" int sphb200_find_halos(sphb200_ctx* ctx, const uint64_t* prefixes, const int32_t* childOffsets,
                                  const int32_t* internalToLeaf, const uint64_t* leaves, const float* radii,
                                  const sphb200_box* box, int firstLeaf, int lastLeaf, int* flags, int* numBoxes_h)
{
    SPHB200_TRY
    int nl = lastLeaf - firstLeaf;
    if (nl <= 0) return 0;
    char* tmp  = static_cast<char*>(ctx->miscTmp.get(size_t(nl) * 8 * sizeof(int) + 64));
    int* cnt   = reinterpret_cast<int*>(tmp);
    int* boxes = reinterpret_cast<int*>(tmp + 64);
    SPHB200_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int), ctx->stream));
    SPHB200_LAUNCH(ctx, haloBoxesKernel, gridFor(nl, 128), 128, 0, leaves, radii, firstLeaf, lastLeaf, *box, boxes, nl,
                   cnt, (const uint32_t*)nullptr, (int*)nullptr);
    uint64_t range[2];
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, cnt, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(&range[0], leaves + firstLeaf, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(&range[1], leaves + lastLeaf, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    int nb = ctx->hostScalars[0];
    if (numBoxes_h) *numBoxes_h = nb;
    if (nb > 0)
        SPHB200_LAUNCH(ctx, markCollisionsKernel, gridFor(nb, kMarkWarps), kMarkWarps * 32, 0, prefixes, childOffsets, internalToLeaf,
                       (const uint32_t*)nullptr, boxes, nb, 1u, range[0], range[1], KeyType(0), nodeRange(0),
                       (uint32_t*)nullptr, flags);
    SPHB200_CATCH
}

/*! @brief halo radii as Halos::discover computes them: radii[leaf] = float(2 * searchExt * max h over
 *         [layout[leaf], layout[leaf+1])), 0 for empty leaves, for the leaves [firstLeaf, lastLeaf) */
extern "C" int sphb200_halo_radii(sphb200_ctx* ctx, const void* h, int dtype, const uint32_t* layout, int firstLeaf,
                                  int lastLeaf, float searchExt, float* radii)
{
    SPHB200_TRY
    int nl = lastLeaf - firstLeaf;
    if (nl <= 0) return 0;
    if (dtype == SPHB200_F64)
        SPHB200_LAUNCH(ctx, haloRadiiKernel<double>, gridFor(size_t(nl) * 32, 256), 256, 0, static_cast<const double*>(h),
                       layout, firstLeaf, lastLeaf, searchExt, radii);
    else
        SPHB200_LAUNCH(ctx, haloRadiiKernel<float>, gridFor(size_t(nl) * 32, 256), 256, 0, static_cast<const float*>(h),
                       layout, firstLeaf, lastLeaf, searchExt, radii);
    SPHB200_CATCH
}

/*! @brief one step on host-resident state for the ranks this process drives (the call a host-resident front-end makes)
 *
 * Same results as upload + step + download, with the PCIe transfers staged beside the kernels like the single-rank
 * sphb200_sim_step_host: uploads in order of first use on a copy stream ({x,y,z} | h | {m,temp} | the other eight
 * fields, which are only reordered right before IAD), m / next-step h / alpha go back as soon as they are final, the
 * Courant step is evaluated with AVswitches and all-reduced before MomentumEnergy, which then runs in 16 particle
 * ranges, each integrated and sent to the host while the next one computes.  n_in[l] particles in, n_out[l] out (the
 * assignment moves particles between ranks); host_out arrays hold hostCapacity elements.
 */
extern "C" int sphb200_domain_step_host(sphb200_sim* const* sims, int numLocal, const size_t* n_in,
                                        const sphb200_fields* host_in, const sphb200_fields* host_out, size_t* n_out,
                                        size_t hostCapacity)
{
    SPHB200_TRY
    Group g = makeGroup(sims, numLocal);
    const int NL = g.numLocal();
    constexpr int kRanges = 16;
    static const int groupA[]  = {F_x, F_y, F_z};
    static const int groupH[]  = {F_h};
    static const int groupMT[] = {F_m, F_temp};
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        if (n_in[l] > s->capacity) throw std::runtime_error("domain_step_host: n exceeds capacity");
        if (!s->copyIn)
        {
            SPHB200_CUDA(cudaStreamCreateWithFlags(&s->copyIn, cudaStreamNonBlocking));
            SPHB200_CUDA(cudaStreamCreateWithFlags(&s->copyOut, cudaStreamNonBlocking));
            for (auto& e : s->evUp) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            for (auto& e : s->evDown) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        }
        if (!s->evRange[0])
            for (auto& e : s->evRange) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        syncStream(s);
        s->ctx->packValid = 0;
        size_t n = n_in[l];
        s->start = 0;
        s->n = s->end = uint32_t(n);
        sphb200_fields hin = host_in[l];
        auto upload = [&](const int* ids, int count, cudaEvent_t done)
        {
            for (int k = 0; k < count; ++k)
            {
                void* src = *fieldSlot(hin, ids[k]);
                if (!src) throw std::runtime_error("domain_step_host: all conserved fields are required in host_in");
                SPHB200_CUDA(cudaMemcpyAsync(s->field[ids[k]], src, n * s->elemBytes(ids[k]), cudaMemcpyHostToDevice,
                                             s->copyIn));
            }
            SPHB200_CUDA(cudaEventRecord(done, s->copyIn));
        };
        upload(groupA, 3, s->evUp[0]);
        upload(groupH, 1, s->evUp[1]);
        upload(groupMT, 2, s->evUp[3]);
        upload(kLateIds, 8, s->evUp[2]);
        waitUpload(s, 0);
    }
    SyncStaging st;
    st.active = true;
    st.lateSrcOff.assign(NL, 0);
    st.latePending.assign(NL, false);
    domainSync(g, &st);

    auto download = [&](sphb200_sim* s, const sphb200_fields& hout, int id, const void* devBase, cudaEvent_t after)
    {
        // owned range [start, end) of devBase (indexed like the fields) -> host array [0, end - start)
        sphb200_fields h = hout;
        void* dst        = *fieldSlot(h, id);
        if (!dst) return;
        size_t eb = s->elemBytes(id);
        SPHB200_CUDA(cudaEventRecord(after, s->ctx->stream));
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, after, 0));
        SPHB200_CUDA(cudaMemcpyAsync(dst, static_cast<const char*>(devBase) + size_t(s->start) * eb,
                                     size_t(s->end - s->start) * eb, cudaMemcpyDeviceToHost, s->copyOut));
    };
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        if (size_t(s->end - s->start) > hostCapacity)
            throw std::runtime_error("domain_step_host: host buffers too small for the owned particles");
        // the uploads of this call must have landed before results overwrite the same host arrays
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evUp[2], 0));
        download(s, host_out[l], F_m, s->field[F_m], s->evDown[0]); // m is final after the reorder
    }
    const bool ranged = !g.sims[0]->cfg.hc.avClean;
    auto phase = [&](int p)
    {
        for (auto* s : g.sims)
            CHECK_RC(sphb200_sim_forces_phase(s, p));
    };
    const int ex1[2] = {F_xm, F_h};
    const int ex2[6] = {F_vx, F_vy, F_vz, F_prho, F_c, F_kx};
    const int ex3[7] = {F_c11, F_c12, F_c13, F_c22, F_c23, F_c33, F_divv};
    const int ex4[1] = {F_alpha};
    phase(0);
    for (int l = 0; l < NL; ++l)
    {
        // next step's h = updateH(ng0, nc, h) is known once the neighbors are counted: into the spare buffer of h and
        // off to the host; the position update recomputes the same value in place
        sphb200_sim* s   = g.sims[l];
        sphb200_fields f = s->table();
        CHECK_RC(sphb200_update_h_into(s->ctx, s->cfg.dtype, s->start, s->end, s->cfg.hc.ng0, f.nc, f.h, s->fieldAlt[F_h]));
        download(s, host_out[l], F_h, s->fieldAlt[F_h], s->evDown[2]);
    }
    exchangeHalos(g, ex1, 2);
    phase(1);
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        if (!st.latePending[l]) continue;
        waitUpload(s, 2);
        gatherInto(s, kLateIds, 8, s->dist->numOwn, st.lateSrcOff[l], s->start, true);
        s->ctx->packValid = 0;
    }
    exchangeHalos(g, ex2, 6);
    phase(2);
    exchangeHalos(g, ex3, 7);
    if (!ranged)
    {
        phase(3);
        for (int l = 0; l < NL; ++l)
            download(g.sims[l], host_out[l], F_alpha, g.sims[l]->field[F_alpha], g.sims[l]->evDown[1]);
        exchangeHalos(g, ex4, 1);
        const int ex5[6] = {F_dV11, F_dV12, F_dV13, F_dV22, F_dV23, F_dV33};
        exchangeHalos(g, ex5, 6);
        phase(4);
        domainIntegrate(g);
        for (int l = 0; l < NL; ++l)
        {
            sphb200_sim* s = g.sims[l];
            SPHB200_CUDA(cudaEventRecord(s->evRange[0], s->ctx->stream));
            SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evRange[0], 0));
            sphb200_fields h = host_out[l];
            for (int id = 0; id < kNumConserved; ++id)
            {
                if (id == F_m || id == F_alpha || id == F_h) continue;
                void* dst = *fieldSlot(h, id);
                if (!dst) continue;
                size_t eb = s->elemBytes(id);
                SPHB200_CUDA(cudaMemcpyAsync(dst, static_cast<char*>(s->field[id]) + size_t(s->start) * eb,
                                             size_t(s->end - s->start) * eb, cudaMemcpyDeviceToHost, s->copyOut));
            }
            // h was sent from the spare buffer before the position update: identical values, nothing to resend
        }
    }
    else
    {
        // AVswitches + the Courant step of momentumEnergy, then computeTimestep (ts_global.hpp:96-112) across the ranks
        std::vector<void*> ptr;
        for (auto* s : g.sims)
        {
            sphb200_fields f       = s->table();
            sphb200_hydro_const hc = s->hydroConst();
            CHECK_RC(sphb200_av_switches_courant(s->ctx, s->cfg.dtype, s->start, s->end, &s->box, &hc, s->neighbors, &f));
            s->mark(7);
            CHECK_RC(sphb200_timestep(s->ctx, s->cfg.dtype, s->start, s->end, &hc, &f, nullptr));
            ptr.push_back(sphb200_timestep_device_ptr(s->ctx) + 2);
        }
        commAllReduce(g.comm, g.rs, ptr.data(), 1, COMM_F64, COMM_MIN);
        for (auto* s : g.sims)
            SPHB200_CUDA(cudaMemcpyAsync(s->ctx->hostDt, sphb200_timestep_device_ptr(s->ctx), 3 * sizeof(double),
                                         cudaMemcpyDeviceToHost, s->ctx->stream));
        for (int l = 0; l < NL; ++l)
            download(g.sims[l], host_out[l], F_alpha, g.sims[l]->field[F_alpha], g.sims[l]->evDown[1]);
        exchangeHalos(g, ex4, 1);
        for (int l = 0; l < NL; ++l)
        {
            sphb200_sim* s = g.sims[l];
            syncStream(s);
            for (int k = 0; k < 3; ++k)
                s->dtLocal[k] = s->ctx->hostDt[k];
            double dt     = s->ctx->hostDt[2];
            double dtStir = s->minDt; // TurbVeProp stirs with the time step of the previous iteration
            s->ttot += dt;
            s->minDt_m1 = s->minDt;
            s->minDt    = dt;
            if (s->turb) CHECK_RC(sphb200_turb_advance(s->turb, s->ctx, dtStir));
        }
        for (int l = 0; l < NL; ++l)
        {
            sphb200_sim* s         = g.sims[l];
            auto* ctx              = s->ctx;
            sphb200_fields f       = s->table();
            sphb200_hydro_const hc = s->hydroConst();
            int dtype              = s->cfg.dtype;
            uint32_t a = s->start, b = s->end;
            uint32_t per = ((b - a + kRanges - 1) / kRanges + 127) / 128 * 128;
            sphb200_fields h = host_out[l];
            for (int r = 0; r < kRanges; ++r)
            {
                uint32_t rb = std::min(b, a + r * per), re = std::min(b, rb + per);
                if (re <= rb) break;
                CHECK_RC(sphb200_momentum_energy_range(ctx, dtype, a, b, rb, re, r == 0, &s->box, &hc, s->neighbors, &f));
                if (s->turb) CHECK_RC(sphb200_turb_stir(s->turb, ctx, dtype, rb, re, f.x, f.y, f.z, f.ax, f.ay, f.az));
                CHECK_RC(sphb200_positions_update_h(ctx, dtype, rb, re, &s->box, &hc, s->minDt, s->minDt_m1, &f));
                SPHB200_CUDA(cudaEventRecord(s->evRange[r], ctx->stream));
                SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evRange[r], 0));
                for (int id = 0; id < kNumConserved; ++id)
                {
                    if (id == F_m || id == F_alpha || id == F_h) continue;
                    void* dst = *fieldSlot(h, id);
                    if (!dst) continue;
                    size_t eb = s->elemBytes(id);
                    SPHB200_CUDA(cudaMemcpyAsync(static_cast<char*>(dst) + size_t(rb - a) * eb,
                                                 static_cast<char*>(s->field[id]) + size_t(rb) * eb, size_t(re - rb) * eb,
                                                 cudaMemcpyDeviceToHost, s->copyOut));
                }
            }
            s->mark(8);
            s->mark(9);
        }
    }
    for (int l = 0; l < NL; ++l)
    {
        sphb200_sim* s = g.sims[l];
        n_out[l]       = s->end - s->start;
        syncStream(s);
        SPHB200_CUDA(cudaStreamSynchronize(s->copyOut));
    }
    SPHB200_CATCH
}
