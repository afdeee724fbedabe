// The locally essential ("focused") octree of cornerstone: particle-count driven inside a focus key range,
// MAC driven outside.
//
// Reference: cstone::CombinedUpdate::updateFocus[Gpu] and FocusedOctreeSingleNode (focus/octree_focus.hpp:81-288),
// rebalanceDecisionEssential / enforceKeys / protectAncestors (focus/rebalance.hpp:49-258), injectKeys
// (focus/inject.hpp:52-73), computeMinMacR2 / evaluateMacPbc / markMacs (traversal/macs.hpp:58-264), findPeersMac
// (traversal/peers.hpp:63-120), FocusedOctree::updateTree / updateCounts / updateMinMac (focus/octree_focus_mpi.hpp).
//
// Everything per-node is one kernel over the linked octree that stays on the device; what the reference does with
// one depth-first traversal per focus leaf (markMacsGpu) is evaluated here per NODE against the few spanning boxes
// of the focus range (<= 294, kept in shared memory): a node is marked iff for one box it and all its ancestors
// violate the MAC -- exactly the set of nodes the reference's traversals reach.  Node counts of internal nodes come
// from one scan of the leaf counts and two binary searches per node instead of a level-by-level upsweep.
// Compiled with --fmad=false: MAC distances are compared bit for bit with the reference's CPU code.
#include <algorithm>
#include <memory>

#include "common.cuh"
#include "hilbert.cuh"
#include "primitives.cuh"

namespace sphb200
{

//! @brief geometric center and half-size of a target box
struct GeoBox
{
    double c[3], s[3];
};

//! @brief hilbertIBox (sfc/hilbert.hpp:264-281) followed by centerAndSize (sfc/box.hpp:334-350), same operation order
__host__ __device__ inline GeoBox nodeGeometry(KeyType start, unsigned level, const sphb200_box& box)
{
    unsigned cube = (1u << kMaxTreeLevel) >> level;
    unsigned mask = ~(cube - 1);
    unsigned ix, iy, iz;
    hilbertDecode(start, ix, iy, iz);
    int lo[3] = {int(ix & mask), int(iy & mask), int(iz & mask)};

    constexpr double uL = 1.0 / double(1u << kMaxTreeLevel);
    GeoBox g;
    for (int d = 0; d < 3; ++d)
    {
        double halfUnit = (0.5 * uL) * box.len[d];
        int hi          = lo[d] + int(cube);
        g.c[d]          = box.lim[2 * d] + double(hi + lo[d]) * halfUnit;
        g.s[d]          = double(hi - lo[d]) * halfUnit;
    }
    return g;
}

//! @brief evaluateMacPbc (traversal/macs.hpp:127-139): is the target box closer to the source center than the radius?
__host__ __device__ inline bool macViolated(const double* src4, const GeoBox& t, const sphb200_box& box)
{
    double r2 = 0;
    for (int d = 0; d < 3; ++d)
    {
        double dX = t.c[d] - src4[d];
        if (box.bc[d] == SPHB200_PERIODIC) dX -= box.len[d] * rint(dX * box.inv[d]);
        dX = fabs(dX);
        dX -= t.s[d];
        dX += fabs(dX);
        dX *= 0.5;
        r2 += dX * dX;
    }
    return r2 < fabs(src4[3]);
}

//! @brief computeMinMacR2 (traversal/macs.hpp:58-71) for every node
__global__ void minMacR2Kernel(const KeyType* __restrict__ prefixes, int numNodes, float invThetaEff, sphb200_box box,
                               double* __restrict__ centers4)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    KeyType prefix = prefixes[i];
    GeoBox g       = nodeGeometry(decodePlaceholderBit(prefix), decodePrefixLength(prefix) / 3, box);
    double l       = 2.0 * fmax(fmax(g.s[0], g.s[1]), g.s[2]);
    double mac     = l * double(invThetaEff);
    centers4[4 * i + 0] = g.c[0];
    centers4[4 * i + 1] = g.c[1];
    centers4[4 * i + 2] = g.c[2];
    centers4[4 * i + 3] = mac * mac;
}

//! @brief target boxes of markMacs (macs.hpp:253-260): the spanSfcRange cells of the focus
__global__ void focusBoxesKernel(const KeyType* __restrict__ codes, int numBoxes, sphb200_box box,
                                 GeoBox* __restrict__ out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numBoxes) return;
    out[i] = nodeGeometry(codes[i], treeLevel(codes[i + 1] - codes[i]), box);
}

/*! @brief markings[i] = 1 iff node i fails the MAC paired with a box of the focus (markMacs, macs.hpp:240-264)
 *
 * markMacPerBox descends from the root while a node is not contained in the focus and violates the MAC w.r.t. the
 * box, marking every node it visits that way.  Per node: not contained, and for one box the node and all its
 * ancestors violate (ancestors of a node that is not contained in the focus are not contained either).
 */
__global__ void __launch_bounds__(256)
    markMacsKernel(const KeyType* __restrict__ prefixes, const TreeNodeIndex* __restrict__ parents,
                   const double* __restrict__ centers4, int numNodes, sphb200_box box, const GeoBox* __restrict__ boxes,
                   int numBoxes, KeyType focusStart, KeyType focusEnd, char* __restrict__ markings)
{
    extern __shared__ double shBoxes[];
    GeoBox* sh = reinterpret_cast<GeoBox*>(shBoxes);
    for (int k = threadIdx.x; k < numBoxes * 6; k += blockDim.x)
        shBoxes[k] = reinterpret_cast<const double*>(boxes)[k];
    __syncthreads();

    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    KeyType prefix    = prefixes[i];
    KeyType nodeStart = decodePlaceholderBit(prefix);
    KeyType nodeEnd   = nodeStart + nodeRange(decodePrefixLength(prefix) / 3);
    char mark         = 0;
    if (nodeStart < focusStart || nodeEnd > focusEnd)
    {
        double own[4] = {centers4[4 * i], centers4[4 * i + 1], centers4[4 * i + 2], centers4[4 * i + 3]};
        for (int b = 0; b < numBoxes && !mark; ++b)
        {
            if (!macViolated(own, sh[b], box)) continue;
            bool reached = true;
            int a        = i;
            while (a != 0)
            {
                a = parents[(a - 1) / 8];
                if (!macViolated(centers4 + 4 * a, sh[b], box))
                {
                    reached = false;
                    break;
                }
            }
            if (reached) mark = 1;
        }
    }
    markings[i] = mark;
}

//! @brief mergeCountAndMacOp for every node (focus/rebalance.hpp:49-90, rebalanceDecisionEssential :144-160)
__global__ void essentialDecisionKernel(const KeyType* __restrict__ prefixes, const TreeNodeIndex* __restrict__ childOffsets,
                                        const TreeNodeIndex* __restrict__ parents, const uint32_t* __restrict__ counts,
                                        const char* __restrict__ macs, KeyType focusStart, KeyType focusEnd,
                                        unsigned bucket, int numNodes, TreeNodeIndex* __restrict__ ops)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    KeyType prefix = prefixes[i];
    unsigned level = decodePrefixLength(prefix) / 3;
    if (i)
    {
        TreeNodeIndex parent = parents[(i - 1) / 8];
        bool countMerge      = counts[parent] <= bucket;
        bool macMerge        = macs[parent] == 0;
        // the sibling group touches the focus: it cannot be removed on behalf of the MAC
        KeyType firstGroupKey = decodePlaceholderBit(prefixes[parent]);
        KeyType lastGroupKey  = firstGroupKey + 8 * nodeRange(level);
        bool inFringe         = lastGroupKey > focusStart && focusEnd > firstGroupKey;
        if (countMerge || (macMerge && !inFringe))
        {
            ops[i] = 0;
            return;
        }
    }
    KeyType nodeStart = decodePlaceholderBit(prefix);
    bool isLeaf       = childOffsets[i] == 0;
    bool inFocus      = nodeStart >= focusStart && nodeStart < focusEnd;
    ops[i] = (isLeaf && level < unsigned(kMaxTreeLevel) && counts[i] > bucket && (macs[i] || inFocus)) ? 8 : 1;
}

/*! @brief enforceKeySingle per mandatory key (focus/rebalance.hpp:189-243); status = max over the keys
 *
 * The writes are monotone (0 -> 1 on the sibling groups of the supporting ancestors, max() on the closest node), so
 * the keys can be handled concurrently: whichever thread cancels a merge first, the others cancel the same set.
 */
__global__ void enforceKeysKernel(const KeyType* __restrict__ keys, int numKeys, const KeyType* __restrict__ prefixes,
                                  const TreeNodeIndex* __restrict__ childOffsets,
                                  const TreeNodeIndex* __restrict__ parents, TreeNodeIndex* ops, int* status)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= numKeys) return;
    KeyType key = keys[k];
    if (key == 0 || key == nodeRange(0)) return;

    // makePrefix (sfc/common.hpp:320-326): the biggest node that starts at key
    int keyPos          = kMaxTreeLevel - int(__ffsll((long long)key) - 1) / 3;
    KeyType nodeKeyWant = encodePlaceholderBit(key, 3 * keyPos);
    // containingNode (tree/octree.hpp:244-260)
    TreeNodeIndex nodeIdx = 0;
    for (int l = 1; l <= keyPos; ++l)
    {
        if (childOffsets[nodeIdx] == 0 || nodeKeyWant == prefixes[nodeIdx]) break;
        nodeIdx = childOffsets[nodeIdx] + TreeNodeIndex(octalDigit(key, unsigned(l)));
    }
    KeyType nodeKeyHave = prefixes[nodeIdx];
    int nodeLevelHave   = int(decodePrefixLength(nodeKeyHave) / 3);

    int st          = 0; // converged
    bool trySplit   = nodeKeyHave != nodeKeyWant && nodeLevelHave < kMaxTreeLevel;
    bool undoMerges = ops[nodeIdx] == 0 || trySplit;
    if (undoMerges && nodeIdx > 0)
    {
        st                   = 1; // cancelMerge
        TreeNodeIndex parent = nodeIdx;
        do
        {
            parent                     = parents[(parent - 1) / 8];
            TreeNodeIndex firstSibling = childOffsets[parent];
            for (TreeNodeIndex s = firstSibling; s < firstSibling + 8; ++s)
                atomicMax(ops + s, 1);
        } while (parent != 0);
    }
    if (trySplit)
    {
        // one level at most: deeper keys are injected afterwards (status failed)
        int levelDiff = keyPos - nodeLevelHave;
        st            = levelDiff > 1 ? 3 : 2;
        atomicMax(ops + nodeIdx, 8);
    }
    if (st) atomicMax(status, st);
}

/*! @brief nzAncestorOp for every node, in place (focus/rebalance.hpp:103-176, protectAncestors)
 *
 * Non-zero entries never change and a zero entry only turns into the value of its closest non-zero ancestor when the
 * node is that ancestor's left-most descendant, so concurrent readers see either state with the same outcome.
 */
__global__ void protectAncestorsKernel(const KeyType* __restrict__ prefixes, const TreeNodeIndex* __restrict__ parents,
                                       int numNodes, volatile TreeNodeIndex* ops, int* numChanges)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    TreeNodeIndex a = i;
    TreeNodeIndex v = ops[a];
    while (v == 0 && a != 0)
    {
        a = parents[(a - 1) / 8];
        v = ops[a];
    }
    TreeNodeIndex decision = 0;
    if (i == 0) decision = ops[0];
    else if (decodePlaceholderBit(prefixes[i]) == decodePlaceholderBit(prefixes[a])) decision = v;
    if (decision != 1) atomicAdd(numChanges, 1);
    ops[i] = decision;
}

//! @brief leafOps[k] = ops of leaf k (cornerstone order), leafOps[numLeaves] = 0; counts the leaves whose op is not 1
__global__ void gatherLeafOpsKernel(const TreeNodeIndex* __restrict__ ops, const TreeNodeIndex* __restrict__ leafToInternal,
                                    int numInternal, int numLeaves, uint32_t* __restrict__ leafOps, int* numNotOne)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > numLeaves) return;
    if (k == numLeaves)
    {
        leafOps[k] = 0;
        return;
    }
    TreeNodeIndex op = ops[leafToInternal[numInternal + k]];
    leafOps[k]       = uint32_t(op);
    if (op != 1) atomicAdd(numNotOne, 1);
}

//! @brief particle count of every node from the scanned leaf counts (scatter + upsweep of FocusedOctree::updateCounts)
__global__ void nodeCountsFromLayoutKernel(const KeyType* __restrict__ prefixes, int numNodes,
                                           const KeyType* __restrict__ leaves, int numLeaves,
                                           const uint32_t* __restrict__ layout, uint32_t* __restrict__ counts)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    KeyType prefix = prefixes[i];
    KeyType start  = decodePlaceholderBit(prefix);
    KeyType end    = start + nodeRange(decodePrefixLength(prefix) / 3);
    size_t a       = lowerBound(leaves, size_t(numLeaves) + 1, start);
    size_t b       = lowerBound(leaves, size_t(numLeaves) + 1, end);
    counts[i]      = layout[b] - layout[a];
}

__global__ void fillU32Kernel(uint32_t* p, int n, uint32_t v)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

} // namespace sphb200

using namespace sphb200;

#define CHECK_RC(call)                                                                                                \
    do                                                                                                                \
    {                                                                                                                 \
        if ((call) != 0) throw std::runtime_error(sphb200_last_error());                                              \
    } while (0)

/*! @brief device twin of cstone::FocusedOctreeSingleNode / the tree part of cstone::FocusedOctree
 *
 * leaves_, tree_ (linked octree), leafCounts_, counts_ (all nodes), macs_ (all nodes), centers_ (all nodes) live in
 * grow-only arenas; an update allocates only when the tree outgrows them.
 */
struct sphb200_focus
{
    sphb200_ctx* ctx;
    unsigned bucket;
    float theta;
    int numLeaves = 1;
    Arena leaves[2]; // current / next
    int cur = 0;
    Arena leafCounts, layout, prefixes, childOffsets, parents, levelRange, internalToLeaf, leafToInternal, counts, macs,
        centers, ops, leafOps, small, boxes;
    int numCountNodes = 1, numMacNodes = 1; // sizes of counts_ / macs_ as of their last update
    int lastStatus    = 0;

    int numInternal() const { return (numLeaves - 1) / 7; }
    int numNodes() const { return numLeaves + numInternal(); }
    KeyType* dLeaves() { return static_cast<KeyType*>(leaves[cur].ptr); }

    void link()
    {
        int nn = numNodes();
        prefixes.get(size_t(nn) * sizeof(KeyType));
        childOffsets.get((size_t(nn) + 1) * sizeof(TreeNodeIndex));
        parents.get(size_t(std::max(1, (nn - 1) / 8)) * sizeof(TreeNodeIndex));
        levelRange.get((kMaxTreeLevel + 2) * sizeof(TreeNodeIndex));
        internalToLeaf.get(size_t(nn) * sizeof(TreeNodeIndex));
        leafToInternal.get(size_t(nn) * sizeof(TreeNodeIndex));
        CHECK_RC(sphb200_build_octree(ctx, dLeaves(), numLeaves, static_cast<uint64_t*>(prefixes.ptr),
                                      static_cast<int32_t*>(childOffsets.ptr), static_cast<int32_t*>(parents.ptr),
                                      static_cast<int32_t*>(levelRange.ptr), static_cast<int32_t*>(internalToLeaf.ptr),
                                      static_cast<int32_t*>(leafToInternal.ptr)));
    }

    ~sphb200_focus()
    {
        for (Arena* a : {&leaves[0], &leaves[1], &leafCounts, &layout, &prefixes, &childOffsets, &parents, &levelRange,
                         &internalToLeaf, &leafToInternal, &counts, &macs, &centers, &ops, &leafOps, &small, &boxes})
            a->release();
    }
};

namespace
{

//! @brief injectKeys (focus/inject.hpp:52-73) on a host copy of the leaves: add the spanning tree of the keys
std::vector<KeyType> injectKeysHost(std::vector<KeyType> tree, std::vector<KeyType> keys)
{
    keys.push_back(0);
    keys.push_back(nodeRange(0));
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
    std::vector<KeyType> cells(512);
    for (size_t k = 0; k + 1 < keys.size(); ++k)
    {
        int nc = sphb200_span_sfc_range(keys[k], keys[k + 1], cells.data(), cells.size());
        if (nc < 0) throw std::runtime_error(sphb200_last_error());
        tree.insert(tree.end(), cells.begin(), cells.begin() + nc);
    }
    std::sort(tree.begin(), tree.end());
    tree.erase(std::unique(tree.begin(), tree.end()), tree.end());
    return tree;
}

//! @brief CombinedUpdate::updateFocusGpu (focus/octree_focus.hpp:158-213)
bool focusUpdateTree(sphb200_focus* f, KeyType focusStart, KeyType focusEnd, const uint64_t* mandatory_h, int numMandatory)
{
    sphb200_ctx* ctx = f->ctx;
    int nn           = f->numNodes();
    int L            = f->numLeaves;
    if (f->numCountNodes != nn || f->numMacNodes != nn)
        throw std::runtime_error("focus: update of counts and MACs required before updating the tree structure");
    if (focusEnd <= focusStart || focusEnd > nodeRange(0)) throw std::runtime_error("focus: empty focus range");

    std::vector<KeyType> allKeys{focusStart, focusEnd};
    allKeys.insert(allKeys.end(), mandatory_h, mandatory_h + numMandatory);
    for (KeyType k : allKeys)
        if (k > nodeRange(0)) throw std::runtime_error("focus: mandatory key out of range");

    auto* ops      = static_cast<TreeNodeIndex*>(f->ops.get(size_t(nn) * sizeof(TreeNodeIndex)));
    auto* leafOps  = static_cast<uint32_t*>(f->leafOps.get((size_t(L) + 2) * sizeof(uint32_t)));
    auto* small    = static_cast<char*>(f->small.get(256 + allKeys.size() * sizeof(KeyType)));
    int* dScalars  = reinterpret_cast<int*>(small); // [0] status, [1] numChanges, [2] leaves with op != 1, [3] new count
    auto* dKeys    = reinterpret_cast<KeyType*>(small + 256);
    auto* prefixes = static_cast<const KeyType*>(f->prefixes.ptr);
    auto* child    = static_cast<const TreeNodeIndex*>(f->childOffsets.ptr);
    auto* parents  = static_cast<const TreeNodeIndex*>(f->parents.ptr);

    SPHB200_CUDA(cudaMemsetAsync(dScalars, 0, 4 * sizeof(int), ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(dKeys, allKeys.data(), allKeys.size() * sizeof(KeyType), cudaMemcpyHostToDevice,
                                 ctx->stream));
    SPHB200_LAUNCH(ctx, essentialDecisionKernel, gridFor(nn, 256), 256, 0, prefixes, child, parents,
                   static_cast<const uint32_t*>(f->counts.ptr), static_cast<const char*>(f->macs.ptr), focusStart,
                   focusEnd, f->bucket, nn, ops);
    SPHB200_LAUNCH(ctx, enforceKeysKernel, gridFor(allKeys.size(), 64), 64, 0, dKeys, int(allKeys.size()), prefixes,
                   child, parents, ops, dScalars);
    SPHB200_LAUNCH(ctx, protectAncestorsKernel, gridFor(nn, 256), 256, 0, prefixes, parents, nn, ops, dScalars + 1);
    SPHB200_LAUNCH(ctx, gatherLeafOpsKernel, gridFor(L + 1, 256), 256, 0, ops,
                   static_cast<const TreeNodeIndex*>(f->leafToInternal.ptr), f->numInternal(), L, leafOps, dScalars + 2);
    exclusiveScan(ctx, leafOps, leafOps, size_t(L) + 1, nullptr);
    // leafOps[L] now holds the new number of leaves
    SPHB200_CUDA(cudaMemcpyAsync(dScalars + 3, leafOps + L, sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    int h[4];
    SPHB200_CUDA(cudaMemcpyAsync(h, dScalars, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    int status = h[0], numChanges = h[1], notOne = h[2], newL = h[3];

    bool converged = numChanges == 0;
    if (status == 1) converged = notOne == 0; // cancelMerge
    else if (status == 2) converged = false;  // rebalance
    f->lastStatus = status;

    int nxt         = 1 - f->cur;
    auto* newLeaves = static_cast<KeyType*>(f->leaves[nxt].get((size_t(newL) + 1) * sizeof(KeyType)));
    rebalanceTree(ctx, f->dLeaves(), leafOps, L, newLeaves);
    f->cur       = nxt;
    f->numLeaves = newL;

    if (status == 3) // failed: the keys lie more than one level below the closest leaf
    {
        converged = false;
        std::vector<KeyType> hostLeaves(size_t(newL) + 1);
        SPHB200_CUDA(cudaMemcpyAsync(hostLeaves.data(), newLeaves, hostLeaves.size() * sizeof(KeyType),
                                     cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        hostLeaves = injectKeysHost(std::move(hostLeaves), allKeys);
        nxt        = 1 - f->cur;
        auto* inj  = static_cast<KeyType*>(f->leaves[nxt].get(hostLeaves.size() * sizeof(KeyType)));
        SPHB200_CUDA(cudaMemcpyAsync(inj, hostLeaves.data(), hostLeaves.size() * sizeof(KeyType), cudaMemcpyHostToDevice,
                                     ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        f->cur       = nxt;
        f->numLeaves = int(hostLeaves.size()) - 1;
    }
    f->link();
    return converged;
}

//! @brief computeMinMacR2 for all nodes + markMacs (FocusedOctree::updateMinMac, octree_focus_mpi.hpp:408-481)
void focusUpdateMacs(sphb200_focus* f, const sphb200_box& box, KeyType focusStart, KeyType focusEnd)
{
    sphb200_ctx* ctx  = f->ctx;
    int nn            = f->numNodes();
    auto* centers     = static_cast<double*>(f->centers.get(size_t(nn) * 4 * sizeof(double)));
    auto* macs        = static_cast<char*>(f->macs.get(size_t(nn)));
    float invThetaEff = 1.0f / f->theta + 0.5f;
    CHECK_RC(sphb200_min_mac_r2(ctx, static_cast<const uint64_t*>(f->prefixes.ptr), nn, invThetaEff, &box, centers));
    CHECK_RC(sphb200_mark_macs(ctx, static_cast<const uint64_t*>(f->prefixes.ptr),
                               static_cast<const int32_t*>(f->parents.ptr), centers, nn, &box, focusStart, focusEnd,
                               macs));
    f->numMacNodes = nn;
}

//! @brief counts of all nodes from the leaf counts in f->leafCounts
void focusUpsweep(sphb200_focus* f)
{
    sphb200_ctx* ctx = f->ctx;
    int nn = f->numNodes(), L = f->numLeaves;
    auto* layout = static_cast<uint32_t*>(f->layout.get((size_t(L) + 1) * sizeof(uint32_t)));
    auto* counts = static_cast<uint32_t*>(f->counts.get(size_t(nn) * sizeof(uint32_t)));
    exclusiveScan(ctx, static_cast<const uint32_t*>(f->leafCounts.ptr), layout, size_t(L), layout + L);
    SPHB200_LAUNCH(ctx, nodeCountsFromLayoutKernel, gridFor(nn, 256), 256, 0, static_cast<const KeyType*>(f->prefixes.ptr),
                   nn, f->dLeaves(), L, layout, counts);
    f->numCountNodes = nn;
}

} // namespace

#define SPHB200_TRY try {
#define SPHB200_CATCH                                                                                                 \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return 1;                                                                                                     \
    }

// ------------------------------------------------------------------------------------------ tier 1: the operators

extern "C" int sphb200_min_mac_r2(sphb200_ctx* ctx, const uint64_t* prefixes, int numNodes, float invThetaEff,
                                  const sphb200_box* box_h, double* centers4)
{
    SPHB200_TRY
    if (numNodes > 0)
        SPHB200_LAUNCH(ctx, minMacR2Kernel, gridFor(numNodes, 256), 256, 0, prefixes, numNodes, invThetaEff, *box_h,
                       centers4);
    SPHB200_CATCH
}

extern "C" int sphb200_mark_macs(sphb200_ctx* ctx, const uint64_t* prefixes, const int32_t* parents,
                                 const double* centers4, int numNodes, const sphb200_box* box_h, uint64_t focusStart,
                                 uint64_t focusEnd, char* markings)
{
    SPHB200_TRY
    if (numNodes <= 0) return 0;
    std::vector<KeyType> codes(512);
    int numBoxes = 0;
    if (focusEnd > focusStart)
    {
        numBoxes = sphb200_span_sfc_range(focusStart, focusEnd, codes.data(), codes.size());
        if (numBoxes < 0) throw std::runtime_error(sphb200_last_error());
    }
    size_t codeBytes = ((size_t(numBoxes) + 1) * sizeof(KeyType) + 255) & ~size_t(255);
    char* tmp        = static_cast<char*>(ctx->miscTmp.get(codeBytes + size_t(std::max(numBoxes, 1)) * sizeof(GeoBox)));
    auto* dCodes     = reinterpret_cast<KeyType*>(tmp);
    auto* dBoxes     = reinterpret_cast<GeoBox*>(tmp + codeBytes);
    if (numBoxes)
    {
        SPHB200_CUDA(cudaMemcpyAsync(dCodes, codes.data(), (size_t(numBoxes) + 1) * sizeof(KeyType),
                                     cudaMemcpyHostToDevice, ctx->stream));
        SPHB200_LAUNCH(ctx, focusBoxesKernel, gridFor(numBoxes, 64), 64, 0, dCodes, numBoxes, *box_h, dBoxes);
    }
    SPHB200_LAUNCH(ctx, markMacsKernel, gridFor(numNodes, 256), 256, size_t(numBoxes) * sizeof(GeoBox), prefixes,
                   parents, centers4, numNodes, *box_h, dBoxes, numBoxes, focusStart, focusEnd, markings);
    // `codes` is pageable: cudaMemcpyAsync has staged it before returning, the vector may go out of scope
    SPHB200_CATCH
}

extern "C" int sphb200_rebalance_decision_essential(sphb200_ctx* ctx, const uint64_t* prefixes,
                                                    const int32_t* childOffsets, const int32_t* parents,
                                                    const uint32_t* counts, const char* macs, uint64_t focusStart,
                                                    uint64_t focusEnd, unsigned bucketSize, int numNodes,
                                                    int32_t* nodeOps)
{
    SPHB200_TRY
    if (numNodes > 0)
        SPHB200_LAUNCH(ctx, essentialDecisionKernel, gridFor(numNodes, 256), 256, 0, prefixes, childOffsets, parents,
                       counts, macs, focusStart, focusEnd, bucketSize, numNodes, nodeOps);
    SPHB200_CATCH
}

extern "C" int sphb200_enforce_keys(sphb200_ctx* ctx, const uint64_t* keys, int numKeys, const uint64_t* prefixes,
                                    const int32_t* childOffsets, const int32_t* parents, int32_t* nodeOps,
                                    int* status_h)
{
    SPHB200_TRY
    SPHB200_CUDA(cudaMemsetAsync(ctx->devScalars, 0, sizeof(int), ctx->stream));
    if (numKeys > 0)
        SPHB200_LAUNCH(ctx, enforceKeysKernel, gridFor(numKeys, 64), 64, 0, keys, numKeys, prefixes, childOffsets,
                       parents, nodeOps, ctx->devScalars);
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, ctx->devScalars, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    *status_h = ctx->hostScalars[0];
    SPHB200_CATCH
}

extern "C" int sphb200_protect_ancestors(sphb200_ctx* ctx, const uint64_t* prefixes, const int32_t* parents,
                                         int numNodes, int32_t* nodeOps, int* converged_h)
{
    SPHB200_TRY
    SPHB200_CUDA(cudaMemsetAsync(ctx->devScalars, 0, sizeof(int), ctx->stream));
    if (numNodes > 0)
        SPHB200_LAUNCH(ctx, protectAncestorsKernel, gridFor(numNodes, 256), 256, 0, prefixes, parents, numNodes, nodeOps,
                       ctx->devScalars);
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, ctx->devScalars, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    *converged_h = ctx->hostScalars[0] == 0;
    SPHB200_CATCH
}

// ------------------------------------------------------------------------------------------ tier 2: the focus tree

extern "C" int sphb200_focus_create(sphb200_ctx* ctx, unsigned bucketSize, float theta, sphb200_focus** out)
{
    SPHB200_TRY
    if (!(theta > 0)) throw std::runtime_error("focus_create: theta must be positive");
    std::unique_ptr<sphb200_focus> f(new sphb200_focus);
    f->ctx    = ctx;
    f->bucket = bucketSize;
    f->theta  = theta;
    // FocusedOctreeSingleNode(bucketSize, theta): the root node with counts_{bucketSize + 1}, macs_{1}
    KeyType root[2] = {0, nodeRange(0)};
    auto* leaves    = static_cast<KeyType*>(f->leaves[0].get(2 * sizeof(KeyType)));
    SPHB200_CUDA(cudaMemcpyAsync(leaves, root, sizeof(root), cudaMemcpyHostToDevice, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    f->link();
    auto* counts     = static_cast<uint32_t*>(f->counts.get(sizeof(uint32_t)));
    auto* leafCounts = static_cast<uint32_t*>(f->leafCounts.get(sizeof(uint32_t)));
    auto* macs       = static_cast<char*>(f->macs.get(1));
    SPHB200_LAUNCH(ctx, fillU32Kernel, 1, 32, 0, counts, 1, bucketSize + 1);
    SPHB200_LAUNCH(ctx, fillU32Kernel, 1, 32, 0, leafCounts, 1, bucketSize + 1);
    SPHB200_CUDA(cudaMemsetAsync(macs, 1, 1, ctx->stream));
    *out = f.release();
    SPHB200_CATCH
}

extern "C" void sphb200_focus_destroy(sphb200_focus* f) { delete f; }

extern "C" int sphb200_focus_update_tree(sphb200_focus* f, uint64_t focusStart, uint64_t focusEnd,
                                         const uint64_t* mandatoryKeys_h, int numMandatory, int* converged_h)
{
    SPHB200_TRY
    *converged_h = focusUpdateTree(f, focusStart, focusEnd, mandatoryKeys_h, numMandatory) ? 1 : 0;
    SPHB200_CATCH
}

extern "C" int sphb200_focus_update_min_mac(sphb200_focus* f, const sphb200_box* box_h, uint64_t focusStart,
                                            uint64_t focusEnd)
{
    SPHB200_TRY
    focusUpdateMacs(f, *box_h, focusStart, focusEnd);
    SPHB200_CATCH
}

extern "C" int sphb200_focus_update_counts(sphb200_focus* f, const uint64_t* keys, size_t n)
{
    SPHB200_TRY
    auto* leafCounts = static_cast<uint32_t*>(f->leafCounts.get(size_t(f->numLeaves) * sizeof(uint32_t)));
    CHECK_RC(sphb200_node_counts(f->ctx, f->dLeaves(), f->numLeaves, keys, n, leafCounts));
    focusUpsweep(f);
    SPHB200_CATCH
}

extern "C" int sphb200_focus_set_leaf_counts(sphb200_focus* f, const uint32_t* leafCounts)
{
    SPHB200_TRY
    auto* dst = static_cast<uint32_t*>(f->leafCounts.get(size_t(f->numLeaves) * sizeof(uint32_t)));
    if (dst != leafCounts)
        SPHB200_CUDA(cudaMemcpyAsync(dst, leafCounts, size_t(f->numLeaves) * sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                                     f->ctx->stream));
    focusUpsweep(f);
    SPHB200_CATCH
}

extern "C" int sphb200_focus_update(sphb200_focus* f, const sphb200_box* box_h, const uint64_t* keys, size_t n,
                                    uint64_t focusStart, uint64_t focusEnd, const uint64_t* mandatoryKeys_h,
                                    int numMandatory, int* converged_h)
{
    SPHB200_TRY
    // FocusedOctreeSingleNode::update (focus/octree_focus.hpp:236-266): tree, then MACs and counts of the new tree
    bool converged = focusUpdateTree(f, focusStart, focusEnd, mandatoryKeys_h, numMandatory);
    focusUpdateMacs(f, *box_h, focusStart, focusEnd);
    CHECK_RC(sphb200_focus_update_counts(f, keys, n));
    *converged_h = converged ? 1 : 0;
    SPHB200_CATCH
}

extern "C" int sphb200_focus_get_view(sphb200_focus* f, sphb200_focus_view* out)
{
    SPHB200_TRY
    out->numLeaves      = f->numLeaves;
    out->numNodes       = f->numNodes();
    out->lastStatus     = f->lastStatus;
    out->leaves         = f->dLeaves();
    out->leafCounts     = static_cast<const uint32_t*>(f->leafCounts.ptr);
    out->prefixes       = static_cast<const uint64_t*>(f->prefixes.ptr);
    out->childOffsets   = static_cast<const int32_t*>(f->childOffsets.ptr);
    out->parents        = static_cast<const int32_t*>(f->parents.ptr);
    out->levelRange     = static_cast<const int32_t*>(f->levelRange.ptr);
    out->internalToLeaf = static_cast<const int32_t*>(f->internalToLeaf.ptr);
    out->leafToInternal = static_cast<const int32_t*>(f->leafToInternal.ptr);
    out->counts         = f->numCountNodes == f->numNodes() ? static_cast<const uint32_t*>(f->counts.ptr) : nullptr;
    out->macs           = f->numMacNodes == f->numNodes() ? static_cast<const char*>(f->macs.ptr) : nullptr;
    out->centers        = f->numMacNodes == f->numNodes() ? static_cast<const double*>(f->centers.ptr) : nullptr;
    SPHB200_CATCH
}

// ------------------------------------------------------------------------------------------ host only: peers

namespace
{

//! @brief minimal host octree over a cornerstone leaf array (the replicated global tree has a few thousand leaves)
struct HostOctree
{
    std::vector<KeyType> start;
    std::vector<int> level, child; // child = index of the first of 8 children, 0 for leaves

    explicit HostOctree(const KeyType* leaves, int numLeaves)
    {
        start.push_back(0);
        level.push_back(0);
        child.push_back(0);
        // breadth first: a node is internal iff the leaf array has a key strictly inside it
        for (size_t i = 0; i < start.size(); ++i)
        {
            KeyType s = start[i], e = s + nodeRange(unsigned(level[i]));
            size_t a = lowerBound(leaves, size_t(numLeaves) + 1, s);
            if (leaves[a + 1] >= e) continue; // leaves[a] == s and the next key ends the node: a leaf
            child[i] = int(start.size());
            for (int o = 0; o < 8; ++o)
            {
                start.push_back(s + KeyType(o) * nodeRange(unsigned(level[i]) + 1));
                level.push_back(level[i] + 1);
                child.push_back(0);
            }
        }
    }
    bool isLeaf(int i) const { return child[i] == 0; }
    KeyType codeEnd(int i) const { return start[i] + nodeRange(unsigned(level[i])); }
    int locate(KeyType s, KeyType e) const
    {
        int i = 0;
        while (!(start[i] == s && codeEnd(i) == e))
        {
            if (isLeaf(i)) throw std::runtime_error("find_peers_mac: the assignment does not follow the tree");
            int l = level[i] + 1;
            i     = child[i] + int(octalDigit(s, unsigned(l)));
        }
        return i;
    }
};

//! @brief minDistance(X, bCenter, bSize, box) (traversal/boxoverlap.hpp:207-217), squared norm
double pointBoxDistSq(const double* X, const GeoBox& b, const sphb200_box& box)
{
    double r2 = 0;
    for (int d = 0; d < 3; ++d)
    {
        double dX = b.c[d] - X[d];
        if (box.bc[d] == SPHB200_PERIODIC) dX -= box.len[d] * std::rint(dX * box.inv[d]);
        dX = std::fabs(dX);
        dX -= b.s[d];
        dX += std::fabs(dX);
        dX *= 0.5;
        r2 += dX * dX;
    }
    return r2;
}

//! @brief minVecMacMutual (traversal/macs.hpp:166-190)
bool minVecMacMutual(const GeoBox& A, const GeoBox& B, const sphb200_box& box, float invThetaEff)
{
    double macB = std::max(std::max(B.s[0], B.s[1]), B.s[2]) * 2 * invThetaEff;
    bool passA  = pointBoxDistSq(B.c, A, box) > macB * macB;
    double macA = std::max(std::max(A.s[0], A.s[1]), A.s[2]) * 2 * invThetaEff;
    bool passB  = pointBoxDistSq(A.c, B, box) > macA * macA;
    return passA && passB;
}

} // namespace

extern "C" int sphb200_find_peers_mac(int myRank, int numRanks, const uint64_t* bounds_h, const uint64_t* leaves_h,
                                      int numLeaves, const sphb200_box* box_h, float invThetaEff, int* peers_h)
{
    SPHB200_TRY
    if (myRank < 0 || myRank >= numRanks) throw std::runtime_error("find_peers_mac: rank out of range");
    const sphb200_box& box = *box_h;
    KeyType domainStart = bounds_h[myRank], domainEnd = bounds_h[myRank + 1];
    for (int r = 0; r < numRanks; ++r)
        peers_h[r] = 0;
    if (domainEnd <= domainStart) return 0;
    HostOctree tree(leaves_h, numLeaves);
    const int numNodes = int(tree.start.size());
    // geometry and the ranks each node touches, once per node (the reference recomputes them for every pair)
    std::vector<GeoBox> geo(numNodes);
    std::vector<int> rankLo(numNodes), rankHi(numNodes);
    auto findRank = [&](KeyType key) // SfcAssignment::findRank (domaindecomp.hpp:104-108)
    { return int(std::upper_bound(bounds_h, bounds_h + numRanks + 1, key) - bounds_h) - 1; };
    for (int i = 0; i < numNodes; ++i)
    {
        geo[i]    = nodeGeometry(tree.start[i], unsigned(tree.level[i]), box);
        rankLo[i] = findRank(tree.start[i]);
        rankHi[i] = std::min(findRank(tree.codeEnd(i) - 1), numRanks - 1);
    }
    // a pair whose source only touches ranks that are peers already cannot change the result: not followed
    auto allMarked = [&](int b)
    {
        for (int r = rankLo[b]; r <= rankHi[b]; ++r)
            if (r != myRank && !peers_h[r]) return false;
        return true;
    };

    // crossFocusPairs (peers.hpp:72-86): continue while a touches the own segment, b is not inside it and the pair
    // fails the mutual min-distance MAC
    auto continuation = [&](int a, int b)
    {
        bool aFocusOverlap = domainEnd > tree.start[a] && tree.codeEnd(a) > domainStart;
        bool bInFocus      = !(tree.start[b] < domainStart || tree.codeEnd(b) > domainEnd);
        if (!aFocusOverlap || bInFocus) return false;
        if (allMarked(b)) return false;
        return !minVecMacMutual(geo[a], geo[b], box, invThetaEff);
    };
    auto p2p = [&](int b) { peers_h[rankLo[b]] = 1; };

    std::vector<KeyType> span(512);
    int numSpan = sphb200_span_sfc_range(domainStart, domainEnd, span.data(), span.size());
    if (numSpan < 0) throw std::runtime_error(sphb200_last_error());
    std::vector<std::pair<int, int>> stack;
    for (int k = 0; k < numSpan; ++k)
    {
        // dualTraversal (traversal/traversal.hpp:135-188) from (spanning node, root)
        int a0 = tree.locate(span[k], span[k + 1]);
        if (tree.isLeaf(a0) && tree.isLeaf(0))
        {
            p2p(0);
            continue;
        }
        stack.assign(1, {a0, 0});
        while (!stack.empty())
        {
            auto [target, source] = stack.back();
            stack.pop_back();
            bool splitTarget =
                (tree.level[target] < tree.level[source] && !tree.isLeaf(target)) || tree.isLeaf(source);
            int numChildren = splitTarget ? (tree.isLeaf(target) ? 0 : 8) : (tree.isLeaf(source) ? 0 : 8);
            for (int o = 0; o < numChildren; ++o)
            {
                int a = splitTarget ? tree.child[target] + o : target;
                int b = splitTarget ? source : tree.child[source] + o;
                if (!continuation(a, b)) continue;
                if (tree.isLeaf(a) && tree.isLeaf(b)) p2p(b);
                else stack.push_back({a, b});
            }
        }
    }
    SPHB200_CATCH
}
