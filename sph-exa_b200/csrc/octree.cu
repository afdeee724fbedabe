// Cornerstone octree: leaf-array node counts, rebalance (computeOctree/updateOctree) and the
// internal linked octree, fully device-resident (the reference ping-pongs through the host and
// thrust, domain/include/cstone/tree/{csarray_gpu.cu,octree_gpu.cu,update_gpu.cuh}).
#include "common.cuh"
#include "primitives.cuh"

namespace sphb200
{

// ---------------------------------------------------------------- node counts
/*! @brief particles per leaf: two lower_bounds in the sorted keys
 *  (calculateNodeCount, tree/csarray.hpp:94-103; result identical to the guess-accelerated updateNodeCount) */
__global__ void nodeCountsKernel(const KeyType* __restrict__ leaves, int numLeaves, const KeyType* __restrict__ keys,
                                 size_t n, uint32_t maxCount, uint32_t* __restrict__ counts)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numLeaves) return;
    size_t a  = lowerBound(keys, n, leaves[i]);
    size_t b  = lowerBound(keys, n, leaves[i + 1]);
    size_t c  = b - a;
    counts[i] = uint32_t(c < size_t(maxCount) ? c : size_t(maxCount));
}

// ---------------------------------------------------------------- rebalance
/*! @brief split/merge/keep decision per leaf
 *
 * siblingAndLevel + calculateNodeOp (tree/csarray.hpp:273-312): 0 = merge (siblings 1..7 of a group whose
 * summed count <= bucket), 8/64/512/4096 = split by 1..4 levels, 1 = keep.  ops has numLeaves+1 entries, the
 * last one is set to 0 so that an exclusive scan yields the new leaf offsets.
 */
__global__ void rebalanceDecisionKernel(const KeyType* __restrict__ leaves, const uint32_t* __restrict__ counts,
                                        int numLeaves, unsigned bucket, uint32_t* __restrict__ ops,
                                        int* __restrict__ changed, const KeyType* __restrict__ guide = nullptr,
                                        int numGuide = 0)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > numLeaves) return;
    if (i == numLeaves)
    {
        ops[i] = 0;
        return;
    }
    KeyType node   = leaves[i];
    KeyType range  = leaves[i + 1] - node;
    unsigned level = treeLevel(range);

    int siblingIdx = -1;
    if (level > 0)
    {
        siblingIdx = int(octalDigit(node, level));
        int g0     = i - siblingIdx;
        // all 8 siblings present at this level <=> leaves[g0 + 8] == leaves[g0] + nodeRange(level - 1)
        bool siblings = (g0 + 8 <= numLeaves) && (leaves[g0 + 8] == leaves[g0] + nodeRange(level - 1));
        if (!siblings) siblingIdx = -1;
    }

    // guide != NULL: the tree must resolve the leaves of a coarser "guide" tree (the replicated global tree of a
    // multi-rank run, whose leaf keys are the rank boundaries): a leaf that spans more than one guide leaf is split
    // whatever it holds, and siblings only merge into a parent that still lies inside one guide leaf.  This is what
    // cstone's focus tree obtains with enforceKeys (focus/rebalance.hpp:189-258) plus exact foreign counts.
    bool forceSplit = false;
    if (guide)
    {
        int gi          = int(upperBound(guide, size_t(numGuide) + 1, node)) - 1;
        KeyType gEnd    = guide[gi + 1];
        unsigned gLevel = treeLevel(gEnd - guide[gi]);
        if (leaves[i + 1] > gEnd) forceSplit = true;
        if (level <= gLevel) siblingIdx = -1; // the parent would reach beyond the guide leaf
    }

    uint32_t op = 1;
    bool done   = false;
    if (forceSplit)
    {
        op   = 8;
        done = true;
        changed[2] = 1;
    }
    if (!done && siblingIdx > 0)
    {
        const uint32_t* g = counts + i - siblingIdx;
        size_t parentCount = size_t(g[0]) + g[1] + g[2] + g[3] + g[4] + g[5] + g[6] + g[7];
        if (parentCount <= size_t(bucket))
        {
            op   = 0;
            done = true;
        }
    }
    if (!done)
    {
        uint32_t c = counts[i];
        if (c > bucket * 512u && level + 3 < kMaxTreeLevel) op = 4096;
        else if (c > bucket * 64u && level + 2 < kMaxTreeLevel) op = 512;
        else if (c > bucket * 8u && level + 1 < kMaxTreeLevel) op = 64;
        else if (c > bucket && level < kMaxTreeLevel) op = 8;
    }
    ops[i] = op;
    if (op != 1) *changed = 1; // benign race, every writer stores the same value
}

//! @brief emit the new leaves of old leaf i at its scanned offset (processNode, tree/csarray.hpp:362-388)
__global__ void processNodesKernel(const KeyType* __restrict__ oldLeaves, const uint32_t* __restrict__ opScan,
                                   int numLeaves, KeyType* __restrict__ newLeaves)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > numLeaves) return;
    if (i == numLeaves)
    {
        newLeaves[opScan[numLeaves]] = oldLeaves[numLeaves]; // 2^63 terminator
        return;
    }
    KeyType node    = oldLeaves[i];
    KeyType range   = oldLeaves[i + 1] - node;
    uint32_t offset = opScan[i];
    uint32_t op     = opScan[i + 1] - offset;
    if (op == 1) { newLeaves[offset] = node; }
    else if (op > 1)
    {
        KeyType childRange = range / op; // op = 8^levelDiff
        for (uint32_t s = 0; s < op; ++s)
            newLeaves[offset + s] = node + s * childRange;
    }
}

// ---------------------------------------------------------------- internal octree
__device__ __forceinline__ int digitWeight(int digit)
{
    int fourGeqMask = -int(digit >= 4);
    return ((7 - digit) & fourGeqMask) - (digit & ~fourGeqMask);
}

//! @brief binary-tree -> octree index offset (binaryKeyWeight, tree/octree.hpp:71-81)
__device__ __forceinline__ int binaryKeyWeight(KeyType key, unsigned level)
{
    int ret = 0;
    for (unsigned l = 1; l <= level + 1; ++l)
        ret += digitWeight(int(octalDigit(key, l)));
    return ret;
}

/*! @brief Warren-Salmon prefixes of all nodes, unsorted: internal nodes first (placed through the
 *  binary/octree index correspondence), leaves behind (createUnsortedLayout, tree/octree.hpp:94-117) */
__global__ void unsortedLayoutKernel(const KeyType* __restrict__ leaves, int numInternal, int numLeaves,
                                     KeyType* __restrict__ prefixes)
{
    int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= numLeaves) return;
    KeyType key                 = leaves[tid];
    KeyType next                = leaves[tid + 1];
    unsigned level              = treeLevel(next - key);
    prefixes[tid + numInternal] = encodePlaceholderBit(key, 3 * int(level));

    int prefixLength = commonPrefix(key, next);
    if (prefixLength % 3 == 0 && tid < numLeaves - 1)
    {
        int octIndex       = (tid + binaryKeyWeight(key, unsigned(prefixLength / 3))) / 7;
        prefixes[octIndex] = encodePlaceholderBit(key, prefixLength);
    }
}

//! @brief order holds the unsorted index of each sorted node: build both maps (invertOrder, octree_gpu.cu:142-151)
__global__ void invertOrderKernel(const uint32_t* __restrict__ order, int numNodes, int numInternal,
                                  TreeNodeIndex* __restrict__ internalToLeaf, TreeNodeIndex* __restrict__ leafToInternal)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    int unsortedIdx             = int(order[i]);
    leafToInternal[unsortedIdx] = i;
    internalToLeaf[i]           = unsortedIdx - numInternal;
}

//! @brief first node index of each level (getLevelRange, tree/octree.hpp:169-178)
__global__ void levelRangeKernel(const KeyType* __restrict__ prefixes, int numNodes, TreeNodeIndex* __restrict__ levelRange)
{
    int level = threadIdx.x;
    if (level <= kMaxTreeLevel)
        levelRange[level] = TreeNodeIndex(lowerBound(prefixes, size_t(numNodes), encodePlaceholderBit(KeyType(0), 3 * level)));
    if (level == kMaxTreeLevel + 1) levelRange[level] = numNodes;
}

//! @brief child/parent links (linkTree, tree/octree.hpp:131-166)
__global__ void linkTreeKernel(const KeyType* __restrict__ prefixes, int numInternal,
                               const TreeNodeIndex* __restrict__ leafToInternal,
                               const TreeNodeIndex* __restrict__ levelRange, TreeNodeIndex* __restrict__ childOffsets,
                               TreeNodeIndex* __restrict__ parents)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numInternal) return;
    TreeNodeIndex idxA    = leafToInternal[i];
    KeyType prefix        = prefixes[idxA];
    KeyType nodeKey       = decodePlaceholderBit(prefix);
    unsigned prefixLength = decodePrefixLength(prefix);
    unsigned level        = prefixLength / 3;
    KeyType childPrefix   = encodePlaceholderBit(nodeKey, int(prefixLength) + 3);

    TreeNodeIndex s = levelRange[level + 1];
    TreeNodeIndex e = levelRange[level + 2];
    TreeNodeIndex childIdx = s + TreeNodeIndex(lowerBound(prefixes + s, size_t(e - s), childPrefix));
    if (childIdx != e && prefixes[childIdx] == childPrefix)
    {
        childOffsets[idxA]          = childIdx;
        parents[(childIdx - 1) / 8] = idxA;
    }
}

__global__ void initRootKernel(KeyType* leaves, uint32_t* counts, uint32_t n)
{
    leaves[0] = 0;
    leaves[1] = nodeRange(0);
    counts[0] = n;
}

static void updateOctreeImpl(sphb200_ctx* ctx, const KeyType* keys, size_t n, unsigned bucket, KeyType* leaves,
                             uint32_t* counts, int capacity, int* numLeaves, int* converged,
                             const KeyType* guide = nullptr, int numGuide = 0, int* forced = nullptr)
{
    int L = *numLeaves;
    // scratch: ops[L+1] | newLeaves (up to capacity+1)
    size_t opsBytes = ((size_t(L) + 2) * sizeof(uint32_t) + 255) & ~size_t(255);
    char* tmp       = static_cast<char*>(ctx->treeTmp.get(opsBytes + (size_t(capacity) + 2) * sizeof(KeyType)));
    uint32_t* ops   = reinterpret_cast<uint32_t*>(tmp);
    KeyType* newLeaves = reinterpret_cast<KeyType*>(tmp + opsBytes);

    int* dFlag      = ctx->devScalars;
    uint32_t* dTotal = reinterpret_cast<uint32_t*>(ctx->devScalars + 1);
    SPHB200_CUDA(cudaMemsetAsync(dFlag, 0, 3 * sizeof(int), ctx->stream));
    SPHB200_LAUNCH(ctx, rebalanceDecisionKernel, gridFor(L + 1, 256), 256, 0, leaves, counts, L, bucket, ops, dFlag,
                   guide, numGuide);
    exclusiveScan(ctx, ops, ops, size_t(L) + 1, nullptr);
    // ops[L] now holds the new leaf count; fetch it together with the change flag
    SPHB200_CUDA(cudaMemcpyAsync(dTotal, ops + L, sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, ctx->devScalars, 3 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    int changed = ctx->hostScalars[0];
    int newL    = ctx->hostScalars[1];
    if (forced) *forced = ctx->hostScalars[2];
    if (newL > capacity)
    {
        // nothing has been modified yet: report the required size so that the caller can retry with larger arrays
        *numLeaves = newL;
        throw CapacityError("update_octree: leaf capacity exceeded (" + std::to_string(newL) + " > " +
                            std::to_string(capacity) + ")");
    }
    if (changed)
    {
        SPHB200_LAUNCH(ctx, processNodesKernel, gridFor(L + 1, 256), 256, 0, leaves, ops, L, newLeaves);
        SPHB200_CUDA(cudaMemcpyAsync(leaves, newLeaves, (size_t(newL) + 1) * sizeof(KeyType), cudaMemcpyDeviceToDevice, ctx->stream));
    }
    SPHB200_LAUNCH(ctx, nodeCountsKernel, gridFor(newL, 256), 256, 0, leaves, newL, keys, n, 0xffffffffu, counts);
    *numLeaves = newL;
    *converged = !changed;
}

void updateOctreeGuided(sphb200_ctx* ctx, const KeyType* keys, size_t n, unsigned bucket, KeyType* leaves,
                        uint32_t* counts, int capacity, int* numLeaves, int* converged, const KeyType* guide,
                        int numGuide, int* forced)
{
    updateOctreeImpl(ctx, keys, n, bucket, leaves, counts, capacity, numLeaves, converged, guide, numGuide, forced);
}

void rebalanceTree(sphb200_ctx* ctx, const KeyType* oldLeaves, const uint32_t* opScan, int numLeaves, KeyType* newLeaves)
{
    SPHB200_LAUNCH(ctx, processNodesKernel, gridFor(numLeaves + 1, 256), 256, 0, oldLeaves, opScan, numLeaves, newLeaves);
}

void initRootTree(sphb200_ctx* ctx, KeyType* leaves, uint32_t* counts, uint32_t n)
{
    SPHB200_LAUNCH(ctx, initRootKernel, 1, 1, 0, leaves, counts, n);
}

} // namespace sphb200

using namespace sphb200;

#define SPHB200_TRY try {
#define SPHB200_CATCH                                                                                                 \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return 1;                                                                                                     \
    }

extern "C" int sphb200_node_counts(sphb200_ctx* ctx, const uint64_t* leaves, int numLeaves, const uint64_t* keys,
                                   size_t n, uint32_t* counts)
{
    SPHB200_TRY
    if (numLeaves > 0)
        SPHB200_LAUNCH(ctx, nodeCountsKernel, gridFor(numLeaves, 256), 256, 0, leaves, numLeaves, keys, n, 0xffffffffu,
                       counts);
    SPHB200_CATCH
}

extern "C" int sphb200_rebalance_decision(sphb200_ctx* ctx, const uint64_t* leaves, const uint32_t* counts,
                                          int numLeaves, unsigned bucket, int32_t* ops, int* converged_h)
{
    SPHB200_TRY
    uint32_t* tmp = static_cast<uint32_t*>(ctx->treeTmp.get((size_t(numLeaves) + 2) * sizeof(uint32_t)));
    SPHB200_CUDA(cudaMemsetAsync(ctx->devScalars, 0, sizeof(int), ctx->stream));
    SPHB200_LAUNCH(ctx, rebalanceDecisionKernel, gridFor(numLeaves + 1, 256), 256, 0, leaves, counts, numLeaves, bucket,
                   tmp, ctx->devScalars);
    SPHB200_CUDA(cudaMemcpyAsync(ops, tmp, size_t(numLeaves) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, ctx->devScalars, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    *converged_h = !ctx->hostScalars[0];
    SPHB200_CATCH
}

extern "C" int sphb200_update_octree(sphb200_ctx* ctx, const uint64_t* keys, size_t n, unsigned bucket,
                                     uint64_t* leaves, uint32_t* counts, int capacity, int* numLeaves_h,
                                     int* converged_h)
{
    try
    {
        updateOctreeImpl(ctx, keys, n, bucket, leaves, counts, capacity, numLeaves_h, converged_h);
        return 0;
    }
    catch (CapacityError& e)
    {
        setError(e.what());
        return SPHB200_ERR_CAPACITY; // *numLeaves_h = required number of leaves, inputs untouched
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_compute_octree(sphb200_ctx* ctx, const uint64_t* keys, size_t n, unsigned bucket,
                                      uint64_t* leaves, uint32_t* counts, int capacity, int* numLeaves_h,
                                      int* iterations_h)
{
    SPHB200_TRY
    if (capacity < 1) throw std::runtime_error("compute_octree: capacity < 1");
    SPHB200_LAUNCH(ctx, initRootKernel, 1, 1, 0, leaves, counts, uint32_t(n));
    int L = 1, conv = 0, it = 0;
    while (!conv)
    {
        updateOctreeImpl(ctx, keys, n, bucket, leaves, counts, capacity, &L, &conv);
        ++it;
        if (it > 64) throw std::runtime_error("compute_octree: no convergence");
    }
    *numLeaves_h = L;
    if (iterations_h) *iterations_h = it;
    SPHB200_CATCH
}

extern "C" int sphb200_build_octree(sphb200_ctx* ctx, const uint64_t* leaves, int numLeaves, uint64_t* prefixes,
                                    int32_t* childOffsets, int32_t* parents, int32_t* levelRange,
                                    int32_t* internalToLeaf, int32_t* leafToInternal)
{
    SPHB200_TRY
    int numInternal = (numLeaves - 1) / 7;
    int numNodes    = numLeaves + numInternal;
    uint32_t* order = static_cast<uint32_t*>(ctx->treeTmp.get(size_t(numNodes) * sizeof(uint32_t)));
    SPHB200_LAUNCH(ctx, unsortedLayoutKernel, gridFor(numLeaves, 256), 256, 0, leaves, numInternal, numLeaves, prefixes);
    // Warren-Salmon keys sort by (level, key); the placeholder bit sits at 3*level <= 63
    radixSortPairs(ctx, prefixes, order, size_t(numNodes), true, 0, 64);
    SPHB200_LAUNCH(ctx, invertOrderKernel, gridFor(numNodes, 256), 256, 0, order, numNodes, numInternal, internalToLeaf,
                   leafToInternal);
    SPHB200_LAUNCH(ctx, levelRangeKernel, 1, 32, 0, prefixes, numNodes, levelRange);
    SPHB200_CUDA(cudaMemsetAsync(childOffsets, 0, (size_t(numNodes) + 1) * sizeof(int32_t), ctx->stream));
    if (numInternal > 0)
        SPHB200_LAUNCH(ctx, linkTreeKernel, gridFor(numInternal, 256), 256, 0, prefixes, numInternal, leafToInternal,
                       levelRange, childOffsets, parents);
    SPHB200_CATCH
}

extern "C" int sphb200_layout(sphb200_ctx* ctx, const uint32_t* counts, int numLeaves, uint32_t* layout)
{
    SPHB200_TRY
    // scan numLeaves counts; layout[numLeaves] receives the total
    exclusiveScan(ctx, counts, layout, size_t(numLeaves), layout + numLeaves);
    SPHB200_CATCH
}
