/*! @file forwarding definitions for the reference's `extern` focus-tree GPU templates (operator level, tier 1)
 *
 * The reference declares rebalanceDecisionEssentialGpu, protectAncestorsGpu and enforceKeysGpu in
 * domain/include/cstone/focus/rebalance_gpu.h:37-60 and defines + explicitly instantiates them in
 * focus/rebalance_gpu.cu; CombinedUpdate::updateFocusGpu (focus/octree_focus.hpp:158-213) calls them on every
 * Domain::sync.  This translation unit is compiled against those SAME declarations and defines the 64-bit
 * instantiations by forwarding to the C ABI of libsphexa_b200.so (csrc/focus.cu).  Linked in place of the reference's
 * rebalance_gpu object it makes the reference's own FocusedOctree -- inside its own cstone::Domain, in its own GPU
 * application -- take its split / merge decisions from this repository's kernels (`sphexa-cuda-b200`,
 * build_sphexa_cuda.sh of the test infrastructure; checked by tests/test_gpu_dropin.py: identical `Focus Tree Nodes` of every iteration).
 *
 * markMacsGpu (traversal/collisions_gpu.h:70) is not forwarded: its signature carries neither the number of nodes nor
 * the parents array that sphb200_mark_macs works from.
 */
#include <cstdint>
#include <stdexcept>
#include <string>

#include "cstone/focus/rebalance.hpp"
#include "cstone/focus/rebalance_gpu.h"

#include "sphexa_b200.h"

namespace
{

void check(int rc)
{
    if (rc != 0) { throw std::runtime_error(std::string("sph-exa_b200: ") + sphb200_last_error()); }
}

//! @brief a context on the legacy default stream, like everything the reference launches
sphb200_ctx* context()
{
    static sphb200_ctx* ctx = nullptr;
    if (!ctx) { check(sphb200_create(0, nullptr, &ctx)); }
    return ctx;
}

} // namespace

namespace cstone
{

template<class KeyType>
void rebalanceDecisionEssentialGpu(const KeyType* prefixes,
                                   const TreeNodeIndex* childOffsets,
                                   const TreeNodeIndex* parents,
                                   const unsigned* counts,
                                   const char* macs,
                                   KeyType focusStart,
                                   KeyType focusEnd,
                                   unsigned bucketSize,
                                   TreeNodeIndex* nodeOps,
                                   TreeNodeIndex numNodes)
{
    static_assert(sizeof(KeyType) == 8, "libsphexa_b200 works on 64-bit keys");
    check(sphb200_rebalance_decision_essential(context(), prefixes, childOffsets, parents, counts, macs, focusStart,
                                               focusEnd, bucketSize, numNodes, nodeOps));
}

template void rebalanceDecisionEssentialGpu(const uint64_t*, const TreeNodeIndex*, const TreeNodeIndex*,
                                            const unsigned*, const char*, uint64_t, uint64_t, unsigned, TreeNodeIndex*,
                                            TreeNodeIndex);

template<class KeyType>
bool protectAncestorsGpu(const KeyType* prefixes, const TreeNodeIndex* parents, TreeNodeIndex* nodeOps,
                         TreeNodeIndex numNodes)
{
    static_assert(sizeof(KeyType) == 8, "libsphexa_b200 works on 64-bit keys");
    int converged = 0;
    check(sphb200_protect_ancestors(context(), prefixes, parents, numNodes, nodeOps, &converged));
    return converged != 0;
}

template bool protectAncestorsGpu(const uint64_t*, const TreeNodeIndex*, TreeNodeIndex*, TreeNodeIndex);

template<class KeyType>
ResolutionStatus enforceKeysGpu(const KeyType* forcedKeys,
                                TreeNodeIndex numForcedKeys,
                                const KeyType* nodeKeys,
                                const TreeNodeIndex* childOffsets,
                                const TreeNodeIndex* parents,
                                TreeNodeIndex* nodeOps)
{
    static_assert(sizeof(KeyType) == 8, "libsphexa_b200 works on 64-bit keys");
    int status = 0;
    check(sphb200_enforce_keys(context(), forcedKeys, numForcedKeys, nodeKeys, childOffsets, parents, nodeOps,
                               &status));
    return ResolutionStatus(status);
}

template ResolutionStatus enforceKeysGpu(const uint64_t*, TreeNodeIndex, const uint64_t*, const TreeNodeIndex*,
                                         const TreeNodeIndex*, TreeNodeIndex*);

} // namespace cstone
