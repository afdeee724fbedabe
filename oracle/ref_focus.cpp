/* TEST INFRASTRUCTURE ONLY -- C veneer over the reference's focus-tree machinery.
 *
 * Compiled by oracle/Makefile together with ref_wrap.cpp into oracle/_ref/libsphexa_ref.so from the reference's
 * headers where they lie.  Every function only CALLS reference code:
 *
 *   computeMinMacR2                traversal/macs.hpp:58-71
 *   markMacs                       traversal/macs.hpp:240-264
 *   rebalanceDecisionEssential     focus/rebalance.hpp:144-160
 *   enforceKeys                    focus/rebalance.hpp:245-258
 *   protectAncestors               focus/rebalance.hpp:162-176
 *   FocusedOctreeSingleNode        focus/octree_focus.hpp:219-288 (CombinedUpdate::updateFocus :81-139)
 *   findPeersMac                   traversal/peers.hpp:63-120
 */
#include <cstdint>
#include <vector>

#include "cstone/domain/domaindecomp.hpp"
#include "cstone/focus/octree_focus.hpp"
#include "cstone/traversal/macs.hpp"
#include "cstone/traversal/peers.hpp"
#include "cstone/tree/octree.hpp"

using KeyType = uint64_t;

static cstone::Box<double> focusBox(const double* lim, const int* bc)
{
    return cstone::Box<double>(lim[0], lim[1], lim[2], lim[3], lim[4], lim[5], cstone::BoundaryType(bc[0]),
                               cstone::BoundaryType(bc[1]), cstone::BoundaryType(bc[2]));
}

extern "C"
{

//! centers4[4 i .. 4 i + 3] = geometric center and squared acceptance radius of node i
void ref_min_mac_r2(const uint64_t* prefixes, int numNodes, float invThetaEff, const double* lim, const int* bc,
                    double* centers4)
{
    auto box = focusBox(lim, bc);
    auto* c  = reinterpret_cast<cstone::Vec4<double>*>(centers4);
    for (int i = 0; i < numNodes; ++i)
        c[i] = cstone::computeMinMacR2(prefixes[i], invThetaEff, box);
}

void ref_mark_macs(const uint64_t* prefixes, const int* childOffsets, const int* parents, const int* levelRange,
                   const int* internalToLeaf, const int* leafToInternal, int numLeaves, int numInternal,
                   const double* centers4, const double* lim, const int* bc, uint64_t focusStart, uint64_t focusEnd,
                   char* markings)
{
    // markMacs takes the mutable view type (it only reads the tree)
    cstone::OctreeView<KeyType> view{numLeaves,
                                     numInternal,
                                     numLeaves + numInternal,
                                     const_cast<KeyType*>(prefixes),
                                     const_cast<int*>(childOffsets),
                                     const_cast<int*>(parents),
                                     const_cast<int*>(levelRange),
                                     const_cast<int*>(internalToLeaf),
                                     const_cast<int*>(leafToInternal)};
    cstone::markMacs(view, reinterpret_cast<const cstone::Vec4<double>*>(centers4), focusBox(lim, bc), focusStart,
                     focusEnd, markings);
}

void ref_rebalance_decision_essential(const uint64_t* prefixes, int numNodes, const int* childOffsets,
                                      const int* parents, const unsigned* counts, const char* macs,
                                      uint64_t focusStart, uint64_t focusEnd, unsigned bucketSize, int* nodeOps)
{
    cstone::rebalanceDecisionEssential<KeyType>({prefixes, size_t(numNodes)}, childOffsets, parents, counts, macs,
                                                focusStart, focusEnd, bucketSize, nodeOps);
}

//! returns the ResolutionStatus as an int (0 converged, 1 cancelMerge, 2 rebalance, 3 failed)
int ref_enforce_keys(const uint64_t* keys, int numKeys, const uint64_t* prefixes, const int* childOffsets,
                     const int* parents, int* nodeOps)
{
    return int(cstone::enforceKeys<KeyType>({keys, size_t(numKeys)}, prefixes, childOffsets, parents, nodeOps));
}

int ref_protect_ancestors(const uint64_t* prefixes, int numNodes, const int* parents, int* nodeOps)
{
    return cstone::protectAncestors<KeyType>({prefixes, size_t(numNodes)}, parents, nodeOps) ? 1 : 0;
}

void* ref_focus_create(unsigned bucketSize, float theta)
{
    return new cstone::FocusedOctreeSingleNode<KeyType>(bucketSize, theta);
}
void ref_focus_destroy(void* f) { delete static_cast<cstone::FocusedOctreeSingleNode<KeyType>*>(f); }

//! one FocusedOctreeSingleNode::update; returns 1 if the tree structure did not change
int ref_focus_update(void* f, const double* lim, const int* bc, const uint64_t* keys, size_t n, uint64_t focusStart,
                     uint64_t focusEnd, const uint64_t* mandatoryKeys, int numMandatory)
{
    auto* tree = static_cast<cstone::FocusedOctreeSingleNode<KeyType>*>(f);
    return tree->update(focusBox(lim, bc), {keys, n}, focusStart, focusEnd, {mandatoryKeys, size_t(numMandatory)}) ? 1
                                                                                                                    : 0;
}
int ref_focus_num_leaves(void* f)
{
    return int(static_cast<cstone::FocusedOctreeSingleNode<KeyType>*>(f)->treeLeaves().size()) - 1;
}
void ref_focus_get(void* f, uint64_t* leaves, unsigned* counts)
{
    auto* tree = static_cast<cstone::FocusedOctreeSingleNode<KeyType>*>(f);
    auto l     = tree->treeLeaves();
    auto c     = tree->leafCounts();
    std::copy(l.begin(), l.end(), leaves);
    std::copy(c.begin(), c.end(), counts);
}

//! peers[r] = 1 for every rank findPeersMac lists for myRank on the global tree `leaves`
void ref_find_peers_mac(int myRank, int numRanks, const uint64_t* bounds, const uint64_t* leaves, int numLeaves,
                        const double* lim, const int* bc, float invThetaEff, int* peers)
{
    cstone::SfcAssignment<KeyType> assignment(numRanks);
    for (int r = 0; r <= numRanks; ++r)
        assignment.set(r, bounds[r], 0);
    cstone::Octree<KeyType> tree;
    tree.update(leaves, numLeaves);
    auto list = cstone::findPeersMac(myRank, assignment, tree, focusBox(lim, bc), invThetaEff);
    for (int r = 0; r < numRanks; ++r)
        peers[r] = 0;
    for (int p : list)
        peers[p] = 1;
}

} // extern "C"
