/*! @file forwarding definitions for GPU functions of the reference's Cornerstone domain layer (operator level, tier 1)
 *
 * cstone::Domain::sync calls these `extern` templates on device arrays (domain/assignment_gpu.cuh, domain/domain.hpp,
 * focus/octree_focus_mpi.hpp, halos/halos.hpp); the reference defines them in sfc/sfc_gpu.cu, tree/octree_gpu.cu,
 * focus/source_center_gpu.cu, traversal/collisions_gpu.cu and primitives/primitives_gpu.cu.  Here the instantiations
 * the SPH front-end uses are defined as explicit SPECIALIZATIONS that forward to the C ABI of libsphexa_b200.so.  A
 * specialization is an ordinary (strong) definition, the reference's explicit instantiations are weak symbols, so
 * linking this object into the reference's GPU application replaces exactly these functions and leaves the other
 * instantiations of the same templates (32-bit keys, float coordinates: gravity) with the reference
 * (`sphexa-cuda-b200`, build_sphexa_cuda.sh of the test infrastructure; tests/test_gpu_dropin.py).
 *
 *   computeSfcKeysGpu<HilbertKey<uint64_t>, double>    sfc/sfc_gpu.h:39            -> sphb200_sfc_keys
 *   buildOctreeGpu<uint64_t>                           tree/octree_gpu.h:50        -> sphb200_build_octree
 *   computeGeoCentersGpu<uint64_t, double>             focus/source_center_gpu.h:75 -> sphb200_node_centers
 *   findHalosGpu<uint64_t, float, double>              traversal/collisions_gpu.h:59 -> sphb200_find_halos
 *   sortByKeyGpu<uint64_t, unsigned> (with buffers)     primitives/primitives_gpu.h:96 -> sphb200_sort_pairs
 *   gatherGpu<double | float | unsigned, unsigned>     primitives/primitives_gpu.h:46 -> sphb200_gather
 */
#include <cstdint>
#include <stdexcept>
#include <string>

#include "cstone/sfc/box.hpp"
#include "cstone/sfc/sfc_gpu.h"
#include "cstone/focus/source_center.hpp"
#include "cstone/focus/source_center_gpu.h"
#include "cstone/primitives/primitives_gpu.h"
#include "cstone/traversal/collisions_gpu.h"
#include "cstone/tree/octree.hpp"
#include "cstone/tree/octree_gpu.h"

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

sphb200_box toBox(const cstone::Box<double>& box)
{
    const double lim[6] = {box.xmin(), box.xmax(), box.ymin(), box.ymax(), box.zmin(), box.zmax()};
    const int bc[3]     = {int(box.boundaryX()), int(box.boundaryY()), int(box.boundaryZ())};
    sphb200_box b;
    sphb200_make_box(lim, bc, &b);
    return b;
}

template<class T>
void gatherOne(const unsigned* ordering, size_t numElements, const T* src, T* buffer)
{
    if (numElements == 0) { return; }
    const void* s[1] = {src};
    void* d[1]       = {buffer};
    const int eb[1]  = {int(sizeof(T))};
    check(sphb200_gather(context(), ordering, numElements, 1, s, d, eb));
}

} // namespace

namespace cstone
{

template<>
void computeSfcKeysGpu(const double* x, const double* y, const double* z, HilbertKey<uint64_t>* keys, size_t numKeys,
                       const Box<double>& box)
{
    static_assert(sizeof(HilbertKey<uint64_t>) == sizeof(uint64_t));
    sphb200_box b = toBox(box);
    check(sphb200_sfc_keys(context(), x, y, z, numKeys, &b, reinterpret_cast<uint64_t*>(keys)));
}

template<>
void buildOctreeGpu(const uint64_t* cstoneTree, OctreeView<uint64_t> d)
{
    // OctreeData::resize keeps one more childOffsets entry than nodes (tree/octree.hpp:334), which this call clears
    check(sphb200_build_octree(context(), cstoneTree, d.numLeafNodes, d.prefixes, d.childOffsets, d.parents,
                               d.levelRange, d.internalToLeaf, d.leafToInternal));
}

template<>
void computeGeoCentersGpu(const uint64_t* prefixes, TreeNodeIndex numNodes, Vec3<double>* centers, Vec3<double>* sizes,
                          const Box<double>& box)
{
    static_assert(sizeof(Vec3<double>) == 3 * sizeof(double));
    sphb200_box b = toBox(box);
    check(sphb200_node_centers(context(), prefixes, numNodes, &b, reinterpret_cast<double*>(centers),
                               reinterpret_cast<double*>(sizes)));
}

template<>
void findHalosGpu(const uint64_t* prefixes, const TreeNodeIndex* childOffsets, const TreeNodeIndex* internalToLeaf,
                  const uint64_t* leaves, const float* interactionRadii, const Box<double>& box,
                  TreeNodeIndex firstNode, TreeNodeIndex lastNode, int* collisionFlags)
{
    sphb200_box b = toBox(box);
    check(sphb200_find_halos(context(), prefixes, childOffsets, internalToLeaf, leaves, interactionRadii, &b, firstNode,
                             lastNode, collisionFlags, nullptr));
}

template<>
void sortByKeyGpu(uint64_t* first, uint64_t* last, unsigned* values, uint64_t* /*keyBuf*/, unsigned* /*valueBuf*/)
{
    // the library sorts in its own grow-only scratch; the caller's buffers stay untouched
    check(sphb200_sort_pairs(context(), first, values, size_t(last - first)));
}

template<>
void gatherGpu(const unsigned* ordering, size_t numElements, const double* src, double* buffer)
{
    gatherOne(ordering, numElements, src, buffer);
}

template<>
void gatherGpu(const unsigned* ordering, size_t numElements, const float* src, float* buffer)
{
    gatherOne(ordering, numElements, src, buffer);
}

template<>
void gatherGpu(const unsigned* ordering, size_t numElements, const unsigned* src, unsigned* buffer)
{
    gatherOne(ordering, numElements, src, buffer);
}

} // namespace cstone
