/* TEST INFRASTRUCTURE ONLY -- C veneer over the reference's GPU-only block-time-step group kernels.
 *
 * cstone::computeGroupSplits (domain/include/cstone/traversal/groups.cuh:150-330, what sph::computeSpatialGroups calls,
 * sph/include/sph/groups.cu:30-47) and sph::groupDivvTimestepGpu / groupAccTimestepGpu (sph/include/sph/ts_groups.cu)
 * exist in the reference only as CUDA code.  oracle/build_ref_gpu.sh compiles this file together with ts_groups.cu,
 * from the sources where they lie, for sm_100a into baseline/_ref/libsphexa_ref_gpu.so.  The veneer only calls the
 * reference's functions on device pointers handed in by the GPU tests (tests/test_gpu_reference_kernels.py); nothing
 * of the product links or loads it.
 */
#include <cstdint>
#include <thrust/device_vector.h>

#include "cstone/cuda/cuda_utils.cuh"
#include "cstone/sfc/box.hpp"
#include "cstone/traversal/find_neighbors.cuh"
#include "cstone/traversal/groups.cuh"
#include "cstone/traversal/groups.hpp"

namespace sph
{
using cstone::GroupView;
template<class T>
extern void groupDivvTimestepGpu(float Krho, const GroupView&, const T* divv, float* groupDt);
template<class T>
extern void groupAccTimestepGpu(float etaAcc, const GroupView&, const T* ax, const T* ay, const T* az, float* groupDt);
} // namespace sph

using cstone::LocalIndex;

template<class T>
static int groupSplits(uint32_t first, uint32_t last, const double* x, const double* y, const double* z, const T* h,
                       const uint64_t* leaves, int numLeaves, const uint32_t* layout, const double* lim, const int* bc,
                       float tolFactor, uint32_t* out, uint32_t capacity)
{
    cstone::Box<double> box(lim[0], lim[1], lim[2], lim[3], lim[4], lim[5], cstone::BoundaryType(bc[0]),
                            cstone::BoundaryType(bc[1]), cstone::BoundaryType(bc[2]));
    thrust::device_vector<util::array<cstone::GpuConfig::ThreadMask, cstone::TravConfig::nwt>> S;
    thrust::device_vector<LocalIndex> scratch, groups;
    cstone::computeGroupSplits<cstone::TravConfig::targetSize>(first, last, x, y, z, h, leaves, numLeaves, layout, box,
                                                               tolFactor, S, scratch, groups);
    if (groups.size() > capacity) return -int(groups.size());
    cudaMemcpy(out, thrust::raw_pointer_cast(groups.data()), groups.size() * sizeof(LocalIndex),
               cudaMemcpyDeviceToDevice);
    return int(groups.size());
}

extern "C"
{

//! @brief group delimiters of [first, last) as sph::computeSpatialGroups makes them; returns their number (groups + 1)
int refgpu_group_splits_f64(uint32_t first, uint32_t last, const double* x, const double* y, const double* z,
                            const double* h, const uint64_t* leaves, int numLeaves, const uint32_t* layout,
                            const double* lim, const int* bc, float tolFactor, uint32_t* out, uint32_t capacity)
{
    return groupSplits(first, last, x, y, z, h, leaves, numLeaves, layout, lim, bc, tolFactor, out, capacity);
}
int refgpu_group_splits_f32(uint32_t first, uint32_t last, const double* x, const double* y, const double* z,
                            const float* h, const uint64_t* leaves, int numLeaves, const uint32_t* layout,
                            const double* lim, const int* bc, float tolFactor, uint32_t* out, uint32_t capacity)
{
    return groupSplits(first, last, x, y, z, h, leaves, numLeaves, layout, lim, bc, tolFactor, out, capacity);
}

void refgpu_group_divv_f64(float Krho, uint32_t first, uint32_t last, uint32_t numGroups, const uint32_t* groupStart,
                           const uint32_t* groupEnd, const double* divv, float* groupDt)
{
    sph::groupDivvTimestepGpu(Krho, cstone::GroupView{first, last, numGroups, groupStart, groupEnd}, divv, groupDt);
    cudaDeviceSynchronize();
}
void refgpu_group_divv_f32(float Krho, uint32_t first, uint32_t last, uint32_t numGroups, const uint32_t* groupStart,
                           const uint32_t* groupEnd, const float* divv, float* groupDt)
{
    sph::groupDivvTimestepGpu(Krho, cstone::GroupView{first, last, numGroups, groupStart, groupEnd}, divv, groupDt);
    cudaDeviceSynchronize();
}
void refgpu_group_acc_f64(float etaAcc, uint32_t first, uint32_t last, uint32_t numGroups, const uint32_t* groupStart,
                          const uint32_t* groupEnd, const double* ax, const double* ay, const double* az,
                          float* groupDt)
{
    sph::groupAccTimestepGpu(etaAcc, cstone::GroupView{first, last, numGroups, groupStart, groupEnd}, ax, ay, az,
                             groupDt);
    cudaDeviceSynchronize();
}
void refgpu_group_acc_f32(float etaAcc, uint32_t first, uint32_t last, uint32_t numGroups, const uint32_t* groupStart,
                          const uint32_t* groupEnd, const float* ax, const float* ay, const float* az, float* groupDt)
{
    sph::groupAccTimestepGpu(etaAcc, cstone::GroupView{first, last, numGroups, groupStart, groupEnd}, ax, ay, az,
                             groupDt);
    cudaDeviceSynchronize();
}

} // extern "C"
