/* TEST INFRASTRUCTURE ONLY -- C veneer over the time-step hierarchy of the reference's block-time-step propagator.
 *
 * sph/include/sph/ts_rungs.hpp holds the host logic of the hierarchy (rungTimestep :131-145, minimumGroupDt :147-157,
 * computeMinTimestep :78-101, findRungRanges :115-129).  Its templates run on a "device vector" and call five
 * cstone primitives that the reference only defines in CUDA translation units (primitives_gpu.cu, cuda_utils).  This
 * file compiles the header where it lies and supplies those five primitives for host memory with the semantics their
 * declarations document (iota, stable sort by key, lower bound, copies), so that the reference's own sequencing
 * arithmetic -- number of rungs, rung boundaries, step length, time left per substep -- runs unmodified on the CPU.
 * Nothing here restates that arithmetic.
 *
 * Only tests/ may load the resulting library (oracle/ref.py).
 */
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

#include "cstone/cuda/cuda_stubs.h"
#include "cstone/util/reallocate.hpp"

namespace hostdev
{
//! @brief host memory that the reference's templates treat as device memory
template<class T>
struct Vec : public std::vector<T>
{
    using std::vector<T>::vector;
};
template<class T>
void reallocateDevice(Vec<T>& v, std::size_t size, double growthRate)
{
    if (size > v.capacity()) v.reserve(std::size_t(double(size) * growthRate));
    v.resize(size);
}
} // namespace hostdev

template<class T>
struct IsDeviceVector<hostdev::Vec<T>> : public std::true_type
{
};

// what sph/sph.hpp has included by the time it reaches ts_rungs.hpp (the header is not self-contained)
#include "cstone/primitives/mpi_wrappers.hpp"
#include "cstone/traversal/groups.hpp"
#include "cstone/tree/accel_switch.hpp"
#include "cstone/tree/definitions.h"
namespace sph
{
using cstone::GroupView;
using cstone::LocalIndex;
} // namespace sph
#include "sph/ts_rungs.hpp"

// ---- the primitives ts_rungs.hpp calls, for host memory (declared in cstone/primitives/primitives_gpu.h:62,84,96 and
// cstone/cuda/cuda_stubs.h:51)
namespace cstone
{
template<class T>
size_t lowerBoundGpu(const T* first, const T* last, T value)
{
    return std::lower_bound(first, last, value) - first;
}
template<class IndexType>
void sequenceGpu(IndexType* input, size_t numElements, IndexType init)
{
    std::iota(input, input + numElements, init);
}
template<class KeyType, class ValueType>
void sortByKeyGpu(KeyType* first, KeyType* last, ValueType* values, KeyType* keyBuf, ValueType* valueBuf)
{
    size_t n = last - first;
    std::vector<size_t> order(n);
    std::iota(order.begin(), order.end(), size_t(0));
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return first[a] < first[b]; });
    for (size_t i = 0; i < n; ++i)
    {
        keyBuf[i]   = first[order[i]];
        valueBuf[i] = values[order[i]];
    }
    std::copy(keyBuf, keyBuf + n, first);
    std::copy(valueBuf, valueBuf + n, values);
}
} // namespace cstone

template<class T>
void memcpyD2H(const T* src, std::size_t n, T* dest)
{
    std::memcpy(dest, src, n * sizeof(T));
}

static void exportTimestep(const sph::Timestep& ts, float* nextDt, float* totDt, int* numRungs, uint32_t* rungRanges)
{
    *nextDt   = ts.nextDt;
    *totDt    = ts.totDt;
    *numRungs = ts.numRungs;
    std::copy(ts.rungRanges.begin(), ts.rungRanges.end(), rungRanges);
}

extern "C"
{

//! @brief sph::findRungRanges<false> (ts_rungs.hpp:115-129): the host branch, std::lower_bound
void ref_find_rung_ranges(float minDt, const float* groupDt, uint32_t numGroups, int numRungs, uint32_t* rungRanges)
{
    auto r = sph::findRungRanges<false>(minDt, groupDt, numGroups, numRungs);
    std::copy(r.begin(), r.end(), rungRanges);
}

//! @brief sph::rungTimestep (ts_rungs.hpp:131-145) on one rank; groupDt is sorted in place, groupIndices = ordering
void ref_rung_timestep(float* groupDt, uint32_t* groupIndices, uint32_t numGroups, float maxDt, float* nextDt,
                       float* totDt, int* numRungs, uint32_t* rungRanges)
{
    hostdev::Vec<char> scratch;
    sph::Timestep ts = sph::rungTimestep(groupDt, groupIndices, numGroups, maxDt, scratch);
    exportTimestep(ts, nextDt, totDt, numRungs, rungRanges);
}

/*! @brief sph::minimumGroupDt (ts_rungs.hpp:147-157) inside a running hierarchy
 *
 * tsRungRanges: Timestep::rungRanges of the hierarchy (5 entries), numGroups = rungRanges[highRung] as the propagator
 * passes it (ve_hydro_bdt.hpp:312-313); groupIndices needs rungRanges[4] entries
 */
void ref_minimum_group_dt(float totDt, float elapsedDt, int numRungs, int substep, const uint32_t* tsRungRanges,
                          float* groupDt, uint32_t* groupIndices, uint32_t numGroups, float* dt, uint32_t* rungRanges)
{
    sph::Timestep ts;
    ts.nextDt = 0;
    ts.totDt = totDt;
    ts.elapsedDt = elapsedDt;
    ts.numRungs = numRungs;
    ts.substep = substep;
    std::copy(tsRungRanges, tsRungRanges + sph::Timestep::maxNumRungs + 1, ts.rungRanges.begin());
    hostdev::Vec<char> scratch;
    auto [d, ranges] = sph::minimumGroupDt(ts, groupDt, groupIndices, numGroups, scratch);
    *dt = d;
    std::copy(ranges.begin(), ranges.end(), rungRanges);
}

} // extern "C"
