// Particle groups and per-group time steps for block time-stepping (HydroVeBdtProp).
//
// Reference: cstone::computeGroupSplits (domain/include/cstone/traversal/groups.cuh:150-330) via
// sph::computeSpatialGroups (sph/include/sph/groups.cu:30-47), groupDivvTimestepGpu / groupAccTimestepGpu /
// storeRungGpu (sph/include/sph/ts_groups.cu:20-103), extractGroupGpu / makeSlicedView (sph/include/sph/groups.hpp:
// 32-58), the group minimum of the Courant step inside momentumAndEnergyGpu (hydro_ve/momentum_energy_gpu.cu:98-104).
//
// A group is a run of SFC-consecutive particles: fixed runs of 32 (TravConfig::targetSize), cut again wherever two
// consecutive particles are further apart than tolFactor (2) x the edge of the smallest leaf cell under the run
// (normalized coordinates).  The reference finds the cuts with warp ballots and rebuilds the delimiters with
// two scans; here every particle decides "do I start a group" on its own and one scan compacts the starts -- the
// same delimiters.
//
// The pair operators of this library work on 32 consecutive list slots per warp, so a set of groups is handed to them
// as a list of target indices in ascending (SFC) order: sphb200_targets_from_groups.
#include "common.cuh"
#include "primitives.cuh"

namespace sphb200
{

constexpr int kMaxRungs = 4;       // Timestep::maxNumRungs (sph/include/sph/timestep.h:44)
constexpr unsigned kGroupSize = 32; // TravConfig::targetSize = GpuConfig::warpSize (traversal/find_neighbors.cuh:62)

/*! @brief mark the particles that start a group
 *
 * One thread per particle of [first, last).  Cut criterion of groupSplitsKernel (groups.cuh:205-232): with
 * w = (i - first) / 32 the fixed run of i, nodeVolume = min over the run's bodies (clamped to last - 1) of the volume
 * of their leaf cell in the unit cube, distCrit = cbrt(nodeVolume) * tolFactor; leaf cells are cubes of edge 2^-level, so cbrt is evaluated exactly as
 * 2^-maxLevel.  i starts a group if it starts a run or |X_i - X_{i-1}| > distCrit in box-normalized coordinates
 * (x * ilx, no minimum image).
 */
__global__ void __launch_bounds__(256)
    groupStartFlagsKernel(LocalIndex first, LocalIndex last, const double* __restrict__ x, const double* __restrict__ y,
                          const double* __restrict__ z, const KeyType* __restrict__ leaves, int numLeaves,
                          const LocalIndex* __restrict__ layout, sphb200_box box, float tolFactor,
                          uint32_t* __restrict__ flags)
{
    LocalIndex t = blockIdx.x * blockDim.x + threadIdx.x;
    LocalIndex n = last - first;
    if (t >= n) return;
    LocalIndex i   = first + t;
    LocalIndex run = t / kGroupSize;
    if (t == run * kGroupSize)
    {
        flags[t] = 1;
        return;
    }
    // deepest leaf under the bodies of the run
    LocalIndex b0 = first + run * kGroupSize;
    LocalIndex b1 = min(b0 + kGroupSize - 1, last - 1);
    int l0 = int(upperBound(layout, size_t(numLeaves), b0)) - 1;
    int l1 = int(upperBound(layout, size_t(numLeaves), b1)) - 1;
    int maxLevel = 0;
    for (int l = l0; l <= l1; ++l)
    {
        // only leaves that hold a body of the run count (the reference takes the leaf of each lane's body): an empty
        // leaf in between can be deeper than all of them when the bucket size is below the run length
        if (layout[l + 1] == layout[l]) continue;
        KeyType range = leaves[l + 1] - leaves[l];
        int level     = 21 - (63 - int(clz64(range))) / 3;
        maxLevel      = max(maxLevel, level);
    }
    double distCrit = ldexp(1.0, -maxLevel) * double(tolFactor);
    // the reference normalizes the positions first (pos_i = x * ilx, kept in registers and exchanged by shuffles) and
    // subtracts afterwards: separately rounded products, no contraction across the subtraction
    double dx = __dsub_rn(__dmul_rn(x[i], box.inv[0]), __dmul_rn(x[i - 1], box.inv[0]));
    double dy = __dsub_rn(__dmul_rn(y[i], box.inv[1]), __dmul_rn(y[i - 1], box.inv[1]));
    double dz = __dsub_rn(__dmul_rn(z[i], box.inv[2]), __dmul_rn(z[i - 1], box.inv[2]));
    flags[t] = (dx * dx + dy * dy + dz * dz > distCrit * distCrit) ? 1u : 0u;
}

//! @brief groups[scan[t]] = first + t for every flagged t; groups[numGroups] = last
__global__ void __launch_bounds__(256)
    compactStartsKernel(LocalIndex first, LocalIndex last, const uint32_t* __restrict__ flags,
                        const uint32_t* __restrict__ scan, uint32_t* __restrict__ groups, uint32_t capacity,
                        const uint32_t* __restrict__ total)
{
    LocalIndex t = blockIdx.x * blockDim.x + threadIdx.x;
    LocalIndex n = last - first;
    if (t == 0 && *total <= capacity) groups[*total] = last;
    if (t >= n || !flags[t]) return;
    if (scan[t] < capacity) groups[scan[t]] = first + t;
}

template<class T>
__global__ void __launch_bounds__(256)
    groupDivvKernel(float Krho, const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd,
                    LocalIndex numGroups, const T* __restrict__ divv, float* __restrict__ groupDt)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= numGroups) return;
    // ts_groups.cu:24-34: a float running maximum that starts at -inf and is taken over divv, not |divv|
    float localMax = -INFINITY;
    for (LocalIndex i = grpStart[g]; i < grpEnd[g]; ++i)
        localMax = max(localMax, float(divv[i]));
    groupDt[g] = min(groupDt[g], Krho / fabsf(localMax));
}

template<class T>
__global__ void __launch_bounds__(256)
    groupAccKernel(float etaAcc, const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd,
                   LocalIndex numGroups, const T* __restrict__ ax, const T* __restrict__ ay, const T* __restrict__ az,
                   float* __restrict__ groupDt)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= numGroups) return;
    float maxAcc = 0;
    for (LocalIndex i = grpStart[g]; i < grpEnd[g]; ++i)
    {
        T a2   = ax[i] * ax[i] + ay[i] * ay[i] + az[i] * az[i];
        maxAcc = max(maxAcc, float(a2));
    }
    groupDt[g] = min(groupDt[g], etaAcc / sqrtf(sqrtf(maxAcc)));
}

//! @brief groupDt[g] = min(groupDt[g], float(min_i dt[i])): what momentumAndEnergyGpu does with its warp minimum
template<class T>
__global__ void __launch_bounds__(256)
    groupMinKernel(const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd, LocalIndex numGroups,
                   const T* __restrict__ dt, float* __restrict__ groupDt)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= numGroups) return;
    T m = INFINITY;
    for (LocalIndex i = grpStart[g]; i < grpEnd[g]; ++i)
        m = min(m, dt[i]);
    groupDt[g] = min(groupDt[g], float(m));
}

__global__ void __launch_bounds__(256)
    storeRungKernel(const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd, LocalIndex numGroups,
                    uint8_t rung, uint8_t* __restrict__ particleRungs)
{
    LocalIndex lane = threadIdx.x & 31;
    LocalIndex g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    // one warp per group like ts_groups.cu:82-92 (groups hold at most 32 particles; the loop also covers longer
    // caller-made groups)
    for (LocalIndex i = grpStart[g] + lane; i < grpEnd[g]; i += 32)
        particleRungs[i] = rung;
}

__global__ void __launch_bounds__(256)
    extractGroupsKernel(const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd,
                        const LocalIndex* __restrict__ indices, LocalIndex numOut, LocalIndex* __restrict__ outStart,
                        LocalIndex* __restrict__ outEnd)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= numOut) return;
    LocalIndex src = indices[g];
    outStart[g]    = grpStart[src];
    outEnd[g]      = grpEnd[src];
}

__global__ void __launch_bounds__(256)
    markGroupMembersKernel(const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd,
                           LocalIndex numGroups, LocalIndex base, uint32_t* __restrict__ flags)
{
    LocalIndex lane = threadIdx.x & 31;
    LocalIndex g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    for (LocalIndex i = grpStart[g] + lane; i < grpEnd[g]; i += 32)
        flags[i - base] = 1;
}

__global__ void __launch_bounds__(256)
    compactTargetsKernel(LocalIndex base, LocalIndex n, const uint32_t* __restrict__ flags,
                         const uint32_t* __restrict__ scan, uint32_t* __restrict__ targets)
{
    LocalIndex t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n || !flags[t]) return;
    targets[scan[t]] = base + t;
}


//! @brief active[i] = 1 for the members of the groups (one warp per group)
__global__ void __launch_bounds__(256)
    markActiveKernel(const LocalIndex* __restrict__ grpStart, const LocalIndex* __restrict__ grpEnd,
                     LocalIndex numGroups, uint8_t* __restrict__ active)
{
    LocalIndex lane = threadIdx.x & 31;
    LocalIndex g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    for (LocalIndex i = grpStart[g] + lane; i < grpEnd[g]; i += 32)
        active[i] = 1;
}

//! @brief positive floats order like their bit patterns: radix-sort keys for sortGroupDt
__global__ void __launch_bounds__(256) dtToKeysKernel(const float* __restrict__ dt, LocalIndex n, KeyType* __restrict__ keys)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n) keys[g] = KeyType(__float_as_uint(dt[g]));
}

__global__ void __launch_bounds__(256) keysToDtKernel(const KeyType* __restrict__ keys, LocalIndex n, float* __restrict__ dt)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n) dt[g] = __uint_as_float(uint32_t(keys[g]));
}

__global__ void __launch_bounds__(256) iotaKernel(LocalIndex* __restrict__ out, LocalIndex n, LocalIndex first)
{
    LocalIndex g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n) out[g] = first + g;
}

/*! @brief findRungRanges (ts_rungs.hpp:115-129) + timestepRangeGpu (:69-75) in one launch
 *
 * out[0] = 0, out[r] = lower_bound(groupDt, float(1 << r) * minDt) for 1 <= r < numRungs, numGroups otherwise.  With
 * minDt < 0 the minimum groupDt[0] itself is used (the caller has not seen it yet).  dtOut = {groupDt[0],
 * groupDt[fastIndex]}.
 */
__global__ void rungRangesKernel(const float* __restrict__ groupDt, LocalIndex numGroups, float minDt, int numRungs,
                                 LocalIndex fastIndex, LocalIndex* __restrict__ out, float* __restrict__ dtOut)
{
    int r = threadIdx.x;
    if (r == 0)
    {
        dtOut[0] = groupDt[0];
        dtOut[1] = groupDt[fastIndex];
    }
    if (r > kMaxRungs) return;
    if (minDt < 0) minDt = groupDt[0];
    LocalIndex res = numGroups;
    if (r == 0) res = 0;
    else if (r < numRungs)
    {
        float maxDtRung = float(1 << r) * minDt;
        LocalIndex lo = 0, hi = numGroups;
        while (lo < hi)
        {
            LocalIndex mid = lo + (hi - lo) / 2;
            if (groupDt[mid] < maxDtRung) lo = mid + 1;
            else hi = mid;
        }
        res = lo;
    }
    out[r] = res;
}

} // namespace sphb200

using namespace sphb200;

#define GROUPS_TRY try {
#define GROUPS_CATCH                                                                                                  \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (CapacityError & e)                                                                                         \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return SPHB200_ERR_CAPACITY;                                                                                  \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return 1;                                                                                                     \
    }

extern "C" int sphb200_compute_groups(sphb200_ctx* ctx, uint32_t first, uint32_t last, const double* x, const double* y,
                                      const double* z, const uint64_t* leaves, int numLeaves, const uint32_t* layout,
                                      const sphb200_box* box, float tolFactor, uint32_t* groups, uint32_t capacity,
                                      uint32_t* numGroups_h)
{
    GROUPS_TRY
    uint32_t n = last - first;
    if (n == 0)
    {
        *numGroups_h = 0;
        return 0;
    }
    auto* flags = static_cast<uint32_t*>(ctx->miscTmp.get((2 * size_t(n) + 2) * sizeof(uint32_t)));
    auto* scan  = flags + n;
    auto* total = scan + n;
    SPHB200_LAUNCH(ctx, groupStartFlagsKernel, gridFor(n, 256), 256, 0, first, last, x, y, z, leaves, numLeaves, layout,
                   *box, tolFactor, flags);
    exclusiveScan(ctx, flags, scan, n, total);
    SPHB200_LAUNCH(ctx, compactStartsKernel, gridFor(n, 256), 256, 0, first, last, flags, scan, groups, capacity, total);
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    *numGroups_h = uint32_t(ctx->hostScalars[0]);
    if (*numGroups_h > capacity) throw CapacityError("compute_groups: group capacity exceeded");
    GROUPS_CATCH
}

extern "C" int sphb200_group_divv_timestep(sphb200_ctx* ctx, int dtype, float Krho, const sphb200_group_view* grp,
                                           const void* divv, float* groupDt)
{
    GROUPS_TRY
    if (grp->numGroups == 0) return 0;
    if (dtype == SPHB200_F64)
        SPHB200_LAUNCH(ctx, groupDivvKernel<double>, gridFor(grp->numGroups, 256), 256, 0, Krho, grp->groupStart,
                       grp->groupEnd, grp->numGroups, static_cast<const double*>(divv), groupDt);
    else
        SPHB200_LAUNCH(ctx, groupDivvKernel<float>, gridFor(grp->numGroups, 256), 256, 0, Krho, grp->groupStart,
                       grp->groupEnd, grp->numGroups, static_cast<const float*>(divv), groupDt);
    GROUPS_CATCH
}

extern "C" int sphb200_group_acc_timestep(sphb200_ctx* ctx, int dtype, float etaAcc, const sphb200_group_view* grp,
                                          const void* ax, const void* ay, const void* az, float* groupDt)
{
    GROUPS_TRY
    if (grp->numGroups == 0) return 0;
    if (dtype == SPHB200_F64)
        SPHB200_LAUNCH(ctx, groupAccKernel<double>, gridFor(grp->numGroups, 256), 256, 0, etaAcc, grp->groupStart,
                       grp->groupEnd, grp->numGroups, static_cast<const double*>(ax), static_cast<const double*>(ay),
                       static_cast<const double*>(az), groupDt);
    else
        SPHB200_LAUNCH(ctx, groupAccKernel<float>, gridFor(grp->numGroups, 256), 256, 0, etaAcc, grp->groupStart,
                       grp->groupEnd, grp->numGroups, static_cast<const float*>(ax), static_cast<const float*>(ay),
                       static_cast<const float*>(az), groupDt);
    GROUPS_CATCH
}

extern "C" int sphb200_group_min_timestep(sphb200_ctx* ctx, int dtype, const sphb200_group_view* grp,
                                          const void* dtParticle, float* groupDt)
{
    GROUPS_TRY
    if (grp->numGroups == 0) return 0;
    if (dtype == SPHB200_F64)
        SPHB200_LAUNCH(ctx, groupMinKernel<double>, gridFor(grp->numGroups, 256), 256, 0, grp->groupStart,
                       grp->groupEnd, grp->numGroups, static_cast<const double*>(dtParticle), groupDt);
    else
        SPHB200_LAUNCH(ctx, groupMinKernel<float>, gridFor(grp->numGroups, 256), 256, 0, grp->groupStart, grp->groupEnd,
                       grp->numGroups, static_cast<const float*>(dtParticle), groupDt);
    GROUPS_CATCH
}

extern "C" int sphb200_store_rung(sphb200_ctx* ctx, const sphb200_group_view* grp, uint8_t rung, uint8_t* particleRungs)
{
    GROUPS_TRY
    if (grp->numGroups == 0) return 0;
    SPHB200_LAUNCH(ctx, storeRungKernel, gridFor(size_t(grp->numGroups) * 32, 256), 256, 0, grp->groupStart,
                   grp->groupEnd, grp->numGroups, rung, particleRungs);
    GROUPS_CATCH
}

extern "C" int sphb200_extract_groups(sphb200_ctx* ctx, const sphb200_group_view* grp, const uint32_t* indices,
                                      uint32_t first, uint32_t last, uint32_t* outStart, uint32_t* outEnd)
{
    GROUPS_TRY
    uint32_t numOut = last - first;
    if (numOut == 0) return 0;
    SPHB200_LAUNCH(ctx, extractGroupsKernel, gridFor(numOut, 256), 256, 0, grp->groupStart, grp->groupEnd,
                   indices + first, numOut, outStart, outEnd);
    GROUPS_CATCH
}

extern "C" int sphb200_targets_from_groups(sphb200_ctx* ctx, const sphb200_group_view* grp, uint32_t* targets,
                                           uint32_t* numTargets_h)
{
    GROUPS_TRY
    *numTargets_h = 0;
    uint32_t n    = grp->lastBody - grp->firstBody;
    if (grp->numGroups == 0 || n == 0) return 0;
    auto* flags = static_cast<uint32_t*>(ctx->miscTmp.get((2 * size_t(n) + 2) * sizeof(uint32_t)));
    auto* scan  = flags + n;
    auto* total = scan + n;
    SPHB200_CUDA(cudaMemsetAsync(flags, 0, size_t(n) * sizeof(uint32_t), ctx->stream));
    SPHB200_LAUNCH(ctx, markGroupMembersKernel, gridFor(size_t(grp->numGroups) * 32, 256), 256, 0, grp->groupStart,
                   grp->groupEnd, grp->numGroups, grp->firstBody, flags);
    exclusiveScan(ctx, flags, scan, n, total);
    SPHB200_LAUNCH(ctx, compactTargetsKernel, gridFor(n, 256), 256, 0, grp->firstBody, n, flags, scan, targets);
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    *numTargets_h = uint32_t(ctx->hostScalars[0]);
    GROUPS_CATCH
}

extern "C" int sphb200_mark_groups(sphb200_ctx* ctx, const sphb200_group_view* grp, uint32_t n, uint8_t* active)
{
    GROUPS_TRY
    if (n) SPHB200_CUDA(cudaMemsetAsync(active, 0, n, ctx->stream));
    if (grp->numGroups == 0) return 0;
    SPHB200_LAUNCH(ctx, markActiveKernel, gridFor(size_t(grp->numGroups) * 32, 256), 256, 0, grp->groupStart,
                   grp->groupEnd, grp->numGroups, active);
    GROUPS_CATCH
}

extern "C" int sphb200_set_scope(sphb200_ctx* ctx, const sphb200_op_scope* scope)
{
    GROUPS_TRY
    ctx->scopeActive     = scope ? scope->active : nullptr;
    ctx->searchExtFactor = (scope && scope->searchExtFactor > 0) ? scope->searchExtFactor : 1.0f;
    ctx->scopeDtParticle = scope ? scope->dtParticle : nullptr;
    ctx->scopeGather     = (scope && scope->active) ? scope->gatherMask : nullptr;
    ctx->packValid       = 0; // records packed under another scope are not complete
    GROUPS_CATCH
}

extern "C" int sphb200_sort_group_dt(sphb200_ctx* ctx, float* groupDt, uint32_t* groupIndices, uint32_t numGroups)
{
    GROUPS_TRY
    if (numGroups == 0) return 0;
    auto* keys = static_cast<KeyType*>(ctx->miscTmp.get(size_t(numGroups) * sizeof(KeyType)));
    SPHB200_LAUNCH(ctx, dtToKeysKernel, gridFor(numGroups, 256), 256, 0, groupDt, numGroups, keys);
    // stable, like the cub radix sort behind sortByKeyGpu: equal time steps keep their SFC order
    radixSortPairs(ctx, keys, groupIndices, numGroups, true, 0, 32);
    SPHB200_LAUNCH(ctx, keysToDtKernel, gridFor(numGroups, 256), 256, 0, keys, numGroups, groupDt);
    GROUPS_CATCH
}

extern "C" int sphb200_sequence(sphb200_ctx* ctx, uint32_t* out, uint32_t n, uint32_t first)
{
    GROUPS_TRY
    if (n) SPHB200_LAUNCH(ctx, iotaKernel, gridFor(n, 256), 256, 0, out, n, first);
    GROUPS_CATCH
}

extern "C" int sphb200_rung_ranges(sphb200_ctx* ctx, float minDt, const float* groupDt, uint32_t numGroups, int numRungs,
                                   float fastFraction, uint32_t* rungRanges_h, float* minGroupDt_h)
{
    GROUPS_TRY
    if (numGroups == 0)
    {
        for (int r = 0; r <= kMaxRungs; ++r)
            rungRanges_h[r] = 0;
        if (minGroupDt_h) minGroupDt_h[0] = minGroupDt_h[1] = INFINITY;
        return 0;
    }
    auto* dOut = reinterpret_cast<uint32_t*>(ctx->devScalars + 8);
    auto* dDt  = reinterpret_cast<float*>(ctx->devScalars + 16);
    SPHB200_LAUNCH(ctx, rungRangesKernel, 1, 32, 0, groupDt, numGroups, minDt, numRungs,
                   LocalIndex(fastFraction * numGroups), dOut, dDt);
    SPHB200_CUDA(cudaMemcpyAsync(ctx->hostScalars + 8, ctx->devScalars + 8, 10 * sizeof(int), cudaMemcpyDeviceToHost,
                                 ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r <= kMaxRungs; ++r)
        rungRanges_h[r] = uint32_t(ctx->hostScalars[8 + r]);
    if (minGroupDt_h) std::memcpy(minGroupDt_h, ctx->hostScalars + 16, 2 * sizeof(float));
    GROUPS_CATCH
}
