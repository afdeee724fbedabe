// Shared helpers for the sm_100a hot-path kernels.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>
#include <cuda_runtime.h>
#include "sphexa_b200.h"

namespace sphb200
{

using KeyType       = uint64_t;
using LocalIndex    = uint32_t;
using TreeNodeIndex = int32_t;

constexpr int kMaxTreeLevel = 21;              // maxTreeLevel<uint64_t>, sfc/common.hpp
constexpr int kNumSms       = 148;             // B200: 2 dies x 74 SMs
constexpr int kTableSize    = 20000;           // sph/table_lookup.hpp:11

// Neighbor-list layout shared by the search (neighbors.cu) and the J-loops (hydro.cu): one lane per target, a warp
// covers 32 SFC-consecutive targets and entry k of all 32 lanes is one 128-byte line.
// (Measured alternatives, Sedov 200^3 on B200: 2/4/8 lanes per target and 16-byte index blocks with 2-4 gathers in
// flight per lane were all slower -- fewer L1 wavefronts or more ILP did not pay for the lost occupancy.)
__host__ __device__ inline size_t nbOffset(size_t t, unsigned k, unsigned ngmax)
{
    return ((t >> 5) * ngmax + k) * 32 + (t & 31);
}

void setError(const std::string& msg);

struct CudaError : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

struct CapacityError : std::runtime_error
{
    using std::runtime_error::runtime_error;
};

#define SPHB200_CUDA(call)                                                                                            \
    do                                                                                                                \
    {                                                                                                                 \
        cudaError_t err__ = (call);                                                                                   \
        if (err__ != cudaSuccess)                                                                                     \
        {                                                                                                             \
            throw ::sphb200::CudaError(std::string(#call) + " failed: " + cudaGetErrorString(err__) + " at " +       \
                                       __FILE__ + ":" + std::to_string(__LINE__));                                   \
        }                                                                                                             \
    } while (0)

// device scratch buffer that only ever grows (zero allocations in steady state)
struct Arena
{
    void* ptr       = nullptr;
    size_t capacity = 0;
    void* get(size_t bytes)
    {
        if (bytes > capacity)
        {
            if (ptr) SPHB200_CUDA(cudaFree(ptr));
            size_t newCap = size_t(double(bytes) * 1.05) + 256;
            SPHB200_CUDA(cudaMalloc(&ptr, newCap));
            capacity = newCap;
        }
        return ptr;
    }
    void release()
    {
        if (ptr) cudaFree(ptr);
        ptr      = nullptr;
        capacity = 0;
    }
};

} // namespace sphb200

struct sphb200_ctx
{
    int device           = 0;
    cudaStream_t stream  = nullptr;
    bool ownStream       = false;
    uint64_t launches    = 0;
    sphb200::Arena sortTmp, treeTmp, scanTmp, miscTmp, nbTmp, obsTmp, machTmp;
    double* devObs = nullptr; // last result of sphb200_conserved_quantities (10 doubles, inside obsTmp)
    double* devMach = nullptr; // last result of sphb200_mach_square_sum (1 double, inside obsTmp)
    // gather records of the pair operators (hydro.cu) and the repacked kernel tables
    static constexpr int kMaxRecs = 12;
    sphb200::Arena rec[kMaxRecs], tabPairs, tabQuads;
    uint32_t packValid = 0;    // bit k: rec[k] mirrors the current field arrays
    bool packReuse     = false; // set by sphb200_sim: records stay valid across operator calls
    bool tabValid      = false;
    const void *tabWh = nullptr, *tabWhd = nullptr;
    int tabElem = 0;
    // pinned host + device scalars
    int* hostScalars     = nullptr; // pinned, 64 ints
    int* devScalars      = nullptr; // 64 ints
    double* devDt        = nullptr; // [0] courant, [1] rho, [2] local min, [3..] scratch
    double* hostDt       = nullptr; // pinned
    // GroupView scope of the target-range operators (sphb200_set_scope)
    const uint8_t* scopeActive = nullptr;
    float searchExtFactor      = 1.0f;
    void* scopeDtParticle      = nullptr;
    uint8_t* scopeGather       = nullptr;
    // kernel-table variant of the pair operators (hydro.cu): -1 decided from timed launches, 0 L1, 1 shared memory
    int tabForce = -1;
    struct TabTune
    {
        int phase   = 0;  // 0: time the L1 variant next, 1: time the shared-memory variant next, 2: decided
        int calls   = 0;  // launches since the decision
        int pending = -1; // variant whose launch is bracketed by ev[0], ev[1] and not read back yet
        bool closed = false;
        size_t pendingTargets = 0;
        size_t tunedTargets   = 0; // launch size the decision was taken at
        double perTarget[2] = {0, 0};
        bool smem = false;
        cudaEvent_t ev[2] = {nullptr, nullptr};
    };
    static constexpr int kMaxTabOps = 8;
    TabTune tabTune[kMaxTabOps];
    //! @brief forget the decisions: the next launches time both variants again (a new simulation uses this context)
    void resetTableTuners()
    {
        for (auto& t : tabTune)
        {
            t.phase   = 0;
            t.pending = -1;
        }
    }
    bool hilbertTabReady       = false; // sfc.cu: the 3-level Hilbert transition table is on this device
};

namespace sphb200
{

inline int gridFor(size_t n, int block) { return int((n + block - 1) / block); }

//! @brief grid sized in multiples of the SM count for grid-stride kernels
inline int persistentGrid(size_t n, int block, int ctasPerSm)
{
    size_t needed = (n + block - 1) / block;
    size_t cap    = size_t(kNumSms) * ctasPerSm;
    return int(needed < cap ? (needed ? needed : 1) : cap);
}

#define SPHB200_LAUNCH(ctx, kernel, grid, block, smem, ...)                                                          \
    do                                                                                                                \
    {                                                                                                                 \
        kernel<<<(grid), (block), (smem), (ctx)->stream>>>(__VA_ARGS__);                                              \
        (ctx)->launches++;                                                                                            \
        SPHB200_CUDA(cudaGetLastError());                                                                             \
    } while (0)

// ---------- device helpers ----------
#ifdef __CUDACC__

__host__ __device__ inline unsigned clz64(uint64_t x)
{
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}

//! @brief nodeRange<uint64_t>(level), sfc/common.hpp:125-132
__host__ __device__ inline KeyType nodeRange(unsigned level) { return KeyType(1) << (3u * (kMaxTreeLevel - level)); }

//! @brief treeLevel(range), sfc/common.hpp:173-178 (1 unused bit for 64-bit keys)
__host__ __device__ inline unsigned treeLevel(KeyType range) { return (clz64(range - 1) - 1) / 3; }

//! @brief octalDigit, sfc/common.hpp:238-242
__host__ __device__ inline unsigned octalDigit(KeyType code, unsigned position)
{
    return (code >> (3u * (kMaxTreeLevel - position))) & 7u;
}

//! @brief Warren-Salmon placeholder-bit format, sfc/common.hpp:189-222
__host__ __device__ inline KeyType encodePlaceholderBit(KeyType code, int prefixLength)
{
    int nShifts = 3 * kMaxTreeLevel - prefixLength;
    return (KeyType(1) << prefixLength) | (code >> nShifts);
}
__host__ __device__ inline unsigned decodePrefixLength(KeyType code) { return 63 - clz64(code); }
__host__ __device__ inline KeyType decodePlaceholderBit(KeyType code)
{
    int prefixLength = decodePrefixLength(code);
    return (code ^ (KeyType(1) << prefixLength)) << (3 * kMaxTreeLevel - prefixLength);
}
//! @brief commonPrefix, sfc/common.hpp:161-165
__host__ __device__ inline int commonPrefix(KeyType a, KeyType b) { return int(clz64(a ^ b)) - 1; }

//! @brief first index in [0,n) with a[idx] >= v  (stl::lower_bound)
template<class T>
__host__ __device__ inline size_t lowerBound(const T* a, size_t n, T v)
{
    size_t lo = 0, hi = n;
    while (lo < hi)
    {
        size_t mid = (lo + hi) >> 1;
        if (a[mid] < v) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

//! @brief first index with a[idx] > v (std::upper_bound)
template<class T>
__host__ __device__ inline size_t upperBound(const T* a, size_t n, T v)
{
    size_t lo = 0, hi = n;
    while (lo < hi)
    {
        size_t mid = (lo + hi) >> 1;
        if (!(v < a[mid])) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

//! @brief order-preserving map double -> uint64 for atomicMin/Max on floating point values
__device__ inline unsigned long long orderedKey(double v)
{
    unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__host__ __device__ inline double fromOrderedKey(unsigned long long k)
{
    unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    double d;
    memcpy(&d, &b, 8);
    return d;
#endif
}

#endif // __CUDACC__

} // namespace sphb200
