// SFC layer: 63-bit Hilbert keys, node geometry, bounding box reduction.
// Compiled with --fmad=false: integer coordinates and node centers must match the
// reference CPU build (compiled -ffp-contract=off) bit for bit.
#include "common.cuh"

namespace sphb200
{

/*! @brief 3D Hilbert key of 21-bit integer coordinates
 *
 * Same bit recurrence as cstone::iHilbert<uint64_t> (domain/include/cstone/sfc/hilbert.hpp:58-110):
 * walk levels 20..0, append the Hilbert digit of the current octant, then transform the
 * remaining low bits (conditional complement, then cyclic rotation or x/z swap).
 */
__device__ __forceinline__ KeyType hilbertKey(unsigned px, unsigned py, unsigned pz)
{
    // octant (x<<2|y<<1|z) -> Hilbert digit, packed 3 bits per entry: {0,1,3,2,7,6,4,5}
    constexpr unsigned digitLut = (0u) | (1u << 3) | (3u << 6) | (2u << 9) | (7u << 12) | (6u << 15) | (4u << 18) | (5u << 21);
    KeyType key = 0;
#pragma unroll
    for (int level = kMaxTreeLevel - 1; level >= 0; --level)
    {
        unsigned xi = (px >> level) & 1u;
        unsigned yi = (py >> level) & 1u;
        unsigned zi = (pz >> level) & 1u;

        unsigned octant = (xi << 2) | (yi << 1) | zi;
        key             = (key << 3) | ((digitLut >> (3 * octant)) & 7u);

        unsigned nyi = yi ^ 1u, nzi = zi ^ 1u;
        px ^= -(xi & (nyi | zi));
        py ^= -((xi & (yi | zi)) | (yi & nzi));
        pz ^= -((xi & nyi & nzi) | (yi & nzi));

        if (zi)
        {
            unsigned pt = px;
            px          = py;
            py          = pz;
            pz          = pt;
        }
        else if (!yi)
        {
            unsigned pt = px;
            px          = pz;
            pz          = pt;
        }
    }
    return key;
}

//! @brief inverse of hilbertKey, cstone::decodeHilbert (hilbert.hpp:146-188)
__device__ __forceinline__ void hilbertDecode(KeyType key, unsigned& ox, unsigned& oy, unsigned& oz)
{
    unsigned px = 0, py = 0, pz = 0;
    for (unsigned level = 0; level < kMaxTreeLevel; ++level)
    {
        unsigned octant = (key >> (3 * level)) & 7u;
        unsigned xi     = octant >> 2u;
        unsigned yi     = (octant >> 1u) & 1u;
        unsigned zi     = octant & 1u;

        if (yi ^ zi)
        {
            unsigned pt = px;
            px          = pz;
            pz          = py;
            py          = pt;
        }
        else if ((!xi & !yi & !zi) || (xi & yi & zi))
        {
            unsigned pt = px;
            px          = pz;
            pz          = pt;
        }

        unsigned mask = (1u << level) - 1;
        px ^= mask & (-(xi & (yi | zi)));
        py ^= mask & (-((xi & ((!yi) | (!zi))) | ((!xi) & yi & zi)));
        pz ^= mask & (-((xi & (!yi) & (!zi)) | (yi & zi)));

        px |= (xi << level);
        py |= ((xi ^ yi) << level);
        pz |= ((yi ^ zi) << level);
    }
    ox = px;
    oy = py;
    oz = pz;
}

struct KeyParams
{
    double xmin, ymin, zmin, mx, my, mz;
};

//! @brief integer coordinate as in cstone::sfc3D (sfc/sfc.hpp:157-175): trunc(floor(x*m) - xmin*m), clamped
__device__ __forceinline__ unsigned iCoord(double x, double xmin, double m)
{
    constexpr int mcoord = (1 << kMaxTreeLevel) - 1;
    double t             = __dsub_rn(floor(__dmul_rn(x, m)), __dmul_rn(xmin, m));
    int ix               = __double2int_rz(t);
    ix                   = min(ix, mcoord);
    return unsigned(max(ix, 0)); // reference asserts ix >= 0
}

/*! @brief one thread per particle, grid-stride; 24 B read + 8 B written per particle (HBM-bound) */
__global__ void __launch_bounds__(256) sfcKeysKernel(const double* __restrict__ x, const double* __restrict__ y,
                                                     const double* __restrict__ z, size_t n, KeyParams p,
                                                     KeyType* __restrict__ keys)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        unsigned ix = iCoord(x[i], p.xmin, p.mx);
        unsigned iy = iCoord(y[i], p.ymin, p.my);
        unsigned iz = iCoord(z[i], p.zmin, p.mz);
        keys[i]     = hilbertKey(ix, iy, iz);
    }
}

/*! @brief node centers and half-sizes from Warren-Salmon prefixes
 *
 * cstone::nodeFpCenters (focus/source_center.hpp:145-158) = decodePlaceholderBit -> hilbertIBox
 * (hilbert.hpp:264-281) -> centerAndSize (sfc/box.hpp:334-359), same operation order, no FMA.
 */
__global__ void nodeCentersKernel(const KeyType* __restrict__ prefixes, int numNodes, sphb200_box box,
                                  double* __restrict__ centers, double* __restrict__ sizes)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numNodes) return;
    KeyType prefix = prefixes[i];
    KeyType start  = decodePlaceholderBit(prefix);
    unsigned level = decodePrefixLength(prefix) / 3;

    unsigned cubeLength = (1u << kMaxTreeLevel) >> level;
    unsigned mask       = ~(cubeLength - 1);
    unsigned ix, iy, iz;
    hilbertDecode(start, ix, iy, iz);
    ix &= mask;
    iy &= mask;
    iz &= mask;

    constexpr double uL = 1.0 / double(1u << kMaxTreeLevel);
    int lo[3]           = {int(ix), int(iy), int(iz)};
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        double halfUnit    = __dmul_rn(__dmul_rn(0.5, uL), box.len[d]);
        int hi             = lo[d] + int(cubeLength);
        centers[3 * i + d] = __dadd_rn(box.lim[2 * d], __dmul_rn(double(hi + lo[d]), halfUnit));
        sizes[3 * i + d]   = __dmul_rn(double(hi - lo[d]), halfUnit);
    }
}

//! @brief per-axis min/max with one ordered-key atomic per block (MinMaxGpu, primitives_gpu.cu:118-128)
__global__ void __launch_bounds__(256) minmax3Kernel(const double* __restrict__ x, const double* __restrict__ y,
                                                     const double* __restrict__ z, size_t n,
                                                     unsigned long long* __restrict__ out)
{
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mxv[3] = {-INFINITY, -INFINITY, -INFINITY};
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        double v[3] = {x[i], y[i], z[i]};
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            mn[d]  = fmin(mn[d], v[d]);
            mxv[d] = fmax(mxv[d], v[d]);
        }
    }
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
        {
            mn[d]  = fmin(mn[d], __shfl_xor_sync(0xffffffffu, mn[d], o));
            mxv[d] = fmax(mxv[d], __shfl_xor_sync(0xffffffffu, mxv[d], o));
        }
    }
    if ((threadIdx.x & 31) == 0)
    {
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            atomicMin(out + 2 * d, orderedKey(mn[d]));
            atomicMax(out + 2 * d + 1, orderedKey(mxv[d]));
        }
    }
}

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_sfc_keys(sphb200_ctx* ctx, const double* x, const double* y, const double* z, size_t n,
                                const sphb200_box* box, uint64_t* keys)
{
    try
    {
        if (n == 0) return 0;
        constexpr double cubeLength = double(1u << kMaxTreeLevel);
        // sfc.hpp:188-192: cubeLength * box.ilx() evaluated once on the host, in double
        KeyParams p{box->lim[0], box->lim[2], box->lim[4], cubeLength * box->inv[0], cubeLength * box->inv[1],
                    cubeLength * box->inv[2]};
        int grid = persistentGrid(n, 256, 16);
        SPHB200_LAUNCH(ctx, sfcKeysKernel, grid, 256, 0, x, y, z, n, p, keys);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_node_centers(sphb200_ctx* ctx, const uint64_t* prefixes, int numNodes, const sphb200_box* box,
                                    double* centers, double* sizes)
{
    try
    {
        if (numNodes == 0) return 0;
        SPHB200_LAUNCH(ctx, nodeCentersKernel, gridFor(numNodes, 128), 128, 0, prefixes, numNodes, *box, centers,
                       sizes);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_minmax3(sphb200_ctx* ctx, const double* x, const double* y, const double* z, size_t n,
                               double* minmax_h)
{
    try
    {
        auto* d = reinterpret_cast<unsigned long long*>(ctx->devDt + 8);
        unsigned long long init[6];
        for (int k = 0; k < 3; ++k)
        {
            init[2 * k]     = ~0ull;
            init[2 * k + 1] = 0ull;
        }
        SPHB200_CUDA(cudaMemcpyAsync(d, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
        if (n) SPHB200_LAUNCH(ctx, minmax3Kernel, persistentGrid(n, 256, 8), 256, 0, x, y, z, n, d);
        unsigned long long res[6];
        SPHB200_CUDA(cudaMemcpyAsync(res, d, sizeof(res), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 6; ++k)
            minmax_h[k] = fromOrderedKey(res[k]);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}
