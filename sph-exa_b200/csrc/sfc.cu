// SFC layer: 63-bit Hilbert keys, node geometry, bounding box reduction.
// Compiled with --fmad=false: integer coordinates and node centers must match the
// reference CPU build (compiled -ffp-contract=off) bit for bit.
#include <algorithm>

#include "common.cuh"
#include "primitives.cuh"
#include "hilbert.cuh"

namespace sphb200
{

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

/*! @brief one thread per particle, grid-stride, bit recurrence: ~650 integer instructions per key, ALU-bound
 *         (ncu: alu pipe 96 %).  Used for short ranges, where loading the transition table does not pay. */
__global__ void __launch_bounds__(256) sfcKeysBitwiseKernel(const double* __restrict__ x, const double* __restrict__ y,
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

// ---------------------------------------------------------------- table-driven Hilbert keys
/* The recurrence of hilbertKey only ever applies an axis permutation and per-axis complements to the low bits, so the
 * remaining work at any level is a function of (state, octant) with state = one of the 24 signed axis permutations
 * reachable from the identity.  Three levels at a time: entry [state][x3 y3 z3] = 9 key bits | next state << 9,
 * 24 * 512 uint16 = 24 KB in shared memory; 7 lookups per key instead of 21 rounds of bit logic. */
//! @brief bank swizzle of the table index: lanes of a warp tend to share y3, z3 (particles along a lattice row, one
//!        leaf) and differ in x3 and the state, which would all fall into one bank (bits 6.. of a uint16 index)
__host__ __device__ inline unsigned hilbertSwizzle(unsigned idx) { return idx ^ ((idx >> 5) & 0x3Eu); }

constexpr int kHilbertStates  = 24;
constexpr int kHilbertTabSize = kHilbertStates * 512;
__device__ uint16_t gHilbertTab[kHilbertTabSize];

namespace
{

struct HilbertState
{
    int perm[3]; // current axis a reads original axis perm[a] ...
    int flip[3]; // ... complemented when flip[a]
    bool operator==(const HilbertState& o) const
    {
        return std::equal(perm, perm + 3, o.perm) && std::equal(flip, flip + 3, o.flip);
    }
};

//! @brief one level of hilbertKey on the original octant bits under @p s: Hilbert digit and the state of the levels below
int hilbertLevel(const HilbertState& s, const int orig[3], HilbertState& next)
{
    constexpr int digitOf[8] = {0, 1, 3, 2, 7, 6, 4, 5};
    int xi = orig[s.perm[0]] ^ s.flip[0], yi = orig[s.perm[1]] ^ s.flip[1], zi = orig[s.perm[2]] ^ s.flip[2];
    int nyi = yi ^ 1, nzi = zi ^ 1;
    int f[3]   = {xi & (nyi | zi), (xi & (yi | zi)) | (yi & nzi), (xi & nyi & nzi) | (yi & nzi)};
    int rot[3] = {1, 2, 0}, swp[3] = {2, 1, 0}, idn[3] = {0, 1, 2};
    const int* sig = zi ? rot : (!yi ? swp : idn); // new axis a = old axis sig[a]
    for (int a = 0; a < 3; ++a)
    {
        next.perm[a] = s.perm[sig[a]];
        next.flip[a] = s.flip[sig[a]] ^ f[sig[a]];
    }
    return digitOf[(xi << 2) | (yi << 1) | zi];
}

std::vector<uint16_t> buildHilbertTable()
{
    std::vector<HilbertState> states{HilbertState{{0, 1, 2}, {0, 0, 0}}};
    auto idOf = [&](const HilbertState& s)
    {
        for (size_t k = 0; k < states.size(); ++k)
            if (states[k] == s) return int(k);
        states.push_back(s);
        return int(states.size()) - 1;
    };
    for (size_t k = 0; k < states.size(); ++k) // closure of the identity under the 8 octant transitions
        for (int o = 0; o < 8; ++o)
        {
            int orig[3] = {(o >> 2) & 1, (o >> 1) & 1, o & 1};
            HilbertState next;
            hilbertLevel(HilbertState(states[k]), orig, next);
            idOf(next);
        }
    if (int(states.size()) != kHilbertStates) throw std::runtime_error("sfc: unexpected number of Hilbert states");

    std::vector<uint16_t> tab(kHilbertTabSize);
    for (int st = 0; st < kHilbertStates; ++st)
        for (int in = 0; in < 512; ++in)
        {
            int x3 = in >> 6, y3 = (in >> 3) & 7, z3 = in & 7;
            HilbertState cur = states[st];
            int digits       = 0;
            for (int b = 2; b >= 0; --b)
            {
                int orig[3] = {(x3 >> b) & 1, (y3 >> b) & 1, (z3 >> b) & 1};
                HilbertState next;
                digits = (digits << 3) | hilbertLevel(cur, orig, next);
                cur    = next;
            }
            tab[hilbertSwizzle(unsigned(st * 512 + in))] = uint16_t(digits | (idOf(cur) << 9));
        }
    return tab;
}

} // namespace

/*! @brief Hilbert keys through the transition table; same keys as sfcKeysBitwiseKernel
 *
 * 24 B read + 8 B written per particle.  Per key: 3 coordinate conversions, 7 shared-memory lookups with ~12 integer
 * instructions each.  The entry keeps the next state in bits 9..15, i.e. already shifted to its place in the index.
 */
__global__ void __launch_bounds__(256) sfcKeysKernel(const double* __restrict__ x, const double* __restrict__ y,
                                                     const double* __restrict__ z, size_t n, KeyParams p,
                                                     KeyType* __restrict__ keys)
{
    __shared__ __align__(16) uint16_t tab[kHilbertTabSize];
    {
        const uint4* src = reinterpret_cast<const uint4*>(gHilbertTab);
        uint4* dst       = reinterpret_cast<uint4*>(tab);
        for (int k = threadIdx.x; k < kHilbertTabSize / 8; k += blockDim.x)
            dst[k] = src[k];
    }
    __syncthreads();
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        unsigned ix = iCoord(x[i], p.xmin, p.mx);
        unsigned iy = iCoord(y[i], p.ymin, p.my);
        unsigned iz = iCoord(z[i], p.zmin, p.mz);
        // groups of three levels, top first; x3 lands in bits 6..8, y3 in 3..5, z3 in 0..2 of the index
        unsigned e, idx;
        unsigned part[3]; // 27 + 27 + 9 key bits
#pragma unroll
        for (int q = 0; q < 3; ++q)
        {
            unsigned acc = 0;
#pragma unroll
            for (int g = (q == 0 ? 6 : (q == 1 ? 3 : 0)); g >= (q == 0 ? 4 : (q == 1 ? 1 : 0)); --g)
            {
                unsigned xs = (3 * g >= 6) ? (ix >> (3 * g - 6)) : (ix << (6 - 3 * g));
                unsigned ys = (3 * g >= 3) ? (iy >> (3 * g - 3)) : (iy << (3 - 3 * g));
                unsigned zs = iz >> (3 * g);
                idx         = (xs & 0x1C0u) | (ys & 0x38u) | (zs & 7u);
                if (g != 6) idx |= e & 0xFE00u;
                e   = tab[hilbertSwizzle(idx)];
                acc = (acc << 9) | (e & 511u);
            }
            part[q] = acc;
        }
        keys[i] = (KeyType(part[0]) << 36) | (KeyType(part[1]) << 9) | KeyType(part[2]);
    }
}

// ---------------------------------------------------------------- glass-block assembly (init/grid.hpp:152-300)
/* A template block with coordinates in [0,1)^3 is stacked multiplicity[d] times per dimension into the global box; a
 * rank keeps the copies that fall into the octree cells spanning its SFC range.  The host lists the (cell, tile)
 * pairs in the reference's loop order, two kernels count and write the selected particles, so that the output has the
 * reference's order: cell, tile (x, y, z nested), position in the block. */
struct AssemblePair
{
    double sel[6];    // xmin xmax ymin ymax zmin zmax of the cell (createFpBox, box.hpp:361-370)
    double origin[3]; // globalOrigin + blockOrigin of the tile (scaleBlockToGlobal, grid.hpp:168-184)
};

constexpr int kAssembleChunk = 256;

//! @brief scaleBlockToGlobal + the half-open box test of extractBlock (grid.hpp:201-222)
__device__ __forceinline__ bool blockToGlobal(const AssemblePair& p, double sx, double sy, double sz, double ux,
                                              double uy, double uz, double& gx, double& gy, double& gz)
{
    gx = __dadd_rn(__dmul_rn(ux, sx), p.origin[0]);
    gy = __dadd_rn(__dmul_rn(uy, sy), p.origin[1]);
    gz = __dadd_rn(__dmul_rn(uz, sz), p.origin[2]);
    return p.sel[0] <= gx && gx < p.sel[1] && p.sel[2] <= gy && gy < p.sel[3] && p.sel[4] <= gz && gz < p.sel[5];
}

template<bool Write>
__global__ void __launch_bounds__(kAssembleChunk)
    assembleKernel(const AssemblePair* __restrict__ pairs, uint32_t numChunks, double sx, double sy, double sz,
                   const double* __restrict__ xb, const double* __restrict__ yb, const double* __restrict__ zb,
                   size_t blockSize, uint32_t* __restrict__ counts, const uint32_t* __restrict__ offsets,
                   double* __restrict__ x, double* __restrict__ y, double* __restrict__ z)
{
    __shared__ uint32_t warpCount[kAssembleChunk / 32];
    uint32_t pairIdx = blockIdx.x / numChunks, chunk = blockIdx.x % numChunks;
    size_t i         = size_t(chunk) * kAssembleChunk + threadIdx.x;
    const AssemblePair& p = pairs[pairIdx];
    double gx = 0, gy = 0, gz = 0;
    bool sel = i < blockSize && blockToGlobal(p, sx, sy, sz, xb[i], yb[i], zb[i], gx, gy, gz);
    unsigned ballot = __ballot_sync(0xffffffffu, sel);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warpCount[warp] = __popc(ballot);
    __syncthreads();
    if (!Write)
    {
        if (threadIdx.x == 0)
        {
            uint32_t c = 0;
            for (int w = 0; w < kAssembleChunk / 32; ++w)
                c += warpCount[w];
            counts[blockIdx.x] = c;
        }
        return;
    }
    if (sel)
    {
        uint32_t pos = offsets[blockIdx.x] + __popc(ballot & ((1u << lane) - 1));
        for (int w = 0; w < warp; ++w)
            pos += warpCount[w];
        x[pos] = gx;
        y[pos] = gy;
        z[pos] = gz;
    }
}

//! @brief cutSphere (grid.hpp:267-300): keep sqrt(x^2 + y^2 + z^2) <= radius, order preserved
template<bool Write>
__global__ void __launch_bounds__(kAssembleChunk)
    cutSphereKernel(double radius, const double* __restrict__ x, const double* __restrict__ y,
                    const double* __restrict__ z, size_t n, uint32_t* __restrict__ counts,
                    const uint32_t* __restrict__ offsets, double* __restrict__ ox, double* __restrict__ oy,
                    double* __restrict__ oz)
{
    __shared__ uint32_t warpCount[kAssembleChunk / 32];
    size_t i  = size_t(blockIdx.x) * kAssembleChunk + threadIdx.x;
    double xi = 0, yi = 0, zi = 0;
    bool sel  = false;
    if (i < n)
    {
        xi  = x[i], yi = y[i], zi = z[i];
        sel = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(xi, xi), __dmul_rn(yi, yi)), __dmul_rn(zi, zi))) <= radius;
    }
    unsigned ballot = __ballot_sync(0xffffffffu, sel);
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) warpCount[warp] = __popc(ballot);
    __syncthreads();
    if (!Write)
    {
        if (threadIdx.x == 0)
        {
            uint32_t c = 0;
            for (int w = 0; w < kAssembleChunk / 32; ++w)
                c += warpCount[w];
            counts[blockIdx.x] = c;
        }
        return;
    }
    if (sel)
    {
        uint32_t pos = offsets[blockIdx.x] + __popc(ballot & ((1u << lane) - 1));
        for (int w = 0; w < warp; ++w)
            pos += warpCount[w];
        ox[pos] = xi;
        oy[pos] = yi;
        oz[pos] = zi;
    }
}

namespace
{

//! @brief cstone::spanSfcRange (sfc/common.hpp:362-400): octree cells covering [a, b), cells.back() == b
std::vector<KeyType> spanKeyRange(KeyType a, KeyType b)
{
    auto digit   = [](KeyType k, int pos) { return unsigned(k >> (3 * (kMaxTreeLevel - pos))) & 7u; };
    auto power   = [](int pos) { return KeyType(1) << (3 * (kMaxTreeLevel - pos)); };
    auto lastNz  = [](KeyType k) { return k ? int(kMaxTreeLevel) - int(__builtin_ctzll(k)) / 3 : int(kMaxTreeLevel); };
    std::vector<KeyType> cells;
    int firstDiff = (int(clz64(a ^ b)) + 3 - 1) / 3; // first octal place where a and b differ
    int aLast = lastNz(a), bLast = lastNz(b);
    for (int pos = aLast; pos > firstDiff; --pos) // climb: finish the octants a starts inside
        for (unsigned k = (8 - digit(a, pos)) % 8; k > 0; --k, a += power(pos))
            cells.push_back(a);
    for (int pos = firstDiff; pos <= bLast; ++pos) // descend towards b
        for (int k = int(digit(b, pos)) - int(digit(a, pos)); k > 0; --k, a += power(pos))
            cells.push_back(a);
    cells.push_back(b);
    return cells;
}

} // namespace

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_assemble_cuboid(sphb200_ctx* ctx, uint64_t keyStart, uint64_t keyEnd, const sphb200_box* box,
                                       const int* multiplicity, const double* xBlock, const double* yBlock,
                                       const double* zBlock, size_t blockSize, double* x, double* y, double* z,
                                       size_t capacity, uint64_t* count_h)
{
    try
    {
        *count_h = 0;
        if (keyEnd <= keyStart || blockSize == 0) return 0;
        const int* m = multiplicity;
        if (m[0] < 1 || m[1] < 1 || m[2] < 1) throw std::runtime_error("assemble_cuboid: multiplicity must be >= 1");
        std::vector<KeyType> cells = spanKeyRange(keyStart, keyEnd);

        constexpr int maxCoord = 1 << kMaxTreeLevel;
        constexpr double uL    = 1.0 / maxCoord;
        std::vector<AssemblePair> pairs;
        for (size_t c = 0; c + 1 < cells.size(); ++c)
        {
            // hilbertIBox (hilbert.hpp:264-281) of the cell
            unsigned level      = treeLevel(cells[c + 1] - cells[c]);
            unsigned cubeLength = (1u << kMaxTreeLevel) >> level;
            unsigned lo[3];
            hilbertDecode(cells[c], lo[0], lo[1], lo[2]);
            double sel[6], unit[6];
            for (int d = 0; d < 3; ++d)
            {
                lo[d] &= ~(cubeLength - 1);
                int imin = int(lo[d]), imax = int(lo[d] + cubeLength);
                // centerAndSize + createFpBox (box.hpp:334-370) in the global box and in the unit cube
                double half   = 0.5 * uL * box->len[d];
                double center = box->lim[2 * d] + (imax + imin) * half;
                double size   = (imax - imin) * half;
                sel[2 * d]     = center - size;
                sel[2 * d + 1] = center + size;
                double halfU   = 0.5 * uL * 1.0;
                double centerU = 0.0 + (imax + imin) * halfU;
                double sizeU   = (imax - imin) * halfU;
                unit[2 * d]     = centerU - sizeU;
                unit[2 * d + 1] = centerU + sizeU;
            }
            // gridIntersection (grid.hpp:141-151)
            int first[3], last[3];
            for (int d = 0; d < 3; ++d)
            {
                first[d] = int(unit[2 * d] * m[d]);
                last[d]  = int(std::ceil(unit[2 * d + 1] * m[d]));
            }
            for (int ix = first[0]; ix < last[0]; ++ix)
                for (int iy = first[1]; iy < last[1]; ++iy)
                    for (int iz = first[2]; iz < last[2]; ++iz)
                    {
                        AssemblePair p;
                        std::copy(sel, sel + 6, p.sel);
                        int g[3] = {ix, iy, iz};
                        for (int d = 0; d < 3; ++d)
                            p.origin[d] = box->lim[2 * d] + g[d] * box->len[d] / m[d];
                        pairs.push_back(p);
                    }
        }
        if (pairs.empty()) return 0;

        uint32_t numChunks = uint32_t((blockSize + kAssembleChunk - 1) / kAssembleChunk);
        size_t numBlocks   = pairs.size() * numChunks;
        if (numBlocks > 0x7fffffffull) throw std::runtime_error("assemble_cuboid: too many (cell, tile) pairs");
        size_t pairBytes = (pairs.size() * sizeof(AssemblePair) + 255) & ~size_t(255);
        char* tmp        = static_cast<char*>(ctx->miscTmp.get(pairBytes + (2 * numBlocks + 2) * sizeof(uint32_t)));
        auto* dPairs     = reinterpret_cast<AssemblePair*>(tmp);
        auto* counts     = reinterpret_cast<uint32_t*>(tmp + pairBytes);
        uint32_t* offs   = counts + numBlocks;
        uint32_t* total  = offs + numBlocks;
        SPHB200_CUDA(cudaMemcpyAsync(dPairs, pairs.data(), pairs.size() * sizeof(AssemblePair), cudaMemcpyHostToDevice,
                                     ctx->stream));
        double sx = box->len[0] / m[0], sy = box->len[1] / m[1], sz = box->len[2] / m[2];
        SPHB200_LAUNCH(ctx, assembleKernel<false>, unsigned(numBlocks), kAssembleChunk, 0, dPairs, numChunks, sx, sy, sz,
                       xBlock, yBlock, zBlock, blockSize, counts, (const uint32_t*)nullptr, (double*)nullptr,
                       (double*)nullptr, (double*)nullptr);
        exclusiveScan(ctx, counts, offs, numBlocks, total);
        uint32_t total_h = 0;
        SPHB200_CUDA(cudaMemcpyAsync(&total_h, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream)); // also keeps `pairs` alive until the upload is done
        *count_h = total_h;
        if (total_h > capacity) return SPHB200_ERR_CAPACITY;
        if (total_h)
            SPHB200_LAUNCH(ctx, assembleKernel<true>, unsigned(numBlocks), kAssembleChunk, 0, dPairs, numChunks, sx, sy,
                           sz, xBlock, yBlock, zBlock, blockSize, counts, offs, x, y, z);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_cut_sphere(sphb200_ctx* ctx, double radius, double* x, double* y, double* z, size_t n,
                                  uint64_t* count_h)
{
    try
    {
        *count_h = 0;
        if (n == 0) return 0;
        size_t numBlocks = (n + kAssembleChunk - 1) / kAssembleChunk;
        size_t coordBytes = (n * sizeof(double) + 255) & ~size_t(255);
        char* tmp        = static_cast<char*>(ctx->miscTmp.get(3 * coordBytes + (2 * numBlocks + 2) * sizeof(uint32_t)));
        double* ox       = reinterpret_cast<double*>(tmp);
        double* oy       = reinterpret_cast<double*>(tmp + coordBytes);
        double* oz       = reinterpret_cast<double*>(tmp + 2 * coordBytes);
        auto* counts     = reinterpret_cast<uint32_t*>(tmp + 3 * coordBytes);
        uint32_t* offs   = counts + numBlocks;
        uint32_t* total  = offs + numBlocks;
        SPHB200_LAUNCH(ctx, cutSphereKernel<false>, unsigned(numBlocks), kAssembleChunk, 0, radius, x, y, z, n, counts,
                       (const uint32_t*)nullptr, (double*)nullptr, (double*)nullptr, (double*)nullptr);
        exclusiveScan(ctx, counts, offs, numBlocks, total);
        SPHB200_LAUNCH(ctx, cutSphereKernel<true>, unsigned(numBlocks), kAssembleChunk, 0, radius, x, y, z, n, counts,
                       offs, ox, oy, oz);
        uint32_t total_h = 0;
        SPHB200_CUDA(cudaMemcpyAsync(&total_h, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        *count_h = total_h;
        if (total_h)
        {
            SPHB200_CUDA(cudaMemcpyAsync(x, ox, total_h * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
            SPHB200_CUDA(cudaMemcpyAsync(y, oy, total_h * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
            SPHB200_CUDA(cudaMemcpyAsync(z, oz, total_h * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_span_sfc_range(uint64_t keyStart, uint64_t keyEnd, uint64_t* cells_h, size_t capacity)
{
    try
    {
        if (keyEnd <= keyStart) return 0;
        std::vector<KeyType> cells = spanKeyRange(keyStart, keyEnd);
        if (cells.size() > capacity) throw std::runtime_error("span_sfc_range: capacity too small");
        std::copy(cells.begin(), cells.end(), cells_h);
        return int(cells.size()) - 1;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return -1;
    }
}

extern "C" int sphb200_hilbert_table(uint16_t* table_h, size_t capacity)
{
    try
    {
        static const std::vector<uint16_t> table = buildHilbertTable();
        if (capacity < table.size()) throw std::runtime_error("hilbert_table: capacity < 24 * 512");
        std::copy(table.begin(), table.end(), table_h);
        return int(kHilbertStates);
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return -1;
    }
}

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
        if (n < size_t(1) << 16)
        {
            SPHB200_LAUNCH(ctx, sfcKeysBitwiseKernel, persistentGrid(n, 256, 16), 256, 0, x, y, z, n, p, keys);
            return 0;
        }
        if (!ctx->hilbertTabReady)
        {
            static const std::vector<uint16_t> table = buildHilbertTable();
            SPHB200_CUDA(cudaMemcpyToSymbolAsync(gHilbertTab, table.data(), kHilbertTabSize * sizeof(uint16_t), 0,
                                                 cudaMemcpyHostToDevice, ctx->stream));
            ctx->hilbertTabReady = true;
        }
        SPHB200_LAUNCH(ctx, sfcKeysKernel, persistentGrid(n, 256, 8), 256, 0, x, y, z, n, p, keys);
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
