// Primitives: stable LSD radix sort (key u64, payload u32), exclusive scan, multi-field gather,
// range pack/unpack.  Hand-written for sm_100a; replaces the reference's cub/thrust calls
// (domain/include/cstone/primitives/primitives_gpu.cu:85-101,228-334).
#include "common.cuh"
#include "primitives.cuh"

namespace sphb200
{

// ---------------------------------------------------------------- exclusive scan (uint32)
constexpr int kScanThreads = 512;
constexpr int kScanItems   = 8;
constexpr int kScanTile    = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t warpInclusiveScan(uint32_t v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        uint32_t t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

//! @brief block-wide exclusive scan of one value per thread; returns exclusive prefix, total in *blockTotal
__device__ __forceinline__ uint32_t blockExclusiveScan(uint32_t v, uint32_t* smem /*>= 33*/, uint32_t* blockTotal)
{
    int lane     = threadIdx.x & 31;
    int warp     = threadIdx.x >> 5;
    int numWarps = blockDim.x >> 5;
    uint32_t inc = warpInclusiveScan(v, lane);
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0)
    {
        uint32_t w  = (lane < numWarps) ? smem[lane] : 0;
        uint32_t wi = warpInclusiveScan(w, lane);
        smem[lane]  = wi - w;
        if (lane == 31) smem[32] = wi;
    }
    __syncthreads();
    uint32_t res = smem[warp] + inc - v;
    *blockTotal  = smem[32];
    __syncthreads();
    return res;
}

__global__ void __launch_bounds__(kScanThreads) scanTileSumsKernel(const uint32_t* __restrict__ in, size_t n,
                                                                   uint32_t* __restrict__ tileSums)
{
    __shared__ uint32_t smem[33];
    size_t base  = size_t(blockIdx.x) * kScanTile + size_t(threadIdx.x) * kScanItems;
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < n) sum += in[base + k];
    uint32_t total;
    blockExclusiveScan(sum, smem, &total);
    if (threadIdx.x == 0) tileSums[blockIdx.x] = total;
}

//! @brief scans each tile; adds tileOffsets[tile] if given; in == out allowed
__global__ void __launch_bounds__(kScanThreads) scanTilesKernel(const uint32_t* __restrict__ in, size_t n,
                                                                const uint32_t* __restrict__ tileOffsets,
                                                                uint32_t* __restrict__ out, uint32_t* totalOut)
{
    __shared__ uint32_t smem[33];
    size_t base = size_t(blockIdx.x) * kScanTile + size_t(threadIdx.x) * kScanItems;
    uint32_t v[kScanItems];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
    {
        v[k] = (base + k < n) ? in[base + k] : 0;
        sum += v[k];
    }
    uint32_t total;
    uint32_t excl = blockExclusiveScan(sum, smem, &total);
    uint32_t off  = excl + (tileOffsets ? tileOffsets[blockIdx.x] : 0);
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
    {
        if (base + k < n) out[base + k] = off;
        off += v[k];
    }
    if (totalOut && blockIdx.x == gridDim.x - 1 && threadIdx.x == blockDim.x - 1) *totalOut = off;
}

void exclusiveScan(sphb200_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n, uint32_t* totalOut)
{
    if (n == 0)
    {
        if (totalOut) SPHB200_CUDA(cudaMemsetAsync(totalOut, 0, sizeof(uint32_t), ctx->stream));
        return;
    }
    size_t numTiles = (n + kScanTile - 1) / kScanTile;
    if (numTiles == 1)
    {
        SPHB200_LAUNCH(ctx, scanTilesKernel, 1, kScanThreads, 0, in, n, (const uint32_t*)nullptr, out, totalOut);
        return;
    }
    // tile sums live in scanTmp: level 0 needs numTiles entries, deeper levels stacked behind
    size_t need = 0;
    for (size_t t = numTiles; t > 1; t = (t + kScanTile - 1) / kScanTile)
        need += t + 1;
    need += 2;
    uint32_t* tmp = static_cast<uint32_t*>(ctx->scanTmp.get(need * sizeof(uint32_t)));

    // recursive lambda via explicit level loop
    std::vector<std::pair<uint32_t*, size_t>> levels; // (sums array, count)
    uint32_t* cursor = tmp;
    size_t cnt       = numTiles;
    const uint32_t* src = in;
    size_t srcN         = n;
    while (true)
    {
        SPHB200_LAUNCH(ctx, scanTileSumsKernel, int(cnt), kScanThreads, 0, src, srcN, cursor);
        levels.push_back({cursor, cnt});
        if (cnt <= size_t(kScanTile)) break;
        src  = cursor;
        srcN = cnt;
        cursor += cnt + 1;
        cnt = (cnt + kScanTile - 1) / kScanTile;
    }
    // scan top level in place (single tile)
    {
        auto [arr, c] = levels.back();
        SPHB200_LAUNCH(ctx, scanTilesKernel, 1, kScanThreads, 0, arr, c, (const uint32_t*)nullptr, arr,
                       (uint32_t*)nullptr);
    }
    for (int l = int(levels.size()) - 2; l >= 0; --l)
    {
        auto [arr, c]   = levels[l];
        auto [offs, oc] = levels[l + 1];
        SPHB200_LAUNCH(ctx, scanTilesKernel, int(oc), kScanThreads, 0, arr, c, offs, arr, (uint32_t*)nullptr);
    }
    SPHB200_LAUNCH(ctx, scanTilesKernel, int(numTiles), kScanThreads, 0, in, n, levels[0].first, out, totalOut);
}

// ---------------------------------------------------------------- radix sort
constexpr int kRadixBits   = 8;
constexpr int kRadix       = 1 << kRadixBits;
constexpr int kSortTile    = 4096; // keys per CTA, count and scatter kernels alike
constexpr int kCountThreads = 256;
constexpr int kCountItems   = kSortTile / kCountThreads;
constexpr int kSortThreads = 512;
constexpr int kSortWarps   = kSortThreads / 32;
constexpr int kSortItems   = kSortTile / kSortThreads; // keys per thread
static_assert(kCountThreads == kRadix, "one histogram bin per thread");

// shared memory of the scatter kernel (dynamic, above the 48 KB static limit)
constexpr size_t kStageKeysOff = 0;
constexpr size_t kStageValsOff = kStageKeysOff + kSortTile * sizeof(KeyType);
constexpr size_t kWarpHistOff  = kStageValsOff + kSortTile * sizeof(uint32_t);
constexpr size_t kBinDeltaOff  = kWarpHistOff + kSortWarps * kRadix * sizeof(uint16_t);
constexpr size_t kScanOff      = kBinDeltaOff + kRadix * sizeof(uint32_t);
constexpr size_t kScatterSmem  = kScanOff + 36 * sizeof(uint32_t);

//! @brief per-tile digit histogram, written digit-major: hist[digit * numTiles + tile]
__global__ void __launch_bounds__(kCountThreads) radixCountKernel(const KeyType* __restrict__ keys, size_t n, int shift,
                                                                  uint32_t* __restrict__ hist, int numTiles)
{
    __shared__ uint32_t sh[kRadix];
    sh[threadIdx.x] = 0;
    __syncthreads();
    size_t base = size_t(blockIdx.x) * kSortTile;
    if (base + kSortTile <= n && (reinterpret_cast<uintptr_t>(keys) & 15) == 0)
    {
        // all loads of the tile in flight before the first shared-memory atomic
        const ulonglong2* k2 = reinterpret_cast<const ulonglong2*>(keys + base);
        ulonglong2 v[kCountItems / 2];
#pragma unroll
        for (int k = 0; k < kCountItems / 2; ++k)
            v[k] = k2[k * kCountThreads + threadIdx.x];
#pragma unroll
        for (int k = 0; k < kCountItems / 2; ++k)
        {
            atomicAdd(&sh[(v[k].x >> shift) & (kRadix - 1)], 1u);
            atomicAdd(&sh[(v[k].y >> shift) & (kRadix - 1)], 1u);
        }
    }
    else
    {
        for (int k = 0; k < kCountItems; ++k)
        {
            size_t i = base + size_t(k) * kCountThreads + threadIdx.x;
            if (i < n) atomicAdd(&sh[(keys[i] >> shift) & (kRadix - 1)], 1u);
        }
    }
    __syncthreads();
    hist[size_t(threadIdx.x) * numTiles + blockIdx.x] = sh[threadIdx.x];
}

/*! @brief stable scatter of one tile, staged through shared memory
 *
 * Order inside a tile is (warp, chunk, lane); each warp ranks its 32-key chunks with
 * __match_any_sync (peers with the same digit), keeping a private digit counter row in
 * shared memory, so no atomics are needed and the ranking is deterministic and stable.
 *
 * Keys and payload are then placed at their slot of the digit-sorted tile in shared memory, and slot s
 * goes to binStart[d] + (s - tileStart[d]): consecutive threads write consecutive addresses inside a
 * digit's run (4096 keys over 256 digits: runs of ~16 keys = 128 B) instead of 32 scattered 8-byte
 * stores per warp instruction.  Full = the tile has all kSortTile keys (every tile but the last).
 */
template<bool Full>
__device__ __forceinline__ void radixScatterTile(const KeyType* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                 KeyType* __restrict__ keysOut, uint32_t* __restrict__ valsOut, size_t n,
                                                 int shift, const uint32_t* __restrict__ rowScan,
                                                 const uint32_t* __restrict__ rowTotal, int numTiles)
{
    extern __shared__ __align__(16) unsigned char sortSmem[];
    KeyType* stageK    = reinterpret_cast<KeyType*>(sortSmem + kStageKeysOff);
    uint32_t* stageV   = reinterpret_cast<uint32_t*>(sortSmem + kStageValsOff);
    uint16_t* warpHist = reinterpret_cast<uint16_t*>(sortSmem + kWarpHistOff); // [kSortWarps][kRadix]
    uint32_t* binDelta = reinterpret_cast<uint32_t*>(sortSmem + kBinDeltaOff); // global run start - start in the tile
    uint32_t* scanSmem = reinterpret_cast<uint32_t*>(sortSmem + kScanOff);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int tile = blockIdx.x;
    // this tile's offset inside the run of digit threadIdx.x and the size of the whole run: needed after the ranking
    uint32_t inRow = 0, rowSize = 0;
    if (threadIdx.x < kRadix)
    {
        inRow   = rowScan[size_t(threadIdx.x) * numTiles + tile];
        rowSize = rowTotal[threadIdx.x];
    }
    {
        uint32_t* z = reinterpret_cast<uint32_t*>(warpHist);
        for (int k = threadIdx.x; k < kSortWarps * kRadix / 2; k += kSortThreads)
            z[k] = 0;
    }

    const size_t tileBase    = size_t(tile) * kSortTile;
    const uint32_t tileCount = Full ? uint32_t(kSortTile) : uint32_t(n - tileBase);
    const uint32_t warpOff   = uint32_t(warp) * (32 * kSortItems) + lane;
    const KeyType* kIn       = keysIn + tileBase;
    KeyType key[kSortItems];
    uint32_t val[kSortItems];
    uint16_t rank[kSortItems];

#pragma unroll
    for (int c = 0; c < kSortItems; ++c)
    {
        uint32_t e = warpOff + c * 32;
        key[c]     = (Full || e < tileCount) ? kIn[e] : ~KeyType(0);
    }
#pragma unroll
    for (int c = 0; c < kSortItems; ++c)
    {
        uint32_t e = warpOff + c * 32;
        if (valsIn) { val[c] = (Full || e < tileCount) ? valsIn[tileBase + e] : 0u; }
        else { val[c] = uint32_t(tileBase) + e; }
    }
    __syncthreads();

    const unsigned ltMask = (1u << lane) - 1;
    uint16_t* myHist      = warpHist + warp * kRadix;
#pragma unroll
    for (int c = 0; c < kSortItems; ++c)
    {
        bool valid    = Full || (warpOff + c * 32 < tileCount);
        unsigned d    = unsigned(key[c] >> shift) & (kRadix - 1);
        // lanes holding the same digit: one ballot per digit bit (__match_any_sync serialises over the distinct
        // values of the warp, ~110 cycles of the ADU pipe when all 32 digits differ)
        unsigned peer = Full ? 0xffffffffu : __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < kRadixBits; ++b)
        {
            bool bit   = (d >> b) & 1u;
            unsigned m = __ballot_sync(0xffffffffu, bit);
            peer &= bit ? m : ~m;
        }
        // the first lane of every peer group bumps the warp's counter of that digit; full-mask shuffle (a shuffle
        // under the peer masks would be issued once per group)
        int leader    = valid ? __ffs(peer) - 1 : lane;
        uint32_t base = 0;
        if (valid && lane == leader)
        {
            base      = myHist[d];
            myHist[d] = uint16_t(base + __popc(peer));
        }
        base    = __shfl_sync(0xffffffffu, base, leader);
        rank[c] = uint16_t(base + __popc(peer & ltMask));
        __syncwarp();
    }
    __syncthreads();
    {
        // digit = threadIdx.x: exclusive scan over the warps of this tile, on top of the digit's start inside the tile
        int d          = threadIdx.x;
        uint32_t total = 0;
        if (d < kRadix)
        {
#pragma unroll
            for (int w = 0; w < kSortWarps; ++w)
                total += warpHist[w * kRadix + d];
        }
        uint32_t blockTotal;
        uint32_t tileStart = blockExclusiveScan(total, scanSmem, &blockTotal);
        uint32_t digitBase = blockExclusiveScan(rowSize, scanSmem, &blockTotal); // start of the digit's run in the output
        if (d < kRadix)
        {
            uint32_t running = tileStart;
#pragma unroll
            for (int w = 0; w < kSortWarps; ++w)
            {
                uint32_t t               = warpHist[w * kRadix + d];
                warpHist[w * kRadix + d] = uint16_t(running);
                running += t;
            }
            binDelta[d] = digitBase + inRow - tileStart;
        }
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < kSortItems; ++c)
    {
        if (Full || (warpOff + c * 32 < tileCount))
        {
            unsigned d    = unsigned(key[c] >> shift) & (kRadix - 1);
            uint32_t slot = uint32_t(myHist[d]) + rank[c];
            stageK[slot]  = key[c];
            stageV[slot]  = val[c];
        }
    }
    __syncthreads();
#pragma unroll
    for (int c = 0; c < kSortItems; ++c)
    {
        uint32_t sl = uint32_t(c) * kSortThreads + threadIdx.x;
        if (Full || sl < tileCount)
        {
            KeyType k    = stageK[sl];
            uint32_t pos = binDelta[unsigned(k >> shift) & (kRadix - 1)] + sl;
            keysOut[pos] = k;
            valsOut[pos] = stageV[sl];
        }
    }
}

__global__ void __launch_bounds__(kSortThreads, 2)
    radixScatterKernel(const KeyType* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                       KeyType* __restrict__ keysOut, uint32_t* __restrict__ valsOut, size_t n, int shift,
                       const uint32_t* __restrict__ rowScan, const uint32_t* __restrict__ rowTotal, int numTiles)
{
    // the last tile may be ragged: same launch, predicated body
    if (size_t(blockIdx.x + 1) * kSortTile <= n)
        radixScatterTile<true>(keysIn, valsIn, keysOut, valsOut, n, shift, rowScan, rowTotal, numTiles);
    else
        radixScatterTile<false>(keysIn, valsIn, keysOut, valsOut, n, shift, rowScan, rowTotal, numTiles);
}

/*! @brief exclusive scan of every digit row of the tile histogram (in place) and the row totals
 *
 * hist is digit-major [kRadix][numTiles]: one CTA per digit.  The scatter kernel turns the 256 row totals into the
 * digit bases itself, so a pass needs three launches (count, row scan, scatter).
 */
__global__ void __launch_bounds__(kScanThreads) radixRowScanKernel(uint32_t* __restrict__ hist, int numTiles,
                                                                   uint32_t* __restrict__ rowTotal)
{
    __shared__ uint32_t smem[33];
    uint32_t* row    = hist + size_t(blockIdx.x) * numTiles;
    uint32_t carry   = 0;
    for (int base = 0; base < numTiles; base += kScanTile)
    {
        int i0 = base + threadIdx.x * kScanItems;
        uint32_t v[kScanItems];
        uint32_t sum = 0;
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
        {
            v[k] = (i0 + k < numTiles) ? row[i0 + k] : 0;
            sum += v[k];
        }
        uint32_t total;
        uint32_t off = carry + blockExclusiveScan(sum, smem, &total);
#pragma unroll
        for (int k = 0; k < kScanItems; ++k)
        {
            if (i0 + k < numTiles) row[i0 + k] = off;
            off += v[k];
        }
        carry += total;
    }
    if (threadIdx.x == 0) rowTotal[blockIdx.x] = carry;
}

void radixSortPairs(sphb200_ctx* ctx, KeyType* keys, uint32_t* vals, size_t n, bool valsAreIota, int beginBit,
                    int endBit)
{
    if (n == 0) return;
    // per device and cheap: no caching
    SPHB200_CUDA(cudaFuncSetAttribute(radixScatterKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kScatterSmem)));
    int numTiles    = int((n + kSortTile - 1) / kSortTile);
    size_t histSize = size_t(kRadix) * numTiles;
    // temp: alt keys, alt vals, hist, row totals
    size_t bytes = n * sizeof(KeyType) + n * sizeof(uint32_t) + (histSize + kRadix) * sizeof(uint32_t) + 512;
    char* tmp    = static_cast<char*>(ctx->sortTmp.get(bytes));
    KeyType* keysAlt  = reinterpret_cast<KeyType*>(tmp);
    uint32_t* valsAlt = reinterpret_cast<uint32_t*>(tmp + n * sizeof(KeyType));
    uint32_t* hist    = reinterpret_cast<uint32_t*>(tmp + n * sizeof(KeyType) + ((n * sizeof(uint32_t) + 255) & ~size_t(255)));
    uint32_t* rowTotal = hist + histSize;

    int numPasses = (endBit - beginBit + kRadixBits - 1) / kRadixBits;
    KeyType* kIn  = keys;
    KeyType* kOut = keysAlt;
    uint32_t* vIn = vals;
    uint32_t* vOut = valsAlt;
    for (int p = 0; p < numPasses; ++p)
    {
        int shift = beginBit + p * kRadixBits;
        SPHB200_LAUNCH(ctx, radixCountKernel, numTiles, kCountThreads, 0, kIn, n, shift, hist, numTiles);
        SPHB200_LAUNCH(ctx, radixRowScanKernel, kRadix, kScanThreads, 0, hist, numTiles, rowTotal);
        const uint32_t* vSrc = (p == 0 && valsAreIota) ? nullptr : vIn;
        SPHB200_LAUNCH(ctx, radixScatterKernel, numTiles, kSortThreads, kScatterSmem, kIn, vSrc, kOut, vOut, n, shift,
                       hist, rowTotal, numTiles);
        std::swap(kIn, kOut);
        std::swap(vIn, vOut);
    }
    if (kIn != keys)
    {
        SPHB200_CUDA(cudaMemcpyAsync(keys, kIn, n * sizeof(KeyType), cudaMemcpyDeviceToDevice, ctx->stream));
        SPHB200_CUDA(cudaMemcpyAsync(vals, vIn, n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, ctx->stream));
    }
}

// ---------------------------------------------------------------- multi-field gather
constexpr int kMaxGatherFields = 24;
struct GatherArgs
{
    const void* src[kMaxGatherFields];
    void* dst[kMaxGatherFields];
    unsigned char is8[kMaxGatherFields];
    int nfields;
};

/*! @brief dst_f[i] = src_f[perm[i]] for all fields in one launch
 *
 * perm is read once per particle (the reference launches one gather per field and re-reads it,
 * primitives_gpu.cu:85-101); all random reads of a particle are issued back-to-back for MLP.
 */
__global__ void __launch_bounds__(256) gatherFieldsKernel(const uint32_t* __restrict__ perm, size_t n, GatherArgs a)
{
    size_t stride = size_t(gridDim.x) * blockDim.x;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    {
        uint32_t s = perm[i];
#pragma unroll 4
        for (int f = 0; f < a.nfields; ++f)
        {
            if (a.is8[f]) { static_cast<uint64_t*>(a.dst[f])[i] = __ldg(static_cast<const uint64_t*>(a.src[f]) + s); }
            else { static_cast<uint32_t*>(a.dst[f])[i] = __ldg(static_cast<const uint32_t*>(a.src[f]) + s); }
        }
    }
}

// ---------------------------------------------------------------- range pack / unpack (halo buffers)
struct RangeArgs
{
    void* field[kMaxGatherFields];
    unsigned char is8[kMaxGatherFields];
    size_t packedOffset[kMaxGatherFields]; // byte offset of field f inside the packed buffer
    int nfields;
};

/*! @brief element e of the concatenated ranges <-> array index rangeBegin[r] + (e - rangeScan[r])
 *  (gatherRangesKernel / scatterRangesKernel, halos/gather_halos_gpu.cu:41-117; here all fields per launch) */
template<bool Pack>
__global__ void __launch_bounds__(256) rangesKernel(const uint32_t* __restrict__ rangeScan,
                                                    const uint32_t* __restrict__ rangeBegin, int numRanges,
                                                    uint32_t total, RangeArgs a, char* __restrict__ packed)
{
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    // last range with rangeScan[r] <= e
    int r        = int(upperBound(rangeScan, size_t(numRanges), e)) - 1;
    uint32_t idx = rangeBegin[r] + (e - rangeScan[r]);
    for (int f = 0; f < a.nfields; ++f)
    {
        if (a.is8[f])
        {
            auto* p = reinterpret_cast<uint64_t*>(packed + a.packedOffset[f]);
            auto* q = static_cast<uint64_t*>(a.field[f]);
            if (Pack) p[e] = q[idx];
            else q[idx] = p[e];
        }
        else
        {
            auto* p = reinterpret_cast<uint32_t*>(packed + a.packedOffset[f]);
            auto* q = static_cast<uint32_t*>(a.field[f]);
            if (Pack) p[e] = q[idx];
            else q[idx] = p[e];
        }
    }
}

constexpr int kMaxPeers = 64;

struct PeerArgs
{
    uint32_t scan[kMaxPeers + 1]; // element prefix over peers
    size_t byteBase[kMaxPeers];   // start of each peer's block in the packed buffer
    int numPeers;
};

/*! @brief all peers of one halo exchange in one launch
 *
 * Packed buffer = per-peer blocks in peer order (what all_to_all_single splits on); inside a block the fields follow
 * one another, each 8-byte aligned (exchange_halos_gpu.cuh:58).  Element e of the concatenated ranges belongs to the
 * peer q with scan[q] <= e < scan[q+1].
 */
template<bool Pack>
__global__ void __launch_bounds__(256) rangesMultiKernel(const uint32_t* __restrict__ rangeScan,
                                                         const uint32_t* __restrict__ rangeBegin, int numRanges,
                                                         uint32_t total, PeerArgs peers, RangeArgs a,
                                                         char* __restrict__ packed)
{
    uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    int r        = int(upperBound(rangeScan, size_t(numRanges), e)) - 1;
    uint32_t idx = rangeBegin[r] + (e - rangeScan[r]);
    int q = 0;
    while (q + 1 < peers.numPeers && peers.scan[q + 1] <= e)
        ++q;
    uint32_t cnt = peers.scan[q + 1] - peers.scan[q];
    uint32_t el  = e - peers.scan[q];
    size_t off   = peers.byteBase[q];
    for (int f = 0; f < a.nfields; ++f)
    {
        if (a.is8[f])
        {
            auto* p = reinterpret_cast<uint64_t*>(packed + off);
            auto* q8 = static_cast<uint64_t*>(a.field[f]);
            if (Pack) p[el] = q8[idx];
            else q8[idx] = p[el];
            off += size_t(cnt) * 8;
        }
        else
        {
            auto* p = reinterpret_cast<uint32_t*>(packed + off);
            auto* q4 = static_cast<uint32_t*>(a.field[f]);
            if (Pack) p[el] = q4[idx];
            else q4[idx] = p[el];
            off += (size_t(cnt) * 4 + 7) & ~size_t(7);
        }
    }
}

} // namespace sphb200

using namespace sphb200;

static int rangesMultiImpl(sphb200_ctx* ctx, bool pack, const uint32_t* rangeScan, const uint32_t* rangeBegin,
                           int numRanges, int numPeers, const uint32_t* peerCount_h, int nfields, void* const* fields,
                           const int* elemBytes, void* packed, size_t* bytesPerPeer_h)
{
    try
    {
        if (nfields > kMaxGatherFields) throw std::runtime_error("ranges_multi: too many fields");
        if (numPeers > kMaxPeers) throw std::runtime_error("ranges_multi: too many peers");
        PeerArgs peers;
        peers.numPeers = numPeers;
        peers.scan[0]  = 0;
        size_t base    = 0;
        for (int q = 0; q < numPeers; ++q)
        {
            peers.scan[q + 1] = peers.scan[q] + peerCount_h[q];
            peers.byteBase[q] = base;
            size_t blk        = 0;
            for (int f = 0; f < nfields; ++f)
                blk += (size_t(peerCount_h[q]) * elemBytes[f] + 7) & ~size_t(7);
            if (bytesPerPeer_h) bytesPerPeer_h[q] = blk;
            base += blk;
        }
        uint32_t total = peers.scan[numPeers];
        if (total == 0 || nfields == 0 || !packed) return 0;
        RangeArgs a;
        a.nfields = nfields;
        for (int f = 0; f < nfields; ++f)
        {
            a.field[f]        = fields[f];
            a.is8[f]          = elemBytes[f] == 8;
            a.packedOffset[f] = 0;
        }
        if (pack)
            SPHB200_LAUNCH(ctx, rangesMultiKernel<true>, gridFor(total, 256), 256, 0, rangeScan, rangeBegin, numRanges,
                           total, peers, a, static_cast<char*>(packed));
        else
            SPHB200_LAUNCH(ctx, rangesMultiKernel<false>, gridFor(total, 256), 256, 0, rangeScan, rangeBegin, numRanges,
                           total, peers, a, static_cast<char*>(packed));
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_pack_ranges_peers(sphb200_ctx* ctx, const uint32_t* rangeScan, const uint32_t* rangeBegin,
                                         int numRanges, int numPeers, const uint32_t* peerCount_h, int nfields,
                                         const void* const* src, const int* elemBytes, void* packed,
                                         size_t* bytesPerPeer_h)
{
    return rangesMultiImpl(ctx, true, rangeScan, rangeBegin, numRanges, numPeers, peerCount_h, nfields,
                           const_cast<void* const*>(src), elemBytes, packed, bytesPerPeer_h);
}

extern "C" int sphb200_unpack_ranges_peers(sphb200_ctx* ctx, const uint32_t* rangeScan, const uint32_t* rangeBegin,
                                           int numRanges, int numPeers, const uint32_t* peerCount_h, int nfields,
                                           void* const* dst, const int* elemBytes, const void* packed,
                                           size_t* bytesPerPeer_h)
{
    ctx->packValid = 0; // halo values changed under the gather records (hydro.cu)
    return rangesMultiImpl(ctx, false, rangeScan, rangeBegin, numRanges, numPeers, peerCount_h, nfields, dst,
                           elemBytes, const_cast<void*>(packed), bytesPerPeer_h);
}

extern "C" int sphb200_sort_by_key(sphb200_ctx* ctx, uint64_t* keys, uint32_t* perm, size_t n)
{
    try
    {
        radixSortPairs(ctx, keys, perm, n, true, 0, 64);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_sort_pairs(sphb200_ctx* ctx, uint64_t* keys, uint32_t* values, size_t n)
{
    try
    {
        radixSortPairs(ctx, keys, values, n, false, 0, 64);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_gather(sphb200_ctx* ctx, const uint32_t* perm, size_t n, int nfields, const void* const* src,
                              void* const* dst, const int* elemBytes)
{
    try
    {
        if (n == 0 || nfields == 0) return 0;
        for (int f0 = 0; f0 < nfields; f0 += kMaxGatherFields)
        {
            GatherArgs a;
            a.nfields = std::min(kMaxGatherFields, nfields - f0);
            for (int f = 0; f < a.nfields; ++f)
            {
                if (elemBytes[f0 + f] != 4 && elemBytes[f0 + f] != 8) throw std::runtime_error("gather: element size must be 4 or 8");
                a.src[f] = src[f0 + f];
                a.dst[f] = dst[f0 + f];
                a.is8[f] = elemBytes[f0 + f] == 8;
            }
            SPHB200_LAUNCH(ctx, gatherFieldsKernel, persistentGrid(n, 256, 8), 256, 0, perm, n, a);
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

static int rangesImpl(sphb200_ctx* ctx, bool pack, const uint32_t* rangeScan, const uint32_t* rangeBegin, int numRanges,
                      uint32_t total, int nfields, void* const* fields, const int* elemBytes, void* packed)
{
    try
    {
        if (total == 0 || nfields == 0) return 0;
        if (nfields > kMaxGatherFields) throw std::runtime_error("pack_ranges: too many fields");
        RangeArgs a;
        a.nfields  = nfields;
        size_t off = 0;
        for (int f = 0; f < nfields; ++f)
        {
            a.field[f]        = fields[f];
            a.is8[f]          = elemBytes[f] == 8;
            a.packedOffset[f] = off;
            off += (size_t(total) * elemBytes[f] + 7) & ~size_t(7); // 8-byte aligned like exchange_halos_gpu.cuh:58
        }
        if (pack)
            SPHB200_LAUNCH(ctx, rangesKernel<true>, gridFor(total, 256), 256, 0, rangeScan, rangeBegin, numRanges, total,
                           a, static_cast<char*>(packed));
        else
            SPHB200_LAUNCH(ctx, rangesKernel<false>, gridFor(total, 256), 256, 0, rangeScan, rangeBegin, numRanges,
                           total, a, static_cast<char*>(packed));
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_pack_ranges(sphb200_ctx* ctx, const uint32_t* rangeScan, const uint32_t* rangeBegin,
                                   int numRanges, uint32_t total, int nfields, const void* const* src,
                                   const int* elemBytes, void* packed)
{
    return rangesImpl(ctx, true, rangeScan, rangeBegin, numRanges, total, nfields, const_cast<void* const*>(src),
                      elemBytes, packed);
}

extern "C" int sphb200_unpack_ranges(sphb200_ctx* ctx, const uint32_t* rangeScan, const uint32_t* rangeBegin,
                                     int numRanges, uint32_t total, int nfields, void* const* dst, const int* elemBytes,
                                     const void* packed)
{
    ctx->packValid = 0; // halo values changed under the gather records (hydro.cu)
    return rangesImpl(ctx, false, rangeScan, rangeBegin, numRanges, total, nfields, dst, elemBytes,
                      const_cast<void*>(packed));
}

// ---------------------------------------------------------------- scan without device allocation
// sphb200_layout_noalloc: the exclusive scan behind sphb200_layout for callers that must not allocate (CUDA graph
// capture, tight serving loops).  No counterpart in the reference.
// The tile sums of the two-level scan live in the file-scope device array below (16 KB + 8 bytes, held by every GPU
// that loads this library until exit) instead of the context's grow-only arena.  One level of tile sums covers
// kNoAllocMaxTiles tiles of kScanTile counts = 16 777 216 leaves; larger inputs are rejected.  Every entry of the
// scratch that a call reads is written by the same call first (scanTileSumsKernel fills tile sums [0, numTiles)
// before the top-level scan reads them), so nothing has to be cleared between calls; the array keeps the last call's
// contents and calls that overlap on the GPU (different streams or contexts of one device) race on it.
namespace sphb200
{
constexpr int kNoAllocMaxTiles       = kScanTile; // the top level is scanned by ONE tile
constexpr size_t kNoAllocMaxElements = size_t(kNoAllocMaxTiles) * kScanTile;
static __device__ __align__(256) uint32_t g_noAllocScanScratch[kNoAllocMaxTiles + 2];

constexpr int kNoAllocMaxDevices = 64;
//! @brief device address of g_noAllocScanScratch on `device` (looked up once per device with that device current);
//!        the symbol's name in host code is a host-side stand-in, never a device address
static uint32_t* noAllocScratch(int device)
{
    static uint32_t* addr[kNoAllocMaxDevices] = {};
    if (device < 0 || device >= kNoAllocMaxDevices) throw std::runtime_error("layout_noalloc: device id out of range");
    uint32_t* p = __atomic_load_n(&addr[device], __ATOMIC_ACQUIRE);
    if (p) return p;
    int current = -1;
    SPHB200_CUDA(cudaGetDevice(&current));
    if (current != device) SPHB200_CUDA(cudaSetDevice(device));
    void* sym       = nullptr;
    cudaError_t err = cudaGetSymbolAddress(&sym, g_noAllocScanScratch);
    if (current != device && current >= 0) cudaSetDevice(current);
    if (err != cudaSuccess || sym == nullptr)
        throw CudaError(std::string("layout_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(err));
    p = static_cast<uint32_t*>(sym);
    __atomic_store_n(&addr[device], p, __ATOMIC_RELEASE); // the address is a property of the device: racing setters agree
    return p;
}
} // namespace sphb200

extern "C" int sphb200_layout_noalloc(sphb200_ctx* ctx, const uint32_t* counts, int numLeaves, uint32_t* layout)
{
    using namespace sphb200;
    try
    {
        if (!ctx) throw std::runtime_error("layout_noalloc: no context");
        if (numLeaves < 0) throw std::runtime_error("layout_noalloc: numLeaves < 0");
        if (size_t(numLeaves) > kNoAllocMaxElements)
            throw std::runtime_error("layout_noalloc: numLeaves = " + std::to_string(numLeaves) +
                                     " exceeds the static scratch limit of " + std::to_string(kNoAllocMaxElements) +
                                     " (use sphb200_layout)");
        size_t n          = size_t(numLeaves);
        uint32_t* total   = layout + n;
        if (n == 0)
        {
            SPHB200_CUDA(cudaMemsetAsync(total, 0, sizeof(uint32_t), ctx->stream));
            return 0;
        }
        size_t numTiles = (n + kScanTile - 1) / kScanTile;
        if (numTiles == 1)
        {
            SPHB200_LAUNCH(ctx, scanTilesKernel, 1, kScanThreads, 0, counts, n, (const uint32_t*)nullptr, layout, total);
            return 0;
        }
        uint32_t* sums = noAllocScratch(ctx->device);
        SPHB200_LAUNCH(ctx, scanTileSumsKernel, int(numTiles), kScanThreads, 0, counts, n, sums);
        SPHB200_LAUNCH(ctx, scanTilesKernel, 1, kScanThreads, 0, sums, numTiles, (const uint32_t*)nullptr, sums,
                       (uint32_t*)nullptr);
        SPHB200_LAUNCH(ctx, scanTilesKernel, int(numTiles), kScanThreads, 0, counts, n, sums, layout, total);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}
