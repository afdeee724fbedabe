// Micro-benchmarks behind the non-HBM roofs quoted in bench.py / DESIGN.md.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o peaks peaks.cu && ./peaks > peaks.json
// 1. DFMA / FFMA issue peak of the CUDA cores (independent chains, all SMs busy).
// 2. Cost of the gather that the SPH pair kernels perform: every lane of a warp loads one 32-byte record
//    {x, y, z, m} of its k-th neighbor.  Index patterns: coalesced, broadcast, uniformly random in a window, and the
//    real thing -- ascending neighbor lists of 32 Morton-consecutive targets on a periodic 16^3 lattice (2h = 2.88
//    spacings, 92 neighbors).  Access variants: one 256-bit load, 2 x 128-bit, 4 x 64-bit of an AoS record, 4 x 64-bit
//    of SoA arrays, and the same records staged in shared memory (AoS 2 x LDS.128, SoA 4 x LDS.64).
//    Output: SM-cycles per warp-level record request at saturation = the L1 (or shared-memory) data-stage wavefronts
//    such a request costs, if that stage is what saturates.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

#define CK(x)                                                                                                         \
    do                                                                                                                \
    {                                                                                                                 \
        cudaError_t e = (x);                                                                                          \
        if (e != cudaSuccess)                                                                                         \
        {                                                                                                             \
            fprintf(stderr, "%s failed: %s\n", #x, cudaGetErrorString(e));                                            \
            exit(1);                                                                                                  \
        }                                                                                                             \
    } while (0)

template<class T>
__global__ void __launch_bounds__(256) fmaKernel(T* out, int iters, T a, T b)
{
    T c[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
        c[k] = T(threadIdx.x + k);
    for (int it = 0; it < iters; ++it)
    {
#pragma unroll
        for (int k = 0; k < 8; ++k)
            c[k] = c[k] * a + b;
    }
    T s = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        s += c[k];
    if (s == T(-1.2345)) out[0] = s;
}

struct alignas(32) Rec
{
    double a, b, c, d;
};

enum Mode { LDG256, LDG128x2, LDG64x4, SOA64x4, LDS128x2, LDS_SOA64x4, kNumModes };
static const char* modeName[] = {"ldg256_aos", "ldg128x2_aos", "ldg64x4_aos", "ldg64x4_soa", "lds128x2_aos", "lds64x4_soa"};

template<int MODE>
__global__ void __launch_bounds__(1024, 1)
    gatherKernel(const Rec* __restrict__ recs, const double* __restrict__ soa, int numRecs, const uint32_t* __restrict__ idx,
                 int numSlabs, int listLen, int repeats, double* out)
{
    extern __shared__ __align__(128) unsigned char smemRaw[];
    if (MODE == LDS128x2)
    {
        Rec* s = reinterpret_cast<Rec*>(smemRaw);
        for (int i = threadIdx.x; i < numRecs; i += blockDim.x)
            s[i] = recs[i];
        __syncthreads();
    }
    if (MODE == LDS_SOA64x4)
    {
        double* s = reinterpret_cast<double*>(smemRaw);
        for (int i = threadIdx.x; i < 4 * numRecs; i += blockDim.x)
            s[i] = soa[i];
        __syncthreads();
    }
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int numWarps = (gridDim.x * blockDim.x) >> 5;
    double acc0 = 0, acc1 = 0, acc2 = 0, acc3 = 0;
    for (int r = 0; r < repeats; ++r)
        for (int slab = warp % numSlabs, cnt = 0; cnt < (numSlabs + numWarps - 1) / numWarps + 1; ++cnt, slab = (slab + numWarps) % numSlabs)
        {
            const uint32_t* l = idx + size_t(slab) * listLen * 32 + lane;
            uint32_t jn = __ldcs(l);
            for (int k = 0; k < listLen; ++k)
            {
                uint32_t j = jn;
                if (k + 1 < listLen) jn = __ldcs(l + size_t(k + 1) * 32);
                double a, b, c, d;
                if (MODE == LDG256)
                {
                    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(recs + j));
                }
                else if (MODE == LDG128x2)
                {
                    const double2* p = reinterpret_cast<const double2*>(recs + j);
                    double2 u = __ldg(p), v = __ldg(p + 1);
                    a = u.x; b = u.y; c = v.x; d = v.y;
                }
                else if (MODE == LDG64x4)
                {
                    const double* p = reinterpret_cast<const double*>(recs + j);
                    a = __ldg(p); b = __ldg(p + 1); c = __ldg(p + 2); d = __ldg(p + 3);
                }
                else if (MODE == SOA64x4)
                {
                    a = __ldg(soa + j); b = __ldg(soa + numRecs + j); c = __ldg(soa + 2 * numRecs + j); d = __ldg(soa + 3 * numRecs + j);
                }
                else if (MODE == LDS128x2)
                {
                    const double2* p = reinterpret_cast<const double2*>(reinterpret_cast<const Rec*>(smemRaw) + j);
                    double2 u = p[0], v = p[1];
                    a = u.x; b = u.y; c = v.x; d = v.y;
                }
                else
                {
                    const double* s = reinterpret_cast<const double*>(smemRaw);
                    a = s[j]; b = s[numRecs + j]; c = s[2 * numRecs + j]; d = s[3 * numRecs + j];
                }
                acc0 += a; acc1 += b; acc2 += c; acc3 += d;
            }
        }
    double s = acc0 + acc1 + acc2 + acc3;
    if (s == -1.2345) out[0] = s;
}

static uint32_t morton3(uint32_t x, uint32_t y, uint32_t z, int bits)
{
    uint32_t k = 0;
    for (int b = bits - 1; b >= 0; --b)
        k = (k << 3) | (((x >> b) & 1) << 2) | (((y >> b) & 1) << 1) | ((z >> b) & 1);
    return k;
}

int main()
{
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int clockKHz = 0;
    CK(cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, 0));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    printf("{\n \"device\": \"%s\", \"sms\": %d, \"sm_clock_max_mhz\": %.1f,\n", prop.name, sms, clockKHz / 1000.0);

    // ---- 1. FMA peaks
    {
        double* out;
        CK(cudaMalloc(&out, 64));
        const int iters = 1 << 14, grid = sms * 16;
        float best64 = 1e30f, best32 = 1e30f;
        for (int rep = 0; rep < 6; ++rep)
        {
            CK(cudaEventRecord(e0));
            fmaKernel<double><<<grid, 256>>>(out, iters, 1.0000001, 1e-9);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep) best64 = std::min(best64, ms);
            CK(cudaEventRecord(e0));
            fmaKernel<float><<<grid, 256>>>(reinterpret_cast<float*>(out), iters, 1.0000001f, 1e-9f);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(&ms, e0, e1));
            if (rep) best32 = std::min(best32, ms);
        }
        double flop = 2.0 * 8 * iters * 256.0 * grid;
        printf(" \"dfma_tflops\": %.3f, \"ffma_tflops\": %.3f,\n", flop / best64 * 1e-9, flop / best32 * 1e-9);
        double perSmClk = flop / 2 / (best64 * 1e-3) / sms / (clockKHz * 1e3);
        printf(" \"dfma_per_sm_per_clk_at_max_clock\": %.2f,\n", perSmClk);
        CK(cudaFree(out));
    }

    // ---- 2. gather patterns
    const int side = 16, numRecs = side * side * side; // 4096 records = 128 KB
    const int listLen = 92;
    const int numSlabs = numRecs / 32;
    std::vector<uint32_t> order(numRecs), rankOf(numRecs);
    {
        std::vector<std::pair<uint32_t, uint32_t>> keyed(numRecs);
        for (int x = 0; x < side; ++x)
            for (int y = 0; y < side; ++y)
                for (int z = 0; z < side; ++z)
                {
                    uint32_t cell = (x * side + y) * side + z;
                    keyed[cell]   = {morton3(x, y, z, 4), cell};
                }
        std::sort(keyed.begin(), keyed.end());
        for (int r = 0; r < numRecs; ++r)
        {
            order[r]                = keyed[r].second;
            rankOf[keyed[r].second] = r;
        }
    }
    struct Pattern
    {
        const char* name;
        std::vector<uint32_t> idx;
    };
    std::vector<Pattern> patterns;
    auto slabIdx = [&](auto f)
    {
        std::vector<uint32_t> v(size_t(numSlabs) * listLen * 32);
        for (int s = 0; s < numSlabs; ++s)
            for (int k = 0; k < listLen; ++k)
                for (int l = 0; l < 32; ++l)
                    v[(size_t(s) * listLen + k) * 32 + l] = f(s, k, l);
        return v;
    };
    patterns.push_back({"coalesced", slabIdx([&](int s, int k, int l) { return uint32_t((s * 32 + k * 32 + l) % numRecs); })});
    patterns.push_back({"broadcast", slabIdx([&](int s, int k, int l) { return uint32_t((s * 32 + k) % numRecs); })});
    for (int W : {32, 128, 512, 4096})
    {
        static char names[4][32];
        static int ni = 0;
        snprintf(names[ni], 32, "random_window_%d", W);
        uint64_t state = 88172645463325252ull;
        patterns.push_back({names[ni++], slabIdx([&](int s, int k, int l) {
                                state ^= state << 13; state ^= state >> 7; state ^= state << 17;
                                return uint32_t((s * 32 + state % W) % numRecs);
                            })});
    }
    {
        // ascending neighbor lists of Morton-consecutive targets on the periodic lattice, r^2 < 8.3 spacings^2
        std::vector<std::vector<uint32_t>> lists(numRecs);
        for (int r = 0; r < numRecs; ++r)
        {
            int cell = order[r], x = cell / (side * side), y = (cell / side) % side, z = cell % side;
            for (int dx = -3; dx <= 3; ++dx)
                for (int dy = -3; dy <= 3; ++dy)
                    for (int dz = -3; dz <= 3; ++dz)
                    {
                        int d2 = dx * dx + dy * dy + dz * dz;
                        if (d2 == 0 || d2 > 8) continue;
                        int xx = (x + dx + side) % side, yy = (y + dy + side) % side, zz = (z + dz + side) % side;
                        lists[r].push_back(rankOf[(xx * side + yy) * side + zz]);
                    }
            std::sort(lists[r].begin(), lists[r].end());
            if (int(lists[r].size()) != listLen)
            {
                fprintf(stderr, "lattice list length %zu != %d\n", lists[r].size(), listLen);
                return 1;
            }
        }
        patterns.push_back({"lattice_neighbor_lists", slabIdx([&](int s, int k, int l) { return lists[s * 32 + l][k]; })});
    }

    std::vector<Rec> recs(numRecs);
    std::vector<double> soa(4 * numRecs);
    for (int i = 0; i < numRecs; ++i)
    {
        recs[i] = {double(i), 1.0, 2.0, 3.0};
        soa[i] = i; soa[numRecs + i] = 1; soa[2 * numRecs + i] = 2; soa[3 * numRecs + i] = 3;
    }
    Rec* dRecs; double* dSoa; uint32_t* dIdx; double* dOut;
    CK(cudaMalloc(&dRecs, numRecs * sizeof(Rec)));
    CK(cudaMalloc(&dSoa, 4 * numRecs * sizeof(double)));
    CK(cudaMalloc(&dIdx, patterns[0].idx.size() * sizeof(uint32_t)));
    CK(cudaMalloc(&dOut, 64));
    CK(cudaMemcpy(dRecs, recs.data(), numRecs * sizeof(Rec), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dSoa, soa.data(), 4 * numRecs * sizeof(double), cudaMemcpyHostToDevice));

    const int smemBytes = numRecs * sizeof(Rec);
    CK(cudaFuncSetAttribute(gatherKernel<LDS128x2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));
    CK(cudaFuncSetAttribute(gatherKernel<LDS_SOA64x4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smemBytes));

    printf(" \"gather_32B_record\": {\n  \"unit\": \"SM cycles per warp-level request (32 lanes x 32 bytes) at saturation, 32 warps per SM, max clock\",\n");
    for (size_t p = 0; p < patterns.size(); ++p)
    {
        CK(cudaMemcpy(dIdx, patterns[p].idx.data(), patterns[p].idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        printf("  \"%s\": {", patterns[p].name);
        for (int m = 0; m < kNumModes; ++m)
        {
            const int grid = sms, block = 1024, repeats = 8;
            auto launch = [&]()
            {
                switch (m)
                {
                    case LDG256: gatherKernel<LDG256><<<grid, block, 0>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                    case LDG128x2: gatherKernel<LDG128x2><<<grid, block, 0>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                    case LDG64x4: gatherKernel<LDG64x4><<<grid, block, 0>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                    case SOA64x4: gatherKernel<SOA64x4><<<grid, block, 0>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                    case LDS128x2: gatherKernel<LDS128x2><<<grid, block, smemBytes>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                    default: gatherKernel<LDS_SOA64x4><<<grid, block, smemBytes>>>(dRecs, dSoa, numRecs, dIdx, numSlabs, listLen, repeats, dOut); break;
                }
            };
            float best = 1e30f;
            for (int rep = 0; rep < 4; ++rep)
            {
                CK(cudaEventRecord(e0));
                launch();
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                CK(cudaGetLastError());
                float ms;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep) best = std::min(best, ms);
            }
            const int numWarps = grid * block / 32;
            double slabsPerWarp = double((numSlabs + numWarps - 1) / numWarps + 1);
            double requestsPerSm = double(block / 32) * repeats * slabsPerWarp * listLen;
            double cycles = best * 1e-3 * clockKHz * 1e3 / requestsPerSm;
            printf("%s\"%s\": %.2f", m ? ", " : "", modeName[m], cycles);
        }
        printf("}%s\n", p + 1 < patterns.size() ? "," : "");
    }
    printf(" }\n}\n");
    return 0;
}
