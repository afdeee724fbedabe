// Halo discovery helpers for SFC-partitioned multi-GPU runs.
//
// segmentMax: per-leaf maximum smoothing length (segmentMaxKernel, primitives_gpu.cu:180-207).
// markHaloLeaves: owner-side halo discovery.  The reference marks, per OWN leaf, the FOREIGN leaves whose integer box
// overlaps the own leaf's halo box (findHalosKernel, traversal/collisions_gpu.cu:39-66) and then asks the owners for
// them.  Here every rank publishes a few hundred axis-aligned "query boxes" (its leaves' halo boxes, unioned per
// coarse SFC block) and each owner marks its own bucket-64 leaves that overlap a peer's query box; what is sent is a
// superset of the reference's halo set, and the neighbor sets of all owned particles are identical.
#include "common.cuh"

namespace sphb200
{

template<class T>
__global__ void segmentMaxKernel(const T* __restrict__ h, const uint32_t* __restrict__ layout, int numSegments,
                                 double* __restrict__ out)
{
    // one warp per segment (leaves hold <= ~64 particles)
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= numSegments) return;
    uint32_t a = layout[warp], b = layout[warp + 1];
    double m = 0.0;
    for (uint32_t i = a + lane; i < b; i += 32)
        m = fmax(m, double(h[i]));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0) out[warp] = m;
}

/*! @brief mask[leaf] bit q set <=> the leaf's box overlaps one of peer q's query boxes (periodic-aware)
 *
 * boxes: [numPeers][maxBoxes][6] = center xyz, half-size xyz; boxCount[numPeers]
 */
__global__ void markHaloLeavesKernel(const double* __restrict__ leafCenters, const double* __restrict__ leafSizes,
                                     int numLeaves, const double* __restrict__ boxes,
                                     const int* __restrict__ boxCount, int numPeers, int maxBoxes, int myRank,
                                     sphb200_box box, uint32_t* __restrict__ mask)
{
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= numLeaves) return;
    double c[3] = {leafCenters[3 * l], leafCenters[3 * l + 1], leafCenters[3 * l + 2]};
    double s[3] = {leafSizes[3 * l], leafSizes[3 * l + 1], leafSizes[3 * l + 2]};
    uint32_t m  = 0;
    for (int q = 0; q < numPeers; ++q)
    {
        if (q == myRank) continue;
        const double* bq = boxes + size_t(q) * maxBoxes * 6;
        int nb           = boxCount[q];
        bool hit         = false;
        for (int b = 0; b < nb && !hit; ++b)
        {
            bool ok = true;
#pragma unroll
            for (int d = 0; d < 3; ++d)
            {
                double dx = bq[6 * b + d] - c[d];
                if (box.bc[d] == SPHB200_PERIODIC) dx -= box.len[d] * rint(dx * box.inv[d]);
                // conservative margin: pruning only
                ok = ok && (fabs(dx) - s[d] - bq[6 * b + 3 + d] <= 1e-12 * box.len[d]);
            }
            hit = ok;
        }
        if (hit) m |= (1u << q);
    }
    mask[l] = m;
}

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_segment_max(sphb200_ctx* ctx, const void* h, int dtype, const uint32_t* layout, int numSegments,
                                   double* out)
{
    try
    {
        if (numSegments == 0) return 0;
        int grid = gridFor(size_t(numSegments) * 32, 256);
        if (dtype == SPHB200_F64)
            SPHB200_LAUNCH(ctx, segmentMaxKernel<double>, grid, 256, 0, static_cast<const double*>(h), layout,
                           numSegments, out);
        else
            SPHB200_LAUNCH(ctx, segmentMaxKernel<float>, grid, 256, 0, static_cast<const float*>(h), layout,
                           numSegments, out);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_mark_halo_leaves(sphb200_ctx* ctx, const double* leafCenters, const double* leafSizes,
                                        int numLeaves, const double* boxes, const int* boxCount, int numPeers,
                                        int maxBoxes, int myRank, const sphb200_box* box, uint32_t* mask)
{
    try
    {
        if (numPeers > 32) throw std::runtime_error("mark_halo_leaves: at most 32 peers");
        if (numLeaves == 0) return 0;
        SPHB200_LAUNCH(ctx, markHaloLeavesKernel, gridFor(numLeaves, 128), 128, 0, leafCenters, leafSizes, numLeaves,
                       boxes, boxCount, numPeers, maxBoxes, myRank, *box, mask);
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}
