// Hilbert curve device functions shared by the key kernels (sfc.cu) and the halo discovery (domain.cu).
#pragma once
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
__host__ __device__ __forceinline__ void hilbertDecode(KeyType key, unsigned& ox, unsigned& oy, unsigned& oz)
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

} // namespace sphb200
