// Internal (non-ABI) declarations of the primitives shared between translation units.
#pragma once
#include "common.cuh"

namespace sphb200
{
//! @brief out[i] = sum(in[0..i)), in == out allowed; optional device total
void exclusiveScan(sphb200_ctx* ctx, const uint32_t* in, uint32_t* out, size_t n, uint32_t* totalOut);
//! @brief stable LSD radix sort of (key, payload) on bits [beginBit, endBit); payload = iota if valsAreIota
void radixSortPairs(sphb200_ctx* ctx, KeyType* keys, uint32_t* vals, size_t n, bool valsAreIota, int beginBit,
                    int endBit);
} // namespace sphb200
