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
//! @brief one updateOctree step (rebalance + recount) that keeps the tree a refinement of the `guide` leaf array
void updateOctreeGuided(sphb200_ctx* ctx, const KeyType* keys, size_t n, unsigned bucket, KeyType* leaves,
                        uint32_t* counts, int capacity, int* numLeaves, int* converged, const KeyType* guide,
                        int numGuide, int* forced);
//! @brief rebalanceTree (tree/csarray.hpp:390-415): emit the leaves of old leaf i at opScan[i]; opScan = exclusive scan
//!        of the per-leaf ops {0, 1, 8, 64, ...} with numLeaves + 1 entries
void rebalanceTree(sphb200_ctx* ctx, const KeyType* oldLeaves, const uint32_t* opScan, int numLeaves, KeyType* newLeaves);
//! @brief leaves = {0, 2^63}, counts = {n}
void initRootTree(sphb200_ctx* ctx, KeyType* leaves, uint32_t* counts, uint32_t n);
} // namespace sphb200
