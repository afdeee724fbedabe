// Transport of the multi-rank domain (domain.cu): the collectives cstone issues through MPI
// (MPI_Allreduce in tree/update_mpi.hpp:95, box_mpi.hpp:104; point-to-point in domaindecomp_mpi_gpu.cuh:85-186 and
// halos/exchange_halos_gpu.cuh:50-143), here either NCCL called from C++ on the compute stream (one process per GPU)
// or a loopback that runs P virtual ranks inside one process on one GPU (exchanges = device-to-device copies), so
// that the multi-rank code paths can be tested on a single GPU.
//
// All operations are "lockstep": the caller passes the per-rank arguments of every rank this process drives
// (one for NCCL, all P for the loopback) and the call is complete, in stream order, for all of them on return.
#pragma once
#include <vector>
#include "common.cuh"

enum CommType { COMM_U32, COMM_U64, COMM_F32, COMM_F64 };
enum CommOp { COMM_SUM, COMM_MIN, COMM_MAX };

struct sphb200_comm
{
    int numRanks   = 1;
    bool loopback  = true;
    int rank       = 0;       // NCCL: the rank of this process
    void* nccl     = nullptr; // ncclComm_t
    int device     = 0;
};

namespace sphb200
{

struct RankStream
{
    int rank;
    cudaStream_t stream;
    sphb200_ctx* ctx;
};

//! @brief in-place all-reduce of count elements at ptr[l] of every local rank
void commAllReduce(sphb200_comm* c, const std::vector<RankStream>& local, void* const* ptr, size_t count, CommType type,
                   CommOp op);
//! @brief recv[l] = concatenation over ranks q of the bytes at send[q] (bytes each)
void commAllGather(sphb200_comm* c, const std::vector<RankStream>& local, const void* const* send, void* const* recv,
                   size_t bytes);
/*! @brief general exchange: local rank l sends sendBytes[l][q] bytes at sendBase[l] + sendOff[l][q] to rank q, which
 *         stores them at recvBase[q] + recvOff[q][rank(l)]; recvBytes[l][q] = what local rank l expects from q.
 *         The self block (q == rank) is skipped.  Offsets/sizes: numLocal x numRanks, row-major.
 */
void commExchange(sphb200_comm* c, const std::vector<RankStream>& local, const void* const* sendBase,
                  const size_t* sendOff, const size_t* sendBytes, void* const* recvBase, const size_t* recvOff,
                  const size_t* recvBytes);

/*! @brief the same exchange for several (send, recv) array pairs in ONE group: array a of local rank l sends
 *         count[l][q] * elemBytes[a] bytes from sendBase[a][l] + sendFirst[l][q] * elemBytes[a] and receives at
 *         recvBase[a][l] + recvFirst[l][q] * elemBytes[a]  (particle migration: contiguous index ranges per field)
 */
void commExchangeArrays(sphb200_comm* c, const std::vector<RankStream>& local, int numArrays, const int* elemBytes,
                        const void* const* sendBase /*[a][l]*/, const size_t* sendFirst, const size_t* sendCount,
                        void* const* recvBase /*[a][l]*/, const size_t* recvFirst, const size_t* recvCount);

} // namespace sphb200
