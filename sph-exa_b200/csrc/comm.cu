// Transport of the multi-rank domain: NCCL (loaded at run time, called from C++ on the compute stream) or a
// single-process loopback over P virtual ranks.  See comm.cuh.
#include <dlfcn.h>
#include <mutex>
#include <nccl.h>
#include "comm.cuh"

namespace sphb200
{

// ---------------------------------------------------------------------------------------------- NCCL, loaded lazily
// The library is not linked: a process that never creates an NCCL communicator (single rank, loopback tests, the
// CPU-side symbol checks) must not need libnccl.  dlopen finds the copy torch already mapped when there is one.
struct NcclApi
{
    ncclResult_t (*GetUniqueId)(ncclUniqueId*)                                                                = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int)                                         = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t)                                                                   = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t)            = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)                   = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t)                         = nullptr;
    ncclResult_t (*GroupStart)()                                                                               = nullptr;
    ncclResult_t (*GroupEnd)()                                                                                 = nullptr;
    const char* (*GetErrorString)(ncclResult_t)                                                                = nullptr;
};

static NcclApi& nccl()
{
    static NcclApi api;
    static std::once_flag once;
    static std::string failure;
    std::call_once(once,
                   []()
                   {
                       void* h = nullptr;
                       for (const char* name : {"libnccl.so.2", "libnccl.so"})
                       {
                           h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
                           if (h) break;
                       }
                       if (!h)
                       {
                           failure = std::string("cannot load libnccl.so.2: ") + dlerror();
                           return;
                       }
                       auto sym = [&](const char* n)
                       {
                           void* p = dlsym(h, n);
                           if (!p) failure = std::string("libnccl lacks ") + n;
                           return p;
                       };
                       api.GetUniqueId    = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
                       api.CommInitRank   = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
                       api.CommDestroy    = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
                       api.AllReduce      = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
                       api.AllGather      = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
                       api.Send           = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
                       api.Recv           = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
                       api.GroupStart     = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
                       api.GroupEnd       = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
                       api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
                   });
    if (!failure.empty()) throw std::runtime_error(failure);
    return api;
}

#define SPHB200_NCCL(call)                                                                                            \
    do                                                                                                                \
    {                                                                                                                 \
        ncclResult_t r__ = (call);                                                                                    \
        if (r__ != ncclSuccess)                                                                                       \
            throw std::runtime_error(std::string(#call) + " failed: " + nccl().GetErrorString(r__));                  \
    } while (0)

// ---------------------------------------------------------------------------------------------- loopback kernels
constexpr int kMaxLoopRanks = 32;
struct PtrPack
{
    void* p[kMaxLoopRanks];
    int n;
};

template<class T, int Op>
__global__ void loopReduceKernel(PtrPack pk, size_t count)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= count) return;
    T v = static_cast<const T*>(pk.p[0])[i];
    for (int r = 1; r < pk.n; ++r)
    {
        T w = static_cast<const T*>(pk.p[r])[i];
        if (Op == COMM_SUM) v = v + w;
        else if (Op == COMM_MIN) v = w < v ? w : v;
        else v = w > v ? w : v;
    }
    for (int r = 0; r < pk.n; ++r)
        static_cast<T*>(pk.p[r])[i] = v;
}

template<class T>
static void loopReduce(const RankStream& rs, const PtrPack& pk, size_t count, CommOp op)
{
    int grid = gridFor(count, 256);
    if (op == COMM_SUM) SPHB200_LAUNCH(rs.ctx, (loopReduceKernel<T, COMM_SUM>), grid, 256, 0, pk, count);
    else if (op == COMM_MIN) SPHB200_LAUNCH(rs.ctx, (loopReduceKernel<T, COMM_MIN>), grid, 256, 0, pk, count);
    else SPHB200_LAUNCH(rs.ctx, (loopReduceKernel<T, COMM_MAX>), grid, 256, 0, pk, count);
}

static void checkLoop(sphb200_comm* c, const std::vector<RankStream>& local)
{
    if (int(local.size()) != c->numRanks)
        throw std::runtime_error("loopback transport: every collective needs all " + std::to_string(c->numRanks) +
                                 " virtual ranks");
    for (int l = 0; l < c->numRanks; ++l)
    {
        if (local[l].rank != l) throw std::runtime_error("loopback transport: pass the ranks in order 0..P-1");
        if (local[l].stream != local[0].stream)
            throw std::runtime_error("loopback transport: all virtual ranks must share one CUDA stream");
    }
}

static ncclDataType_t ncclType(CommType t)
{
    switch (t)
    {
        case COMM_U32: return ncclUint32;
        case COMM_U64: return ncclUint64;
        case COMM_F32: return ncclFloat32;
        default: return ncclFloat64;
    }
}

void commAllReduce(sphb200_comm* c, const std::vector<RankStream>& local, void* const* ptr, size_t count, CommType type,
                   CommOp op)
{
    if (c->numRanks == 1 || count == 0) return;
    if (c->loopback)
    {
        checkLoop(c, local);
        PtrPack pk;
        pk.n = c->numRanks;
        for (int l = 0; l < pk.n; ++l)
            pk.p[l] = ptr[l];
        switch (type)
        {
            case COMM_U32: loopReduce<uint32_t>(local[0], pk, count, op); break;
            case COMM_U64: loopReduce<unsigned long long>(local[0], pk, count, op); break;
            case COMM_F32: loopReduce<float>(local[0], pk, count, op); break;
            default: loopReduce<double>(local[0], pk, count, op); break;
        }
        return;
    }
    ncclRedOp_t rop = op == COMM_SUM ? ncclSum : (op == COMM_MIN ? ncclMin : ncclMax);
    SPHB200_NCCL(nccl().AllReduce(ptr[0], ptr[0], count, ncclType(type), rop, static_cast<ncclComm_t>(c->nccl),
                                  local[0].stream));
}

void commAllGather(sphb200_comm* c, const std::vector<RankStream>& local, const void* const* send, void* const* recv,
                   size_t bytes)
{
    if (bytes == 0) return;
    if (c->loopback)
    {
        checkLoop(c, local);
        for (int l = 0; l < c->numRanks; ++l)
            for (int q = 0; q < c->numRanks; ++q)
                SPHB200_CUDA(cudaMemcpyAsync(static_cast<char*>(recv[l]) + size_t(q) * bytes, send[q], bytes,
                                             cudaMemcpyDeviceToDevice, local[0].stream));
        return;
    }
    SPHB200_NCCL(nccl().AllGather(send[0], recv[0], bytes, ncclChar, static_cast<ncclComm_t>(c->nccl), local[0].stream));
}

void commExchange(sphb200_comm* c, const std::vector<RankStream>& local, const void* const* sendBase,
                  const size_t* sendOff, const size_t* sendBytes, void* const* recvBase, const size_t* recvOff,
                  const size_t* recvBytes)
{
    const int P = c->numRanks;
    if (c->loopback)
    {
        checkLoop(c, local);
        for (int a = 0; a < P; ++a)
            for (int q = 0; q < P; ++q)
            {
                if (q == a) continue;
                size_t bytes = sendBytes[a * P + q];
                if (bytes != recvBytes[q * P + a])
                    throw std::runtime_error("loopback exchange: rank " + std::to_string(a) + " sends " +
                                             std::to_string(bytes) + " bytes to " + std::to_string(q) + " which expects " +
                                             std::to_string(recvBytes[q * P + a]));
                if (bytes == 0) continue;
                SPHB200_CUDA(cudaMemcpyAsync(static_cast<char*>(recvBase[q]) + recvOff[q * P + a],
                                             static_cast<const char*>(sendBase[a]) + sendOff[a * P + q], bytes,
                                             cudaMemcpyDeviceToDevice, local[0].stream));
            }
        return;
    }
    auto comm = static_cast<ncclComm_t>(c->nccl);
    SPHB200_NCCL(nccl().GroupStart());
    for (int q = 0; q < P; ++q)
    {
        if (q == c->rank) continue;
        if (sendBytes[q])
            SPHB200_NCCL(nccl().Send(static_cast<const char*>(sendBase[0]) + sendOff[q], sendBytes[q], ncclChar, q, comm,
                                     local[0].stream));
        if (recvBytes[q])
            SPHB200_NCCL(nccl().Recv(static_cast<char*>(recvBase[0]) + recvOff[q], recvBytes[q], ncclChar, q, comm,
                                     local[0].stream));
    }
    SPHB200_NCCL(nccl().GroupEnd());
}

void commExchangeArrays(sphb200_comm* c, const std::vector<RankStream>& local, int numArrays, const int* elemBytes,
                        const void* const* sendBase, const size_t* sendFirst, const size_t* sendCount,
                        void* const* recvBase, const size_t* recvFirst, const size_t* recvCount)
{
    const int P = c->numRanks;
    const int numLocal = int(local.size());
    if (c->loopback)
    {
        checkLoop(c, local);
        for (int a = 0; a < P; ++a)
            for (int q = 0; q < P; ++q)
            {
                if (q == a) continue;
                size_t cnt = sendCount[a * P + q];
                if (cnt != recvCount[q * P + a]) throw std::runtime_error("loopback exchange: count mismatch");
                if (cnt == 0) continue;
                for (int f = 0; f < numArrays; ++f)
                {
                    size_t eb = elemBytes[f];
                    SPHB200_CUDA(cudaMemcpyAsync(
                        static_cast<char*>(recvBase[f * numLocal + q]) + recvFirst[q * P + a] * eb,
                        static_cast<const char*>(sendBase[f * numLocal + a]) + sendFirst[a * P + q] * eb, cnt * eb,
                        cudaMemcpyDeviceToDevice, local[0].stream));
                }
            }
        return;
    }
    auto comm = static_cast<ncclComm_t>(c->nccl);
    SPHB200_NCCL(nccl().GroupStart());
    for (int q = 0; q < P; ++q)
    {
        if (q == c->rank) continue;
        for (int f = 0; f < numArrays; ++f)
        {
            size_t eb = elemBytes[f];
            if (sendCount[q])
                SPHB200_NCCL(nccl().Send(static_cast<const char*>(sendBase[f]) + sendFirst[q] * eb, sendCount[q] * eb,
                                         ncclChar, q, comm, local[0].stream));
            if (recvCount[q])
                SPHB200_NCCL(nccl().Recv(static_cast<char*>(recvBase[f]) + recvFirst[q] * eb, recvCount[q] * eb, ncclChar,
                                         q, comm, local[0].stream));
        }
    }
    SPHB200_NCCL(nccl().GroupEnd());
}

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_comm_create_loopback(int numRanks, sphb200_comm** out)
{
    try
    {
        if (numRanks < 1 || numRanks > kMaxLoopRanks)
            throw std::runtime_error("comm_create_loopback: 1 <= numRanks <= " + std::to_string(kMaxLoopRanks));
        auto* c     = new sphb200_comm;
        c->numRanks = numRanks;
        c->loopback = true;
        *out        = c;
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_comm_unique_id(void* id128)
{
    try
    {
        ncclUniqueId id;
        SPHB200_NCCL(nccl().GetUniqueId(&id));
        static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
        std::memcpy(id128, &id, sizeof(id));
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_comm_create_nccl(int device, int rank, int numRanks, const void* id128, sphb200_comm** out)
{
    try
    {
        if (numRanks < 1 || numRanks > 32) throw std::runtime_error("comm_create_nccl: 1 <= numRanks <= 32");
        if (rank < 0 || rank >= numRanks) throw std::runtime_error("comm_create_nccl: bad rank");
        SPHB200_CUDA(cudaSetDevice(device));
        ncclUniqueId id;
        std::memcpy(&id, id128, sizeof(id));
        ncclComm_t comm = nullptr;
        SPHB200_NCCL(nccl().CommInitRank(&comm, numRanks, id, rank));
        auto* c     = new sphb200_comm;
        c->numRanks = numRanks;
        c->loopback = false;
        c->rank     = rank;
        c->nccl     = comm;
        c->device   = device;
        *out        = c;
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" void sphb200_comm_destroy(sphb200_comm* c)
{
    if (!c) return;
    if (c->nccl)
    {
        try
        {
            nccl().CommDestroy(static_cast<ncclComm_t>(c->nccl));
        }
        catch (...)
        {
        }
    }
    delete c;
}

extern "C" int sphb200_comm_num_ranks(const sphb200_comm* c) { return c->numRanks; }
extern "C" int sphb200_comm_is_loopback(const sphb200_comm* c) { return c->loopback ? 1 : 0; }
