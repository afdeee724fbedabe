// sphb200_sim: cstone::Domain<uint64_t,double,GpuTag> + sphexa::HydroVeProp state of one rank (shared by api.cu, which
// sequences one rank, and domain.cu, which sequences the ranks of an SFC-partitioned run).
#pragma once
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>
#include "common.cuh"

struct sphb200_comm;

namespace
{

enum FieldId
{
    // conserved (reordered by sync)
    F_x, F_y, F_z, F_h, F_m, F_vx, F_vy, F_vz, F_x_m1, F_y_m1, F_z_m1, F_temp, F_du_m1, F_alpha,
    kNumConserved,
    // dependent
    F_xm = kNumConserved, F_kx, F_gradh, F_prho, F_c, F_c11, F_c12, F_c13, F_c22, F_c23, F_c33, F_divv, F_curlv, F_ax,
    F_ay, F_az, F_du, F_nc, F_dV11, F_dV12, F_dV13, F_dV22, F_dV23, F_dV33,
    kNumFields
};

// element kind: 0 = always double, 1 = real (dtype dependent), 2 = uint32
constexpr int kKind[kNumFields] = {0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 1, 0, 1, 1,            // conserved
                                   1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 2, // xm .. nc
                                   1, 1, 1, 1, 1, 1};

void** fieldSlot(sphb200_fields& f, int id)
{
    switch (id)
    {
        case F_x: return reinterpret_cast<void**>(&f.x);
        case F_y: return reinterpret_cast<void**>(&f.y);
        case F_z: return reinterpret_cast<void**>(&f.z);
        case F_h: return &f.h;
        case F_m: return &f.m;
        case F_vx: return &f.vx;
        case F_vy: return &f.vy;
        case F_vz: return &f.vz;
        case F_x_m1: return &f.x_m1;
        case F_y_m1: return &f.y_m1;
        case F_z_m1: return &f.z_m1;
        case F_temp: return reinterpret_cast<void**>(&f.temp);
        case F_du_m1: return &f.du_m1;
        case F_alpha: return &f.alpha;
        case F_xm: return &f.xm;
        case F_kx: return &f.kx;
        case F_gradh: return &f.gradh;
        case F_prho: return &f.prho;
        case F_c: return &f.c;
        case F_c11: return &f.c11;
        case F_c12: return &f.c12;
        case F_c13: return &f.c13;
        case F_c22: return &f.c22;
        case F_c23: return &f.c23;
        case F_c33: return &f.c33;
        case F_divv: return &f.divv;
        case F_curlv: return &f.curlv;
        case F_ax: return &f.ax;
        case F_ay: return &f.ay;
        case F_az: return &f.az;
        case F_du: return reinterpret_cast<void**>(&f.du);
        case F_nc: return reinterpret_cast<void**>(&f.nc);
        case F_dV11: return &f.dV11;
        case F_dV12: return &f.dV12;
        case F_dV13: return &f.dV13;
        case F_dV22: return &f.dV22;
        case F_dV23: return &f.dV23;
        case F_dV33: return &f.dV33;
    }
    return nullptr;
}

template<class T>
T* devAlloc(size_t n)
{
    void* p = nullptr;
    SPHB200_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    return static_cast<T*>(p);
}

struct Tree
{
    int capacity  = 0; // leaves
    int numLeaves = 0;
    bool first    = true;
    uint64_t* leaves = nullptr;
    uint32_t* counts = nullptr;
};

enum Phase { P_sync, P_neighbors, P_xmass, P_gradh, P_eos, P_iad, P_av, P_momentum, P_update, kNumPhases };

} // namespace

/*! @brief multi-rank state of a sim (domain.cu): SFC assignment, halo exchange pattern, the bucket-64 tree over
 *         owned + halo particles that the neighbor search walks.  cstone keeps the same things in
 *         GlobalAssignment (domain/assignment.hpp), Halos (halos/halos.hpp) and FocusedOctree.
 */
struct DistState
{
    // halo exchange timers: event pairs around the exchange inside the sync and the four of compute_forces
    cudaEvent_t haloEv[10] = {};
    bool haloEvValid[5]    = {};
    sphb200_comm* comm = nullptr;
    int rank = 0, P = 1;
    // SfcAssignment (domaindecomp.hpp:74-113): rank r owns keys [bounds[r], bounds[r+1])
    std::vector<uint64_t> bounds, oldBounds;
    std::vector<uint32_t> rankCounts;
    std::vector<uint64_t> gLeavesH;
    std::vector<uint32_t> gCountsH;
    bool globalFirst = true;
    float haloFactor = 1.0f; // Domain::setHaloFactor (haloSearchExt_)
    Tree search;             // bucketSizeFocus tree over all local keys [0, n)
    // linked octree of the focus tree (own particles): walked by the halo marking
    uint64_t* fPrefixes = nullptr;
    int32_t *fChild = nullptr, *fParents = nullptr, *fLevelRange = nullptr, *fI2L = nullptr, *fL2I = nullptr;
    uint32_t* fLayout  = nullptr;
    uint64_t* keysTmp  = nullptr; // capacity
    uint32_t* leafMask = nullptr; // focus.capacity
    float* radii       = nullptr; // focus.capacity
    sphb200::Arena boxArena, allBoxArena, listArena, sendArena, recvArena, scanArena;
    uint32_t* pinned   = nullptr; // host, kPinnedInts
    uint32_t* dSmall   = nullptr; // device, kPinnedInts
    // halo exchange pattern (SendList of halos.hpp / exchange_halos_gpu.cuh)
    int numSendRanges  = 0;
    uint32_t *dSendScan = nullptr, *dSendBegin = nullptr; // inside listArena
    uint32_t *dRecvScan = nullptr, *dRecvBegin = nullptr; // P entries each, inside dSmall
    std::vector<uint32_t> sendCount, recvCount, haloOff;
    uint32_t numOwn = 0;
    // bookkeeping for tests and the bench
    uint64_t haloBytes = 0;
    uint32_t numSurfaceBoxes = 0, numMigratedOut = 0, numMigratedIn = 0;
    int numFocusLeavesOwned = 0;
    // the reference-style focus tree (focus.cu; MAC-coarsened outside the own key range), built on request by
    // sphb200_domain_update_focus
    sphb200_focus* macTree = nullptr;
    float macTheta         = 0;
    bool macFirst          = true;
    std::vector<int> macPeers;
    sphb200::Arena macSend, macRecv, macCounts;
    static constexpr int kPinnedInts = 8192;
};

void destroyDist(sphb200_sim* s); // domain.cu

struct sphb200_sim
{
    sphb200_ctx* ctx;
    sphb200_sim_config cfg;
    size_t capacity;
    uint32_t n = 0, start = 0, end = 0;

    void* field[kNumFields]       = {};
    void* fieldAlt[kNumConserved] = {};
    uint64_t* keys                = nullptr;
    uint32_t* perm                = nullptr;
    uint32_t* neighbors           = nullptr;
    void *wh = nullptr, *whd = nullptr;

    Tree global, focus;
    // internal octree of the focus tree
    uint64_t* prefixes = nullptr;
    int32_t *childOffsets = nullptr, *parents = nullptr, *levelRange = nullptr, *internalToLeaf = nullptr,
            *leafToInternal = nullptr;
    double *centers = nullptr, *sizes = nullptr;
    uint32_t* layout = nullptr;

    sphb200_box box;
    double minDt, minDt_m1, ttot = 0;
    double dtLocal[3] = {0, 0, 0};
    unsigned numFails = 0;

    cudaEvent_t ev[kNumPhases + 1];
    bool evValid[kNumPhases + 1] = {};

    // sphb200_sim_step_host: copy engines run beside the compute stream
    sphb200_turb* turb = nullptr; // TurbVeProp: stirring after the hydro forces


    // ---- HydroVeBdtProp state (block time steps), see the bdt section at the end of this file
    struct Timestep // sph::Timestep (sph/include/sph/timestep.h:42-66)
    {
        float nextDt = 0, elapsedDt = 0, totDt = 0;
        int numRungs = 1, substep = 0;
        uint32_t rungRanges[5] = {0, 0, 0, 0, 0};
        float dt_m1[4]         = {0, 0, 0, 0};
        float dt_drift[4]      = {0, 0, 0, 0};
    };
    struct Bdt
    {
        bool enabled    = false;
        int safetySteps = 0;
        float etaAccEps = 0;
        Timestep ts, prev;
        uint32_t capGroups = 0;
        uint32_t* buf[2]   = {nullptr, nullptr}; // group delimiters / [start | end] pairs, 2 * capGroups + 2 each
        sphb200_group_view groups{}, tsGroups{}, activeRungs{}, rungs[4]{};
        float* groupDt         = nullptr;
        uint32_t* groupIndices = nullptr;
        uint8_t *rung = nullptr, *rungAlt = nullptr, *active = nullptr, *gather = nullptr;
        void* dtParticle     = nullptr;
        float searchExt      = 1.0f;
        bool allActive       = true;
    } bdt;

    DistState* dist = nullptr; // multi-rank state (domain.cu), owned by the sim

    cudaStream_t copyIn = nullptr, copyOut = nullptr;
    cudaEvent_t evUp[4] = {}, evDown[3] = {}, evRange[16] = {};

    int elemBytes(int id) const
    {
        int k = kKind[id];
        if (k == 0) return 8;
        if (k == 2) return 4;
        return cfg.dtype == SPHB200_F64 ? 8 : 4;
    }

    sphb200_fields table() const
    {
        sphb200_fields f;
        std::memset(&f, 0, sizeof(f));
        for (int id = 0; id < kNumFields; ++id)
            *fieldSlot(f, id) = field[id];
        f.n = n;
        return f;
    }

    sphb200_hydro_const hydroConst() const
    {
        sphb200_hydro_const hc = cfg.hc;
        hc.wh                  = wh;
        hc.whd                 = whd;
        hc.minDt               = minDt;
        hc.minDt_m1            = minDt_m1;
        return hc;
    }

    sphb200_tree_view treeView() const
    {
        // one rank: the focus tree covers everything; P ranks: the tree over owned + halo particles
        const Tree& focus = dist ? dist->search : this->focus;
        sphb200_tree_view t;
        t.numLeafNodes   = focus.numLeaves;
        t.numNodes       = focus.numLeaves + (focus.numLeaves - 1) / 7;
        t.prefixes       = prefixes;
        t.childOffsets   = childOffsets;
        t.internalToLeaf = internalToLeaf;
        t.levelRange     = levelRange;
        t.leaves         = focus.leaves;
        t.layout         = layout;
        t.centers        = centers;
        t.sizes          = sizes;
        return t;
    }

    void mark(int phase)
    {
        SPHB200_CUDA(cudaEventRecord(ev[phase], ctx->stream));
        evValid[phase] = true;
    }
};


#ifndef SPHB200_TRY
#define SPHB200_TRY try {
#define SPHB200_CATCH                                                                                                 \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        sphb200::setError(e.what());                                                                                  \
        return 1;                                                                                                     \
    }
#endif
#ifndef CHECK_RC
#define CHECK_RC(call)                                                                                                \
    do                                                                                                                \
    {                                                                                                                 \
        if ((call) != 0) throw std::runtime_error(sphb200_last_error());                                              \
    } while (0)
#endif
