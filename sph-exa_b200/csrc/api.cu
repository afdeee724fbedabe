// Context management and the tier-2 (Domain + Propagator) layer of the C ABI.
//
// sphb200_sim plays the role of cstone::Domain<uint64_t,double,GpuTag> (domain/include/cstone/domain/domain.hpp)
// plus sphexa::HydroVeProp (main/src/propagator/ve_hydro.hpp) for one rank: it owns the SFC-sorted SoA
// particle arrays, both cornerstone trees, the neighbor lists, and sequences the operators of tier 1.
#include <cmath>
#include <limits>
#include <algorithm>
#include <array>
#include <mutex>
#include <cstdlib>
#include <utility>
#include <vector>
#include "common.cuh"

namespace sphb200
{

static thread_local std::string g_lastError;
void setError(const std::string& msg) { g_lastError = msg; }

} // namespace sphb200

using namespace sphb200;

#define SPHB200_TRY try {
#define SPHB200_CATCH                                                                                                 \
    return 0;                                                                                                         \
    }                                                                                                                 \
    catch (std::exception & e)                                                                                        \
    {                                                                                                                 \
        setError(e.what());                                                                                           \
        return 1;                                                                                                     \
    }

#define CHECK_RC(call)                                                                                                \
    do                                                                                                                \
    {                                                                                                                 \
        if ((call) != 0) throw std::runtime_error(sphb200_last_error());                                              \
    } while (0)

extern "C" const char* sphb200_last_error(void) { return g_lastError.c_str(); }

extern "C" int sphb200_create(int device, void* stream, sphb200_ctx** out)
{
    SPHB200_TRY
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0)
        throw std::runtime_error("sphb200_create: no CUDA device available (this library has no CPU fallback)");
    SPHB200_CUDA(cudaSetDevice(device));
    auto* ctx   = new sphb200_ctx;
    ctx->device = device;
    if (stream != SPHB200_PRIVATE_STREAM) { ctx->stream = static_cast<cudaStream_t>(stream); }
    else
    {
        SPHB200_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->ownStream = true;
    }
    SPHB200_CUDA(cudaMallocHost(&ctx->hostScalars, 64 * sizeof(int)));
    SPHB200_CUDA(cudaMalloc(&ctx->devScalars, 64 * sizeof(int)));
    SPHB200_CUDA(cudaMalloc(&ctx->devDt, 32 * sizeof(double)));
    SPHB200_CUDA(cudaMallocHost(&ctx->hostDt, 32 * sizeof(double)));
    SPHB200_CUDA(cudaMemset(ctx->devScalars, 0, 64 * sizeof(int)));
    SPHB200_CUDA(cudaMemset(ctx->devDt, 0, 32 * sizeof(double)));
    *out = ctx;
    SPHB200_CATCH
}

extern "C" void sphb200_destroy(sphb200_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    ctx->sortTmp.release();
    ctx->treeTmp.release();
    ctx->scanTmp.release();
    ctx->miscTmp.release();
    ctx->nbTmp.release();
    ctx->obsTmp.release();
    ctx->machTmp.release();
    for (auto& a : ctx->rec)
        a.release();
    ctx->tabPairs.release();
    for (auto& t : ctx->tabTune)
        for (auto& e : t.ev)
            if (e) cudaEventDestroy(e);
    ctx->tabQuads.release();
    cudaFreeHost(ctx->hostScalars);
    cudaFree(ctx->devScalars);
    cudaFree(ctx->devDt);
    cudaFreeHost(ctx->hostDt);
    if (ctx->ownStream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int sphb200_sync(sphb200_ctx* ctx)
{
    SPHB200_TRY
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    SPHB200_CATCH
}

extern "C" uint64_t sphb200_launch_count(const sphb200_ctx* ctx) { return ctx->launches; }

extern "C" void sphb200_make_box(const double lim[6], const int bc[3], sphb200_box* out)
{
    for (int k = 0; k < 6; ++k)
        out->lim[k] = lim[k];
    for (int d = 0; d < 3; ++d)
    {
        out->len[d] = lim[2 * d + 1] - lim[2 * d];
        out->inv[d] = 1.0 / (lim[2 * d + 1] - lim[2 * d]);
        out->bc[d]  = bc[d];
    }
}

// =====================================================================================================
// tier 2
// =====================================================================================================

#include "sim.cuh"

extern "C" int sphb200_sim_create(sphb200_ctx* ctx, const sphb200_sim_config* cfg, sphb200_sim** out)
{
    SPHB200_TRY
    if (cfg->dtype != SPHB200_F64 && cfg->dtype != SPHB200_MIXED) throw std::runtime_error("sim_create: bad dtype");
    if (cfg->hc.ng0 > cfg->hc.ngmax) throw std::runtime_error("ng0 should be smaller than ngmax\n"); // find_neighbors.hpp:52
    auto* s     = new sphb200_sim;
    s->ctx      = ctx;
    // the sim owns the field arrays: gather records (hydro.cu) stay valid between operator calls and are
    // invalidated below wherever a packed field is rewritten
    ctx->packReuse = true;
    ctx->packValid = 0;
    ctx->tabValid  = false;
    ctx->resetTableTuners(); // a new simulation: time the table variants on its data
    s->cfg      = *cfg;
    s->capacity = cfg->capacity;
    s->box      = cfg->box;
    s->minDt    = cfg->hc.minDt;
    s->minDt_m1 = cfg->hc.minDt_m1;
    size_t cap  = s->capacity;
    for (int id = 0; id < kNumFields; ++id)
    {
        if (id >= F_dV11 && !cfg->hc.avClean) continue;
        s->field[id] = devAlloc<char>(cap * s->elemBytes(id));
    }
    for (int id = 0; id < kNumConserved; ++id)
        s->fieldAlt[id] = devAlloc<char>(cap * s->elemBytes(id));
    s->keys      = devAlloc<uint64_t>(cap);
    s->perm      = devAlloc<uint32_t>(cap);
    s->neighbors = devAlloc<uint32_t>(sphb200_neighbors_size(cap, cfg->hc.ngmax));

    size_t tb = size_t(kTableSize) * (cfg->dtype == SPHB200_F64 ? 8 : 4);
    s->wh     = devAlloc<char>(tb);
    s->whd    = devAlloc<char>(tb);
    SPHB200_CUDA(cudaMemcpy(s->wh, cfg->hc.wh, tb, cudaMemcpyHostToDevice));
    SPHB200_CUDA(cudaMemcpy(s->whd, cfg->hc.whd, tb, cudaMemcpyHostToDevice));

    // leaf capacities: a converged bucket-b tree over n particles has < 8 n / b + 8 leaves in practice; be generous
    auto leafCap = [cap](unsigned bucket) { return int(std::min<size_t>(64 + 16 * (cap / std::max(1u, bucket) + 1), size_t(1) << 30)); };
    s->global.capacity = leafCap(cfg->bucketSize) + 4096;
    s->focus.capacity  = leafCap(cfg->bucketSizeFocus) + 4096;
    for (Tree* t : {&s->global, &s->focus})
    {
        t->leaves = devAlloc<uint64_t>(size_t(t->capacity) + 2);
        t->counts = devAlloc<uint32_t>(size_t(t->capacity) + 2);
    }
    size_t maxNodes   = size_t(s->focus.capacity) + s->focus.capacity / 7 + 2;
    s->prefixes       = devAlloc<uint64_t>(maxNodes);
    s->childOffsets   = devAlloc<int32_t>(maxNodes + 1);
    s->parents        = devAlloc<int32_t>(maxNodes / 8 + 2);
    s->levelRange     = devAlloc<int32_t>(32);
    s->internalToLeaf = devAlloc<int32_t>(maxNodes);
    s->leafToInternal = devAlloc<int32_t>(maxNodes);
    s->centers        = devAlloc<double>(3 * maxNodes);
    s->sizes          = devAlloc<double>(3 * maxNodes);
    s->layout         = devAlloc<uint32_t>(size_t(s->focus.capacity) + 2);
    for (auto& e : s->ev)
        SPHB200_CUDA(cudaEventCreate(&e));
    *out = s;
    SPHB200_CATCH
}

extern "C" void sphb200_sim_destroy(sphb200_sim* s)
{
    if (!s) return;
    cudaStreamSynchronize(s->ctx->stream);
    s->ctx->packReuse = false;
    s->ctx->packValid = 0;
    s->ctx->tabValid  = false;
    for (auto p : s->field) cudaFree(p);
    for (auto p : s->fieldAlt) cudaFree(p);
    cudaFree(s->keys); cudaFree(s->perm); cudaFree(s->neighbors); cudaFree(s->wh); cudaFree(s->whd);
    for (Tree* t : {&s->global, &s->focus}) { cudaFree(t->leaves); cudaFree(t->counts); }
    cudaFree(s->prefixes); cudaFree(s->childOffsets); cudaFree(s->parents); cudaFree(s->levelRange);
    cudaFree(s->internalToLeaf); cudaFree(s->leafToInternal); cudaFree(s->centers); cudaFree(s->sizes);
    cudaFree(s->layout);
    for (auto p : s->bdt.buf) cudaFree(p);
    cudaFree(s->bdt.groupDt); cudaFree(s->bdt.groupIndices); cudaFree(s->bdt.rung); cudaFree(s->bdt.rungAlt);
    cudaFree(s->bdt.active); cudaFree(s->bdt.gather); cudaFree(s->bdt.dtParticle);
    if (s->dist) destroyDist(s);
    for (auto& e : s->ev) cudaEventDestroy(e);
    if (s->copyIn)
    {
        cudaStreamDestroy(s->copyIn);
        cudaStreamDestroy(s->copyOut);
        for (auto& e : s->evUp) cudaEventDestroy(e);
        for (auto& e : s->evDown) cudaEventDestroy(e);
        for (auto& e : s->evRange)
            if (e) cudaEventDestroy(e);
    }
    delete s;
}

extern "C" int sphb200_sim_upload(sphb200_sim* s, size_t n, const sphb200_fields* host)
{
    SPHB200_TRY
    if (n > s->capacity) throw std::runtime_error("sim_upload: n exceeds capacity");
    s->ctx->packValid = 0;
    sphb200_fields h = *host;
    for (int id = 0; id < kNumFields; ++id)
    {
        void* src = *fieldSlot(h, id);
        if (!src || !s->field[id]) continue;
        SPHB200_CUDA(cudaMemcpyAsync(s->field[id], src, n * s->elemBytes(id), cudaMemcpyHostToDevice, s->ctx->stream));
    }
    s->n     = uint32_t(n);
    s->start = 0;
    s->end   = uint32_t(n);
    SPHB200_CATCH
}

extern "C" int sphb200_sim_download(sphb200_sim* s, const sphb200_fields* host, uint64_t* keys_h)
{
    SPHB200_TRY
    sphb200_fields h = *host;
    size_t cnt       = s->end - s->start;
    for (int id = 0; id < kNumFields; ++id)
    {
        void* dst = *fieldSlot(h, id);
        if (!dst || !s->field[id]) continue;
        size_t eb = s->elemBytes(id);
        SPHB200_CUDA(cudaMemcpyAsync(dst, static_cast<char*>(s->field[id]) + s->start * eb, cnt * eb,
                                     cudaMemcpyDeviceToHost, s->ctx->stream));
    }
    if (keys_h)
        SPHB200_CUDA(cudaMemcpyAsync(keys_h, s->keys + s->start, cnt * sizeof(uint64_t), cudaMemcpyDeviceToHost, s->ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
    SPHB200_CATCH
}

static void updateTree(sphb200_sim* s, Tree& t, unsigned bucket, const uint64_t* keys, size_t n)
{
    if (t.first)
    {
        // first call: converge (assignment.hpp:108-113, domain.hpp:220-224)
        int it = 0;
        CHECK_RC(sphb200_compute_octree(s->ctx, keys, n, bucket, t.leaves, t.counts, t.capacity, &t.numLeaves, &it));
        t.first = false;
    }
    else
    {
        int conv = 0;
        CHECK_RC(sphb200_update_octree(s->ctx, keys, n, bucket, t.leaves, t.counts, t.capacity, &t.numLeaves, &conv));
    }
}

// ---- Domain::sync split into the parts that sphb200_sim_step_host overlaps with the upload of later field groups
//! @brief global box, SFC keys, sort permutation, global tree (needs x, y, z only)
static void syncKeys(sphb200_sim* s)
{
    auto* ctx = s->ctx;
    size_t numAssigned = s->end - s->start;
    auto* x = static_cast<double*>(s->field[F_x]) + s->start;
    auto* y = static_cast<double*>(s->field[F_y]) + s->start;
    auto* z = static_cast<double*>(s->field[F_z]) + s->start;

    // makeGlobalBox (box_mpi.hpp:84-119): open axes follow the particle extrema
    bool anyOpen = false;
    for (int d = 0; d < 3; ++d)
        anyOpen = anyOpen || s->box.bc[d] != SPHB200_PERIODIC;
    if (anyOpen)
    {
        double mm[6];
        CHECK_RC(sphb200_minmax3(ctx, x, y, z, numAssigned, mm));
        double lim[6];
        for (int d = 0; d < 3; ++d)
        {
            bool pbc       = s->box.bc[d] == SPHB200_PERIODIC;
            lim[2 * d]     = pbc ? s->box.lim[2 * d] : mm[2 * d];
            lim[2 * d + 1] = pbc ? s->box.lim[2 * d + 1] : mm[2 * d + 1];
        }
        sphb200_make_box(lim, s->box.bc, &s->box);
    }

    uint64_t* keys = s->keys + s->start;
    CHECK_RC(sphb200_sfc_keys(ctx, x, y, z, numAssigned, &s->box, keys));
    CHECK_RC(sphb200_sort_by_key(ctx, keys, s->perm, numAssigned));

    // low-res global tree (assignment, load balance); single rank: no reduction needed
    updateTree(s, s->global, s->cfg.bucketSize, keys, numAssigned);
}

//! @brief reorder the given conserved fields into SFC order (gatherArrays, domain/layout.hpp:209)
static void gatherConserved(sphb200_sim* s, const int* ids, int count)
{
    size_t numAssigned = s->end - s->start;
    const void* src[kNumConserved];
    void* dst[kNumConserved];
    int eb[kNumConserved];
    for (int k = 0; k < count; ++k)
    {
        int id = ids[k];
        eb[k]  = s->elemBytes(id);
        src[k] = static_cast<char*>(s->field[id]) + size_t(s->start) * eb[k];
        dst[k] = static_cast<char*>(s->fieldAlt[id]) + size_t(s->start) * eb[k];
    }
    CHECK_RC(sphb200_gather(s->ctx, s->perm, numAssigned, count, src, dst, eb));
    for (int k = 0; k < count; ++k)
        std::swap(s->field[ids[k]], s->fieldAlt[ids[k]]);
}

//! @brief focus tree = bucketSizeFocus tree over the assigned keys (single rank: whole SFC range is in focus)
static void syncFocus(sphb200_sim* s)
{
    auto* ctx = s->ctx;
    size_t numAssigned = s->end - s->start;
    uint64_t* keys     = s->keys + s->start;
    updateTree(s, s->focus, s->cfg.bucketSizeFocus, keys, numAssigned);
    int L = s->focus.numLeaves;
    CHECK_RC(sphb200_build_octree(ctx, s->focus.leaves, L, s->prefixes, s->childOffsets, s->parents, s->levelRange,
                                  s->internalToLeaf, s->leafToInternal));
    CHECK_RC(sphb200_node_centers(ctx, s->prefixes, L + (L - 1) / 7, &s->box, s->centers, s->sizes));
    CHECK_RC(sphb200_layout(ctx, s->focus.counts, L, s->layout));
    s->n = s->end; // single rank: no halos
}

extern "C" int sphb200_sim_sync(sphb200_sim* s)
{
    SPHB200_TRY
    s->ctx->packValid = 0;
    s->mark(0);
    syncKeys(s);
    int all[kNumConserved];
    for (int id = 0; id < kNumConserved; ++id)
        all[id] = id;
    gatherConserved(s, all, kNumConserved);
    syncFocus(s);
    s->mark(1);
    SPHB200_CATCH
}

extern "C" int sphb200_sim_forces_phase(sphb200_sim* s, int phase)
{
    SPHB200_TRY
    auto* ctx               = s->ctx;
    sphb200_fields f        = s->table();
    sphb200_hydro_const hc  = s->hydroConst();
    sphb200_tree_view tv    = s->treeView();
    int dt                  = s->cfg.dtype;
    uint32_t a = s->start, b = s->end;
    switch (phase)
    {
        case 0:
            CHECK_RC(sphb200_find_neighbors(ctx, &tv, f.x, f.y, f.z, f.h, dt, a, b, &s->box, hc.ng0, hc.ngmax, 1,
                                            s->neighbors, f.nc + a, nullptr));
            ctx->packValid = 0; // h changed
            s->mark(2);
            CHECK_RC(sphb200_xmass(ctx, dt, a, b, &s->box, &hc, s->neighbors, &f));
            s->mark(3);
            break;
        case 1:
            CHECK_RC(sphb200_ve_def_gradh(ctx, dt, a, b, &s->box, &hc, s->neighbors, &f));
            s->mark(4);
            CHECK_RC(sphb200_eos(ctx, dt, a, b, &hc, &f));
            s->mark(5);
            break;
        case 2:
            CHECK_RC(sphb200_iad_divv_curlv(ctx, dt, a, b, &s->box, &hc, s->neighbors, &f));
            s->mark(6);
            break;
        case 3:
            CHECK_RC(sphb200_av_switches(ctx, dt, a, b, &s->box, &hc, s->neighbors, &f));
            s->mark(7);
            break;
        case 4:
            CHECK_RC(sphb200_momentum_energy(ctx, dt, a, b, &s->box, &hc, s->neighbors, &f, nullptr));
            s->mark(8);
            // TurbVeProp::computeForces (turb_ve.hpp:113-118): stirring with the current minDt
            if (s->turb) CHECK_RC(sphb200_turb_drive(s->turb, ctx, dt, a, b, f.x, f.y, f.z, f.ax, f.ay, f.az, s->minDt));
            break;
        default: throw std::runtime_error("forces_phase: phase must be 0..4");
    }
    SPHB200_CATCH
}

extern "C" int sphb200_sim_compute_forces(sphb200_sim* s)
{
    SPHB200_TRY
    for (int p = 0; p < 5; ++p)
        CHECK_RC(sphb200_sim_forces_phase(s, p));
    SPHB200_CATCH
}

extern "C" int sphb200_sim_integrate(sphb200_sim* s, double globalMinDt)
{
    SPHB200_TRY
    auto* ctx              = s->ctx;
    sphb200_fields f       = s->table();
    sphb200_hydro_const hc = s->hydroConst();
    double dt;
    if (globalMinDt < 0)
    {
        CHECK_RC(sphb200_timestep(ctx, s->cfg.dtype, s->start, s->end, &hc, &f, s->dtLocal));
        dt = s->dtLocal[2];
    }
    else { dt = globalMinDt; }
    // computeTimestep (ts_global.hpp:108-111)
    s->ttot += dt;
    s->minDt_m1 = s->minDt;
    s->minDt    = dt;
    CHECK_RC(sphb200_positions_update_h(ctx, s->cfg.dtype, s->start, s->end, &s->box, &hc, s->minDt, s->minDt_m1, &f));
    s->mark(9);
    SPHB200_CATCH
}

extern "C" int sphb200_sim_local_timestep(sphb200_sim* s, double* out3)
{
    SPHB200_TRY
    sphb200_fields f       = s->table();
    sphb200_hydro_const hc = s->hydroConst();
    CHECK_RC(sphb200_timestep(s->ctx, s->cfg.dtype, s->start, s->end, &hc, &f, s->dtLocal));
    for (int k = 0; k < 3; ++k)
        out3[k] = s->dtLocal[k];
    SPHB200_CATCH
}

extern "C" int sphb200_sim_step(sphb200_sim* s)
{
    SPHB200_TRY
    CHECK_RC(sphb200_sim_sync(s));
    CHECK_RC(sphb200_sim_compute_forces(s));
    CHECK_RC(sphb200_sim_integrate(s, -1.0));
    SPHB200_CATCH
}

/*! @brief one time step on host-resident state: upload, Propagator::sync/computeForces/integrate, download
 *
 * Same work and same results as upload + step + download, but the three transfers are staged so that the PCIe copies
 * run beside the kernels: {x,y,z} go first (keys and sort need nothing else), {h,m,temp} next (tree, neighbor search,
 * XMass, gradh, EOS), the velocity/history/alpha group is only needed from IAD on and rides behind ~20 ms of compute.
 * m is final after the reorder and alpha after AVswitches, so they are sent back while MomentumEnergy runs; the other
 * fields are rewritten by the position update, the last kernel of the step (its time step is a global minimum over
 * MomentumEnergy), and follow after it.
 */
extern "C" int sphb200_sim_step_host(sphb200_sim* s, size_t n, const sphb200_fields* host_in,
                                     const sphb200_fields* host_out)
{
    SPHB200_TRY
    if (n > s->capacity) throw std::runtime_error("sim_step_host: n exceeds capacity");
    auto* ctx = s->ctx;
    if (!s->copyIn)
    {
        SPHB200_CUDA(cudaStreamCreateWithFlags(&s->copyIn, cudaStreamNonBlocking));
        SPHB200_CUDA(cudaStreamCreateWithFlags(&s->copyOut, cudaStreamNonBlocking));
        for (auto& e : s->evUp) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        for (auto& e : s->evDown) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    // everything issued earlier on the compute stream (e.g. a previous step) must be done before fields are overwritten
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    // SPHB200_HOST_PROFILE=1: timeline of the transfers and kernels of this call on stderr
    static const bool profile = std::getenv("SPHB200_HOST_PROFILE") != nullptr;
    std::vector<std::pair<const char*, cudaEvent_t>> tl;
    auto stamp = [&](const char* name, cudaStream_t st)
    {
        if (!profile) return;
        cudaEvent_t e;
        SPHB200_CUDA(cudaEventCreate(&e));
        SPHB200_CUDA(cudaEventRecord(e, st));
        tl.emplace_back(name, e);
    };
    stamp("start", ctx->stream);
    ctx->packValid = 0;
    s->n = s->end = uint32_t(n);
    s->start      = 0;

    static const int groupA[] = {F_x, F_y, F_z};
    static const int groupH[]  = {F_h};
    static const int groupMT[] = {F_m, F_temp};
    static const int groupAH[] = {F_x, F_y, F_z, F_h};
    static const int groupC[] = {F_vx, F_vy, F_vz, F_x_m1, F_y_m1, F_z_m1, F_du_m1, F_alpha};
    sphb200_fields hin = *host_in, hout = *host_out;
    auto upload = [&](const int* ids, int count, cudaEvent_t done)
    {
        for (int k = 0; k < count; ++k)
        {
            void* src = *fieldSlot(hin, ids[k]);
            if (!src) throw std::runtime_error("sim_step_host: all conserved fields are required in host_in");
            SPHB200_CUDA(cudaMemcpyAsync(s->field[ids[k]], src, n * s->elemBytes(ids[k]), cudaMemcpyHostToDevice,
                                         s->copyIn));
        }
        SPHB200_CUDA(cudaEventRecord(done, s->copyIn));
    };
    auto download = [&](int id, cudaStream_t stream)
    {
        void* dst = *fieldSlot(hout, id);
        if (dst)
            SPHB200_CUDA(cudaMemcpyAsync(dst, s->field[id], n * s->elemBytes(id), cudaMemcpyDeviceToHost, stream));
    };
    // upload order = order of first use: {x,y,z} (keys, sort), h (tree search), {m,temp} (XMass, EOS), the rest (IAD on)
    upload(groupA, 3, s->evUp[0]);
    stamp("up xyz", s->copyIn);
    upload(groupH, 1, s->evUp[1]);
    stamp("up h", s->copyIn);
    upload(groupMT, 2, s->evUp[3]);
    stamp("up m,temp", s->copyIn);
    upload(groupC, 8, s->evUp[2]);
    stamp("up rest", s->copyIn);

    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[0], 0));
    s->mark(0);
    syncKeys(s);
    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[1], 0));
    gatherConserved(s, groupAH, 4);
    syncFocus(s);
    s->mark(1);
    {
        // forces phase 0 (FindNeighbors, XMass) with the reorder of m and temp slipped in between: the search needs
        // positions and h only
        sphb200_fields f       = s->table();
        sphb200_hydro_const hc = s->hydroConst();
        sphb200_tree_view tv   = s->treeView();
        CHECK_RC(sphb200_find_neighbors(ctx, &tv, f.x, f.y, f.z, f.h, s->cfg.dtype, s->start, s->end, &s->box, hc.ng0,
                                        hc.ngmax, 1, s->neighbors, f.nc + s->start, nullptr));
        ctx->packValid = 0; // h changed
        s->mark(2);
        SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[3], 0));
        gatherConserved(s, groupMT, 2);
        SPHB200_CUDA(cudaEventRecord(s->evDown[0], ctx->stream));
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[0], 0));
        download(F_m, s->copyOut);
        f = s->table();
        CHECK_RC(sphb200_xmass(ctx, s->cfg.dtype, s->start, s->end, &s->box, &hc, s->neighbors, &f));
        s->mark(3);
    }
    const bool ranged = !s->cfg.hc.avClean && n >= 16 * 128;
    if (ranged && *fieldSlot(hout, F_h))
    {
        // the next step's h = updateH(ng0, nc, h) is known once the neighbors are counted: evaluate it into the spare
        // buffer of h (free since the reorder) and send it now; the position update recomputes the same value in place
        sphb200_fields f = s->table();
        CHECK_RC(sphb200_update_h_into(ctx, s->cfg.dtype, s->start, s->end, s->cfg.hc.ng0, f.nc, f.h, s->fieldAlt[F_h]));
        SPHB200_CUDA(cudaEventRecord(s->evDown[2], ctx->stream));
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[2], 0));
        SPHB200_CUDA(cudaMemcpyAsync(*fieldSlot(hout, F_h), s->fieldAlt[F_h], n * s->elemBytes(F_h),
                                     cudaMemcpyDeviceToHost, s->copyOut));
    }
    CHECK_RC(sphb200_sim_forces_phase(s, 1));
    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[2], 0));
    gatherConserved(s, groupC, 8);
    ctx->packValid &= 1u; // only the position record (x, y, z, m) is still current
    stamp("gradh+eos done", ctx->stream);
    CHECK_RC(sphb200_sim_forces_phase(s, 2));
    stamp("iad done", ctx->stream);
    // Early Courant step: AVswitches evaluates momentumEnergy's signal velocity over the same pairs, so dt is known
    // before the last pair kernel starts; that kernel then runs range by range, each range is integrated right away
    // and leaves for the host while the next one computes (the final download no longer waits for the whole kernel).
    if (!ranged)
    {
        CHECK_RC(sphb200_sim_forces_phase(s, 3));
        SPHB200_CUDA(cudaEventRecord(s->evDown[1], ctx->stream));
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[1], 0));
        download(F_alpha, s->copyOut);
        CHECK_RC(sphb200_sim_forces_phase(s, 4));
        CHECK_RC(sphb200_sim_integrate(s, -1.0));
        for (int id = 0; id < kNumConserved; ++id)
            if (id != F_m && id != F_alpha) download(id, ctx->stream);
    }
    else
    {
        sphb200_fields f       = s->table();
        sphb200_hydro_const hc = s->hydroConst();
        int dtype              = s->cfg.dtype;
        uint32_t a = s->start, b = s->end;
        CHECK_RC(sphb200_av_switches_courant(ctx, dtype, a, b, &s->box, &hc, s->neighbors, &f));
        s->mark(7);
        stamp("av+courant done", ctx->stream);
        // computeTimestep (ts_global.hpp:96-112): Courant key from AVswitches, divv limit, growth cap.  Read back
        // before alpha is queued on the device-to-host engine: the 24 bytes would wait behind its 64 MB
        CHECK_RC(sphb200_timestep(ctx, dtype, a, b, &hc, &f, s->dtLocal));
        stamp("dt known", ctx->stream);
        SPHB200_CUDA(cudaEventRecord(s->evDown[1], ctx->stream));
        SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[1], 0));
        download(F_alpha, s->copyOut);
        double dt   = s->dtLocal[2];
        s->ttot += dt;
        s->minDt_m1 = s->minDt;
        double dtStir = s->minDt; // TurbVeProp stirs with the time step of the previous iteration (turb_ve.hpp:113-118)
        s->minDt    = dt;

        constexpr int kRanges = 16;
        if (!s->evRange[0])
            for (auto& e : s->evRange) SPHB200_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        uint32_t per = ((b - a + kRanges - 1) / kRanges + 127) / 128 * 128;
        if (s->turb) CHECK_RC(sphb200_turb_advance(s->turb, ctx, dtStir)); // one OU step per time step
        for (int r = 0; r < kRanges; ++r)
        {
            uint32_t rb = std::min(b, a + r * per), re = std::min(b, rb + per);
            if (re <= rb) break;
            CHECK_RC(sphb200_momentum_energy_range(ctx, dtype, a, b, rb, re, r == 0, &s->box, &hc, s->neighbors, &f));
            if (s->turb)
                CHECK_RC(sphb200_turb_stir(s->turb, ctx, dtype, rb, re, f.x, f.y, f.z, f.ax, f.ay, f.az));
            CHECK_RC(sphb200_positions_update_h(ctx, dtype, rb, re, &s->box, &hc, s->minDt, s->minDt_m1, &f));
            SPHB200_CUDA(cudaEventRecord(s->evRange[r], ctx->stream));
            if (r == 0) stamp("range 0 computed", ctx->stream);
            SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evRange[r], 0));
            for (int id = 0; id < kNumConserved; ++id)
            {
                if (id == F_m || id == F_alpha || id == F_h) continue;
                void* dst = *fieldSlot(hout, id);
                if (!dst) continue;
                size_t eb = s->elemBytes(id);
                SPHB200_CUDA(cudaMemcpyAsync(static_cast<char*>(dst) + size_t(rb) * eb,
                                             static_cast<char*>(s->field[id]) + size_t(rb) * eb, size_t(re - rb) * eb,
                                             cudaMemcpyDeviceToHost, s->copyOut));
            }
        }
        s->mark(8);
        s->mark(9);
        stamp("last range computed", ctx->stream);
        stamp("last range down", s->copyOut);
    }
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(s->copyOut));
    if (profile)
    {
        std::string line = "step_host timeline (ms):";
        for (auto& [name, e] : tl)
        {
            float ms = 0;
            cudaEventSynchronize(e);
            cudaEventElapsedTime(&ms, tl[0].second, e);
            char buf[96];
            std::snprintf(buf, sizeof(buf), " [%s %.2f]", name, ms);
            line += buf;
            if (e != tl[0].second) cudaEventDestroy(e);
        }
        cudaEventDestroy(tl[0].second);
        std::fprintf(stderr, "%s\n", line.c_str());
    }
    SPHB200_CATCH
}

extern "C" int sphb200_sim_attach_turbulence(sphb200_sim* s, sphb200_turb* t)
{
    s->turb = t;
    return 0;
}

//! @brief computeConservedQuantities for the owned particles of this rank (out10 may be NULL: result stays on device)
extern "C" int sphb200_sim_conserved(sphb200_sim* s, double* out10)
{
    SPHB200_TRY
    sphb200_fields f       = s->table();
    sphb200_hydro_const hc = s->hydroConst();
    CHECK_RC(sphb200_conserved_quantities(s->ctx, s->cfg.dtype, s->start, s->end, &hc, &f, out10));
    SPHB200_CATCH
}

extern "C" int sphb200_sim_mach_rms(sphb200_sim* s, uint64_t numParticlesGlobal, double* out_h)
{
    SPHB200_TRY
    sphb200_fields f = s->table();
    double sum       = 0;
    CHECK_RC(sphb200_mach_square_sum(s->ctx, s->cfg.dtype, s->start, s->end, &f, &sum));
    uint64_t ng = numParticlesGlobal ? numParticlesGlobal : uint64_t(s->end - s->start);
    *out_h      = std::sqrt(sum / double(ng));
    SPHB200_CATCH
}

extern "C" uint32_t sphb200_sim_start_index(const sphb200_sim* s) { return s->start; }
extern "C" uint32_t sphb200_sim_end_index(const sphb200_sim* s) { return s->end; }
extern "C" uint32_t sphb200_sim_n_with_halos(const sphb200_sim* s) { return s->n; }
extern "C" int sphb200_sim_num_leaves(const sphb200_sim* s) { return s->focus.numLeaves; }
extern "C" int sphb200_sim_num_global_leaves(const sphb200_sim* s) { return s->global.numLeaves; }
extern "C" double sphb200_sim_min_dt(const sphb200_sim* s) { return s->minDt; }
extern "C" double sphb200_sim_ttot(const sphb200_sim* s) { return s->ttot; }
extern "C" const uint64_t* sphb200_sim_keys(const sphb200_sim* s) { return s->keys; }
extern "C" const uint32_t* sphb200_sim_neighbors(const sphb200_sim* s) { return s->neighbors; }
extern "C" const uint64_t* sphb200_sim_global_leaves(const sphb200_sim* s) { return s->global.leaves; }
extern "C" const uint32_t* sphb200_sim_global_counts(const sphb200_sim* s) { return s->global.counts; }
extern "C" const uint32_t* sphb200_sim_leaf_counts(const sphb200_sim* s) { return s->focus.counts; }
extern "C" void sphb200_sim_get_box(const sphb200_sim* s, sphb200_box* out) { *out = s->box; }
extern "C" void sphb200_sim_set_dt(sphb200_sim* s, double minDt, double minDt_m1)
{
    s->minDt    = minDt;
    s->minDt_m1 = minDt_m1;
}

extern "C" int sphb200_sim_get_config(const sphb200_sim* s, sphb200_sim_config* out)
{
    *out     = s->cfg;
    out->box = s->box;
    out->hc  = s->hydroConst();
    return 0;
}

extern "C" int sphb200_sim_set_constants(sphb200_sim* s, const sphb200_hydro_const* hc)
{
    SPHB200_TRY
    sphb200_hydro_const& c = s->cfg.hc;
    if (hc->ngmax != c.ngmax) throw std::runtime_error("sim_set_constants: ngmax is fixed at creation");
    if (hc->K != c.K) throw std::runtime_error("sim_set_constants: kernel normalisation differs from the tables");
    if (hc->avClean != c.avClean) throw std::runtime_error("sim_set_constants: avClean is fixed at creation");
    const void *wh = c.wh, *whd = c.whd;
    c     = *hc;
    c.wh  = wh;
    c.whd = whd;
    SPHB200_CATCH
}

extern "C" void sphb200_sim_set_time(sphb200_sim* s, double ttot) { s->ttot = ttot; }
extern "C" void sphb200_sim_set_box(sphb200_sim* s, const sphb200_box* box) { s->box = *box; }
extern "C" sphb200_turb* sphb200_sim_turbulence(const sphb200_sim* s) { return s->turb; }

extern "C" int sphb200_sim_fields(sphb200_sim* s, sphb200_fields* out)
{
    *out = s->table();
    s->ctx->packValid = 0; // the caller may write through these pointers
    return 0;
}

extern "C" int sphb200_sim_tree(sphb200_sim* s, sphb200_tree_view* out)
{
    *out = s->treeView();
    return 0;
}

extern "C" int sphb200_sim_set_layout(sphb200_sim* s, uint32_t start, uint32_t end, uint32_t nWithHalos)
{
    SPHB200_TRY
    if (nWithHalos > s->capacity || end > nWithHalos || start > end) throw std::runtime_error("set_layout: bad layout");
    s->start = start;
    s->end   = end;
    s->n     = nWithHalos;
    s->ctx->packValid = 0;
    SPHB200_CATCH
}

extern "C" int sphb200_sim_phase_times(sphb200_sim* s, float* ms)
{
    SPHB200_TRY
    SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
    for (int p = 0; p < kNumPhases; ++p)
    {
        ms[p] = 0;
        if (s->evValid[p] && s->evValid[p + 1]) SPHB200_CUDA(cudaEventElapsedTime(&ms[p], s->ev[p], s->ev[p + 1]));
    }
    SPHB200_CATCH
}

namespace sphb200
{
__global__ void sumNcKernel(const uint32_t* __restrict__ nc, uint32_t first, uint32_t last, unsigned long long* out)
{
    unsigned long long s = 0;
    for (size_t i = size_t(first) + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < last;
         i += size_t(gridDim.x) * blockDim.x)
        s += nc[i];
    for (int o = 16; o > 0; o >>= 1)
        s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, s);
}
} // namespace sphb200

extern "C" uint64_t sphb200_sim_total_neighbors(sphb200_sim* s)
{
    try
    {
        auto* d = reinterpret_cast<unsigned long long*>(s->ctx->devDt + 16);
        SPHB200_CUDA(cudaMemsetAsync(d, 0, 8, s->ctx->stream));
        if (s->end > s->start)
            SPHB200_LAUNCH(s->ctx, sumNcKernel, persistentGrid(s->end - s->start, 256, 4), 256, 0,
                           static_cast<const uint32_t*>(s->field[F_nc]), s->start, s->end, d);
        unsigned long long r = 0;
        SPHB200_CUDA(cudaMemcpyAsync(&r, d, 8, cudaMemcpyDeviceToHost, s->ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
        return r;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 0;
    }
}

// =====================================================================================================
// HydroVeBdtProp (main/src/propagator/ve_hydro_bdt.hpp:51-420) for one rank, g = 0
// =====================================================================================================
namespace
{

//! @brief cstone::butterfly (domain/include/cstone/primitives/math.hpp:26-31)
int butterfly(int i) { return i == 0 ? 0 : __builtin_ctz(unsigned(i)) + 1; }

//! @brief HydroVeBdtProp::activeRung (ve_hydro_bdt.hpp:106-110)
int activeRung(int substep, int numRungs)
{
    if (substep == 0 || substep >= (1 << (numRungs - 1))) return 0;
    return butterfly(substep);
}

//! @brief makeSlicedView (sph/include/sph/groups.hpp:50-58)
sphb200_group_view slicedView(const sphb200_group_view& g, uint32_t first, uint32_t last)
{
    sphb200_group_view r = g;
    r.numGroups          = last - first;
    r.groupStart += first;
    r.groupEnd += first;
    return r;
}

__global__ void gatherBytesKernel(const uint32_t* __restrict__ perm, size_t n, const uint8_t* __restrict__ src,
                                  uint8_t* __restrict__ dst)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[perm[i]];
}

__global__ void fillFloatKernel(float* __restrict__ a, size_t n, float v)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) a[i] = v;
}

void bdtAllocGroups(sphb200_sim* s, uint32_t cap)
{
    auto& b = s->bdt;
    SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
    for (auto& p : b.buf)
    {
        cudaFree(p);
        p = devAlloc<uint32_t>(2 * size_t(cap) + 2);
    }
    cudaFree(b.groupDt);
    cudaFree(b.groupIndices);
    b.groupDt      = devAlloc<float>(cap);
    b.groupIndices = devAlloc<uint32_t>(cap);
    b.capGroups    = cap;
}

//! @brief HydroVeBdtProp::fullSync (ve_hydro_bdt.hpp:172-194)
void bdtFullSync(sphb200_sim* s)
{
    auto* ctx = s->ctx;
    auto& b   = s->bdt;
    CHECK_RC(sphb200_sim_sync(s));
    // "rung" is a conserved field of this propagator (ve_hydro_bdt.hpp:90): it follows the particles
    uint32_t numAssigned = s->end - s->start;
    if (numAssigned)
        SPHB200_LAUNCH(ctx, gatherBytesKernel, gridFor(numAssigned, 256), 256, 0, s->perm, size_t(numAssigned),
                       b.rung + s->start, b.rungAlt + s->start);
    std::swap(b.rung, b.rungAlt);
    b.searchExt = 1.0f; // d.treeView = domain.octreeProperties()

    sphb200_fields f = s->table();
    uint32_t numGroups = 0;
    for (int attempt = 0; attempt < 2; ++attempt)
    {
        int rc = sphb200_compute_groups(ctx, s->start, s->end, f.x, f.y, f.z, s->focus.leaves, s->focus.numLeaves,
                                        s->layout, &s->box, 2.0f, b.buf[0], 2 * b.capGroups + 1, &numGroups);
        if (rc == SPHB200_ERR_CAPACITY || numGroups > b.capGroups)
        {
            if (attempt == 1) throw std::runtime_error("bdt: group capacity");
            bdtAllocGroups(s, numGroups + numGroups / 16 + 64);
            continue;
        }
        if (rc != 0) throw std::runtime_error(sphb200_last_error());
        break;
    }
    b.groups      = sphb200_group_view{s->start, s->end, numGroups, b.buf[0], b.buf[0] + 1};
    b.tsGroups    = sphb200_group_view{0, 0, 0, b.buf[1], b.buf[1] + numGroups};
    b.activeRungs = b.groups;
    b.allActive   = true;
    if (numGroups)
        SPHB200_LAUNCH(ctx, fillFloatKernel, gridFor(numGroups, 256), 256, 0, b.groupDt, size_t(numGroups),
                       std::numeric_limits<float>::max());
}

//! @brief HydroVeBdtProp::partialSync (ve_hydro_bdt.hpp:196-211); one rank: no halo exchange
void bdtPartialSync(sphb200_sim* s)
{
    auto& b = s->bdt;
    s->ctx->packValid = 0;
    s->mark(0);
    b.searchExt *= 1.012f;
    int highestRung = butterfly(b.ts.substep);
    b.activeRungs   = slicedView(b.tsGroups, b.ts.rungRanges[0], b.ts.rungRanges[highestRung]);
    b.allActive     = false;
    s->mark(1);
}

//! @brief rungTimestep (sph/include/sph/ts_rungs.hpp:131-145), one rank
sphb200_sim::Timestep rungTimestep(sphb200_sim* s, uint32_t numGroups, float maxDt)
{
    auto& b = s->bdt;
    CHECK_RC(sphb200_sort_group_dt(s->ctx, b.groupDt, b.groupIndices, numGroups));
    uint32_t rr[5];
    float mdt[2];
    CHECK_RC(sphb200_rung_ranges(s->ctx, -1.0f, b.groupDt, numGroups, 1, 0.4f, rr, mdt));
    int numRungs = std::min(int(log2f(mdt[1] / mdt[0])) + 1, 4);
    CHECK_RC(sphb200_rung_ranges(s->ctx, mdt[0], b.groupDt, numGroups, numRungs, 0.4f, rr, nullptr));
    float minDt = std::min(maxDt, mdt[0]);
    sphb200_sim::Timestep ret;
    ret.nextDt   = minDt;
    ret.totDt    = minDt * float(1 << numRungs);
    ret.numRungs = numRungs;
    std::copy(rr, rr + 5, ret.rungRanges);
    return ret;
}

//! @brief HydroVeBdtProp::computeRungs (ve_hydro_bdt.hpp:285-327)
void bdtComputeRungs(sphb200_sim* s)
{
    auto* ctx    = s->ctx;
    auto& b      = s->bdt;
    auto& ts     = b.ts;
    int highRung = activeRung(ts.substep, ts.numRungs);

    if (highRung == 0)
    {
        b.prev      = ts;
        float maxDt = ts.dt_m1[0] * 1.1f; // d.maxDtIncrease
        ts          = rungTimestep(s, b.groups.numGroups, maxDt);
        if (b.safetySteps > 0)
        {
            ts.numRungs = 1;
            std::fill(ts.rungRanges + 1, ts.rungRanges + 5, b.groups.numGroups);
            b.safetySteps--;
        }
    }
    else
    {
        // minimumGroupDt (ts_rungs.hpp:147-157)
        uint32_t numGroups = ts.rungRanges[highRung], numTot = ts.rungRanges[4];
        CHECK_RC(sphb200_sort_group_dt(ctx, b.groupDt, b.groupIndices, numGroups));
        CHECK_RC(sphb200_sequence(ctx, b.groupIndices + numGroups, numTot - numGroups, numGroups));
        uint32_t rr[5];
        float mdt[2];
        CHECK_RC(sphb200_rung_ranges(ctx, -1.0f, b.groupDt, numGroups, 4, 0.4f, rr, mdt));
        float timeLeft   = ts.totDt - ts.elapsedDt;
        int substepsLeft = (1 << ts.numRungs) - ts.substep;
        ts.nextDt        = std::min(mdt[0], timeLeft / float(substepsLeft));
        std::copy(rr, rr + highRung, ts.rungRanges);
    }

    if (highRung == 0 || highRung > 1)
    {
        if (highRung > 1) std::swap(b.groups, b.tsGroups);
        uint32_t numOut = ts.rungRanges[4];
        // the output must not overwrite its input: take the buffer the input does not live in
        uint32_t* out = (b.groups.groupStart >= b.buf[0] && b.groups.groupStart < b.buf[0] + 2 * size_t(b.capGroups) + 2)
                            ? b.buf[1]
                            : b.buf[0];
        b.tsGroups = sphb200_group_view{0, 0, numOut, out, out + numOut};
        CHECK_RC(sphb200_extract_groups(ctx, &b.groups, b.groupIndices, 0, numOut, out, out + numOut));
    }
    for (int r = 0; r < ts.numRungs; ++r)
        b.rungs[r] = slicedView(b.tsGroups, ts.rungRanges[r], ts.rungRanges[r + 1]);
}

} // namespace

extern "C" int sphb200_sim_bdt_enable(sphb200_sim* s, int safetySteps, double etaAcc, double eps)
{
    SPHB200_TRY
    auto& b = s->bdt;
    if (!b.enabled)
    {
        b.rung       = devAlloc<uint8_t>(s->capacity);
        b.rungAlt    = devAlloc<uint8_t>(s->capacity);
        b.active     = devAlloc<uint8_t>(s->capacity);
        b.gather     = devAlloc<uint8_t>(s->capacity);
        b.dtParticle = devAlloc<char>(s->capacity * size_t(s->cfg.dtype == SPHB200_F64 ? 8 : 4));
        SPHB200_CUDA(cudaMemset(b.rung, 0, s->capacity));
        SPHB200_CUDA(cudaMemset(b.rungAlt, 0, s->capacity));
        bdtAllocGroups(s, uint32_t(s->capacity / 16 + 1024));
    }
    b.enabled     = true;
    b.safetySteps = safetySteps;
    if (etaAcc <= 0) etaAcc = 0.2; // ParticlesData::etaAcc, eps (particles_data.hpp:111-113)
    if (eps <= 0) eps = 0.005;
    b.etaAccEps = float(etaAcc * std::sqrt(eps)); // groupAccTimestep (ts_rungs.hpp:59)
    b.ts        = sphb200_sim::Timestep{};
    for (float& v : b.ts.dt_m1)
        v = float(s->minDt); // HydroVeBdtProp(): timestep_.dt_m1[0] = settings.at("minDt")
    b.prev = b.ts;
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_compute_forces(sphb200_sim* s)
{
    SPHB200_TRY
    auto* ctx = s->ctx;
    auto& b   = s->bdt;
    if (!b.enabled) throw std::runtime_error("bdt_compute_forces: call sphb200_sim_bdt_enable first");
    // HydroVeBdtProp::sync (ve_hydro_bdt.hpp:213-218)
    if (activeRung(b.ts.substep, b.ts.numRungs) == 0) bdtFullSync(s);
    else bdtPartialSync(s);

    sphb200_op_scope scope{nullptr, b.searchExt, b.dtParticle, nullptr};
    if (!b.allActive)
    {
        CHECK_RC(sphb200_mark_groups(ctx, &b.activeRungs, s->n, b.active));
        scope.active = b.active;
        SPHB200_CUDA(cudaMemsetAsync(b.gather, 0, s->n, ctx->stream));
        scope.gatherMask = b.gather;
    }
    CHECK_RC(sphb200_set_scope(ctx, &scope));
    // TurbVeBdtProp (turb_ve.hpp:68-73) stirs the active rungs after the whole of HydroVeBdtProp::computeForces, i.e.
    // after the acceleration time step has been taken: keep the stirring out of forces phase 4
    sphb200_turb* turb = s->turb;
    s->turb            = nullptr;
    int rc = 0;
    try
    {
        sphb200_fields f = s->table();
        int dt           = s->cfg.dtype;
        for (int p = 0; p < 5; ++p)
        {
            CHECK_RC(sphb200_sim_forces_phase(s, p));
            if (p == 2) // groupDivvTimestep after IadVelocityDivCurl (ve_hydro_bdt.hpp:252-253)
                CHECK_RC(sphb200_group_divv_timestep(ctx, dt, float(s->cfg.hc.Krho), &b.activeRungs, f.divv, b.groupDt));
        }
        // Courant minimum per group (momentum_energy_gpu.cu:98-104), then groupAccTimestep (ve_hydro_bdt.hpp:282)
        CHECK_RC(sphb200_group_min_timestep(ctx, dt, &b.activeRungs, b.dtParticle, b.groupDt));
        CHECK_RC(sphb200_group_acc_timestep(ctx, dt, b.etaAccEps, &b.activeRungs, f.ax, f.ay, f.az, b.groupDt));
        if (turb)
            CHECK_RC(sphb200_turb_drive(turb, ctx, dt, s->start, s->end, f.x, f.y, f.z, f.ax, f.ay, f.az, s->minDt));
    }
    catch (std::exception& e)
    {
        setError(e.what());
        rc = 1;
    }
    s->turb = turb;
    sphb200_set_scope(ctx, nullptr);
    if (rc) throw std::runtime_error(sphb200_last_error());
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_integrate(sphb200_sim* s)
{
    SPHB200_TRY
    auto* ctx = s->ctx;
    auto& b   = s->bdt;
    if (!b.enabled) throw std::runtime_error("bdt_integrate: call sphb200_sim_bdt_enable first");
    bdtComputeRungs(s);
    auto& ts = b.ts;

    // HydroVeBdtProp::integrate (ve_hydro_bdt.hpp:329-376)
    sphb200_fields f       = s->table();
    sphb200_hydro_const hc = s->hydroConst();
    int dtype              = s->cfg.dtype;
    auto driftBack         = [](int subStep, int rung) { return subStep % (1 << rung); };
    int lowestDriftRung    = butterfly(ts.substep + 1);
    bool isLastSubstep     = activeRung(ts.substep + 1, ts.numRungs) == 0;
    sphb200_box openBox;
    {
        double lim[6] = {0, 1, 0, 1, 0, 1};
        int bc[3]     = {SPHB200_OPEN, SPHB200_OPEN, SPHB200_OPEN};
        sphb200_make_box(lim, bc, &openBox);
    }
    const sphb200_box* substepBox = isLastSubstep ? &s->box : &openBox;

    for (int i = 0; i < ts.numRungs; ++i)
    {
        bool useRung = ts.substep == driftBack(ts.substep, i); // drift back to the start of the hierarchy
        bool advance = i < lowestDriftRung;
        float dt     = ts.nextDt;
        float dt_m1[4];
        std::copy(useRung ? b.prev.dt_m1 : ts.dt_m1, (useRung ? b.prev.dt_m1 : ts.dt_m1) + 4, dt_m1);
        if (advance)
        {
            if (ts.dt_drift[i] > 0)
                CHECK_RC(sphb200_drift_positions(ctx, dtype, &b.rungs[i], 0.0f, ts.dt_drift[i], dt_m1, b.rung, &hc, &f));
            CHECK_RC(sphb200_compute_positions_groups(ctx, dtype, &b.rungs[i], ts.dt_drift[i] + dt, dt_m1, b.rung,
                                                      substepBox, &hc, &f));
            ts.dt_m1[i]    = ts.dt_drift[i] + dt;
            ts.dt_drift[i] = 0;
            CHECK_RC(sphb200_store_rung(ctx, &b.rungs[i], uint8_t(i), b.rung));
        }
        else
        {
            CHECK_RC(sphb200_drift_positions(ctx, dtype, &b.rungs[i], ts.dt_drift[i] + dt, ts.dt_drift[i], dt_m1, b.rung,
                                             &hc, &f));
            ts.dt_drift[i] += dt;
        }
    }
    CHECK_RC(sphb200_update_h_groups(ctx, dtype, &b.activeRungs, hc.ng0, f.nc, f.h));

    ts.substep++;
    ts.elapsedDt += ts.nextDt;
    s->ttot += ts.nextDt;
    s->minDt_m1 = s->minDt;
    s->minDt    = ts.nextDt;
    s->mark(9);
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_step(sphb200_sim* s)
{
    SPHB200_TRY
    CHECK_RC(sphb200_sim_bdt_compute_forces(s));
    CHECK_RC(sphb200_sim_bdt_integrate(s));
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_info(const sphb200_sim* s, sphb200_bdt_info* out)
{
    SPHB200_TRY
    const auto& b = s->bdt;
    out->nextDt = b.ts.nextDt; out->elapsedDt = b.ts.elapsedDt; out->totDt = b.ts.totDt;
    out->numRungs = b.ts.numRungs; out->substep = b.ts.substep;
    std::copy(b.ts.rungRanges, b.ts.rungRanges + 5, out->rungRanges);
    std::copy(b.ts.dt_m1, b.ts.dt_m1 + 4, out->dt_m1);
    std::copy(b.ts.dt_drift, b.ts.dt_drift + 4, out->dt_drift);
    out->numGroups       = b.tsGroups.numGroups;
    out->numActiveGroups = b.activeRungs.numGroups;
    out->searchExtFactor = b.searchExt;
    out->isSynced        = activeRung(b.ts.substep, b.ts.numRungs) == 0;
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_restore(sphb200_sim* s, int numRungs, const float dt_m1[4])
{
    SPHB200_TRY
    auto& b = s->bdt;
    if (!b.enabled) throw std::runtime_error("bdt_restore: call sphb200_sim_bdt_enable first");
    if (numRungs < 1 || numRungs > 4) throw std::runtime_error("bdt_restore: numRungs out of range");
    // Timestep::loadOrStore restores numRungs and dt_m1 only; everything else keeps its initial value and the next
    // computeForces starts a new hierarchy (substep 0)
    b.ts          = sphb200_sim::Timestep{};
    b.ts.numRungs = numRungs;
    std::copy(dt_m1, dt_m1 + 4, b.ts.dt_m1);
    b.prev = b.ts;
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_upload_rungs(sphb200_sim* s, const uint8_t* host, size_t n)
{
    SPHB200_TRY
    auto& b = s->bdt;
    if (!b.enabled) throw std::runtime_error("bdt_upload_rungs: call sphb200_sim_bdt_enable first");
    if (n > s->capacity) throw std::runtime_error("bdt_upload_rungs: n exceeds capacity");
    SPHB200_CUDA(cudaMemcpyAsync(b.rung, host, n, cudaMemcpyHostToDevice, s->ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
    SPHB200_CATCH
}

extern "C" int sphb200_sim_bdt_download_rungs(sphb200_sim* s, uint8_t* host)
{
    SPHB200_TRY
    auto& b = s->bdt;
    if (!b.enabled) throw std::runtime_error("bdt_download_rungs: call sphb200_sim_bdt_enable first");
    SPHB200_CUDA(cudaMemcpyAsync(host, b.rung + s->start, s->end - s->start, cudaMemcpyDeviceToHost, s->ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(s->ctx->stream));
    SPHB200_CATCH
}

extern "C" const uint8_t* sphb200_sim_bdt_rungs(const sphb200_sim* s) { return s->bdt.rung; }
extern "C" const float* sphb200_sim_bdt_group_dt(const sphb200_sim* s) { return s->bdt.groupDt; }
extern "C" const uint8_t* sphb200_sim_bdt_active(const sphb200_sim* s) { return s->bdt.allActive ? nullptr : s->bdt.active; }
extern "C" int sphb200_sim_bdt_groups(const sphb200_sim* s, sphb200_group_view* out)
{
    SPHB200_TRY
    *out = s->bdt.tsGroups;
    SPHB200_CATCH
}
