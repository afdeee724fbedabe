// Context management and the tier-2 (Domain + Propagator) layer of the C ABI.
//
// sphb200_sim plays the role of cstone::Domain<uint64_t,double,GpuTag> (domain/include/cstone/domain/domain.hpp)
// plus sphexa::HydroVeProp (main/src/propagator/ve_hydro.hpp) for one rank: it owns the SFC-sorted SoA
// particle arrays, both cornerstone trees, the neighbor lists, and sequences the operators of tier 1.
#include <cmath>
#include <array>
#include <mutex>
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
    for (auto& a : ctx->rec)
        a.release();
    ctx->tabPairs.release();
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

    cudaStream_t copyIn = nullptr, copyOut = nullptr;
    cudaEvent_t evUp[3] = {}, evDown[2] = {};

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
    for (auto& e : s->ev) cudaEventDestroy(e);
    if (s->copyIn)
    {
        cudaStreamDestroy(s->copyIn);
        cudaStreamDestroy(s->copyOut);
        for (auto& e : s->evUp) cudaEventDestroy(e);
        for (auto& e : s->evDown) cudaEventDestroy(e);
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
    ctx->packValid = 0;
    s->n = s->end = uint32_t(n);
    s->start      = 0;

    static const int groupA[] = {F_x, F_y, F_z};
    static const int groupB[] = {F_h, F_m, F_temp};
    static const int groupC[] = {F_vx, F_vy, F_vz, F_x_m1, F_y_m1, F_z_m1, F_du_m1, F_alpha};
    static const int groupAB[] = {F_x, F_y, F_z, F_h, F_m, F_temp};
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
    upload(groupA, 3, s->evUp[0]);
    upload(groupB, 3, s->evUp[1]);
    upload(groupC, 8, s->evUp[2]);

    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[0], 0));
    s->mark(0);
    syncKeys(s);
    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[1], 0));
    gatherConserved(s, groupAB, 6);
    syncFocus(s);
    s->mark(1);
    SPHB200_CUDA(cudaEventRecord(s->evDown[0], ctx->stream));
    SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[0], 0));
    download(F_m, s->copyOut);

    CHECK_RC(sphb200_sim_forces_phase(s, 0));
    CHECK_RC(sphb200_sim_forces_phase(s, 1));
    SPHB200_CUDA(cudaStreamWaitEvent(ctx->stream, s->evUp[2], 0));
    gatherConserved(s, groupC, 8);
    ctx->packValid &= 1u; // only the position record (x, y, z, m) is still current
    CHECK_RC(sphb200_sim_forces_phase(s, 2));
    CHECK_RC(sphb200_sim_forces_phase(s, 3));
    SPHB200_CUDA(cudaEventRecord(s->evDown[1], ctx->stream));
    SPHB200_CUDA(cudaStreamWaitEvent(s->copyOut, s->evDown[1], 0));
    download(F_alpha, s->copyOut);
    CHECK_RC(sphb200_sim_forces_phase(s, 4));
    CHECK_RC(sphb200_sim_integrate(s, -1.0));

    for (int id = 0; id < kNumConserved; ++id)
        if (id != F_m && id != F_alpha) download(id, ctx->stream);
    SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
    SPHB200_CUDA(cudaStreamSynchronize(s->copyOut));
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
