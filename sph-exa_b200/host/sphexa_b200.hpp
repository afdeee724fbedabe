/*! @file C++20 host shim over the C ABI (include/sphexa_b200.h).
 *
 * Mirrors the reference's boundary for the hot path so that front-end code written against
 * cstone::Domain<KeyType,T,GpuTag> (domain/include/cstone/domain/domain.hpp:65-683) and
 * sphexa::HydroVeProp (main/src/propagator/ve_hydro.hpp:54-219) reads the same:
 *
 *   sphb200::host::HydroVeProp prop(ctx, cfg);      // propagatorFactory("ve"), factory.hpp:52
 *   prop.upload(n, hostFields);                     // transferAllocatedToDevice, sphexa.cpp:122
 *   prop.sync(); prop.computeForces(); prop.integrate();
 *   prop.startIndex(); prop.endIndex(); prop.nParticlesWithHalos();   // domain.hpp:382-429
 *
 * Error behaviour: every non-zero C status becomes std::runtime_error carrying sphb200_last_error(),
 * like the reference's throws (domain.hpp:110,449; xmass_gpu.cu:126-127).
 */
#pragma once

#include <stdexcept>
#include <string>
#include <array>
#include <memory>
#include <vector>
#include <algorithm>
#include <utility>

#include "sphexa_b200.h"

namespace sphb200::host
{

inline void check(int rc)
{
    if (rc != 0) { throw std::runtime_error(sphb200_last_error()); }
}

//! @brief cstone::Box<double> twin (sfc/box.hpp:111-191)
struct Box : sphb200_box
{
    Box(double xmin, double xmax, double ymin, double ymax, double zmin, double zmax, int bx = SPHB200_OPEN,
        int by = SPHB200_OPEN, int bz = SPHB200_OPEN)
    {
        const double l[6] = {xmin, xmax, ymin, ymax, zmin, zmax};
        const int b[3]    = {bx, by, bz};
        sphb200_make_box(l, b, this);
    }
    Box(double lo, double hi, int b = SPHB200_OPEN)
        : Box(lo, hi, lo, hi, lo, hi, b, b, b)
    {
    }
    double xmin() const { return lim[0]; }
    double xmax() const { return lim[1]; }
    double lx() const { return len[0]; }
    double ilx() const { return inv[0]; }
};

class Context
{
public:
    explicit Context(int device = 0, void* stream = SPHB200_PRIVATE_STREAM) { check(sphb200_create(device, stream, &ctx_)); }
    ~Context() { sphb200_destroy(ctx_); }
    Context(const Context&)            = delete;
    Context& operator=(const Context&) = delete;
    sphb200_ctx* get() const { return ctx_; }
    void sync() const { check(sphb200_sync(ctx_)); }
    uint64_t launches() const { return sphb200_launch_count(ctx_); }

private:
    sphb200_ctx* ctx_{nullptr};
};

// ---- operator level: same names as the reference's extern GPU functions -------------------------------------------

//! @brief cstone::computeSfcKeysGpu (sfc/sfc_gpu.h:39)
inline void computeSfcKeysGpu(const Context& c, const double* x, const double* y, const double* z, uint64_t* keys,
                              size_t n, const Box& box)
{
    check(sphb200_sfc_keys(c.get(), x, y, z, n, &box, keys));
}

/*! @brief sphexa::assembleCuboid (main/src/init/grid.hpp:238-264) on device arrays: the copies of the template block
 *         (coordinates in [0,1)^3) that tile the global box and fall into the SFC range [keyStart, keyEnd).
 *  @return number of particles written to x, y, z (reference order); throws when capacity is too small */
inline size_t assembleCuboid(const Context& c, uint64_t keyStart, uint64_t keyEnd, const Box& globalBox,
                             const int multiplicity[3], const double* xBlock, const double* yBlock,
                             const double* zBlock, size_t blockSize, double* x, double* y, double* z, size_t capacity)
{
    uint64_t count = 0;
    check(sphb200_assemble_cuboid(c.get(), keyStart, keyEnd, &globalBox, multiplicity, xBlock, yBlock, zBlock,
                                  blockSize, x, y, z, capacity, &count));
    return size_t(count);
}

//! @brief sphexa::cutSphere (grid.hpp:267-300), in place on device arrays; returns the new size
inline size_t cutSphere(const Context& c, double radius, double* x, double* y, double* z, size_t n)
{
    uint64_t count = 0;
    check(sphb200_cut_sphere(c.get(), radius, x, y, z, n, &count));
    return size_t(count);
}

//! @brief cstone::sortByKeyGpu with a sequence payload (primitives_gpu.h:87-111)
inline void sortByKeyGpu(const Context& c, uint64_t* keys, uint32_t* ordering, size_t n)
{
    check(sphb200_sort_by_key(c.get(), keys, ordering, n));
}

//! @brief cstone::computeNodeCountsGpu (tree/csarray_gpu.h:56)
inline void computeNodeCountsGpu(const Context& c, const uint64_t* tree, uint32_t* counts, int numNodes,
                                 const uint64_t* keys, size_t n)
{
    check(sphb200_node_counts(c.get(), tree, numNodes, keys, n, counts));
}

//! @brief cstone::updateOctreeGpu (tree/update_gpu.cuh:58-82); returns converged
inline bool updateOctreeGpu(const Context& c, const uint64_t* keys, size_t n, unsigned bucketSize, uint64_t* tree,
                            uint32_t* counts, int capacity, int& numLeaves)
{
    int conv = 0;
    check(sphb200_update_octree(c.get(), keys, n, bucketSize, tree, counts, capacity, &numLeaves, &conv));
    return conv != 0;
}

//! @brief cstone::buildOctreeGpu (tree/octree_gpu.h:50)
inline void buildOctreeGpu(const Context& c, const uint64_t* leaves, int numLeaves, uint64_t* prefixes,
                           int32_t* childOffsets, int32_t* parents, int32_t* levelRange, int32_t* internalToLeaf,
                           int32_t* leafToInternal)
{
    check(sphb200_build_octree(c.get(), leaves, numLeaves, prefixes, childOffsets, parents, levelRange, internalToLeaf,
                               leafToInternal));
}

//! @brief cstone::findPeersMac (traversal/peers.hpp:63-120) on host arrays; assignment = numRanks + 1 keys
inline std::vector<int> findPeersMac(int myRank, const std::vector<uint64_t>& assignment,
                                     const std::vector<uint64_t>& globalLeaves, const Box& box, float invThetaEff)
{
    int numRanks = int(assignment.size()) - 1;
    std::vector<int> flags(numRanks), peers;
    check(sphb200_find_peers_mac(myRank, numRanks, assignment.data(), globalLeaves.data(), int(globalLeaves.size()) - 1,
                                 &box, invThetaEff, flags.data()));
    for (int r = 0; r < numRanks; ++r)
        if (flags[r]) peers.push_back(r);
    return peers;
}

//! @brief cstone::FocusedOctreeSingleNode / the tree part of cstone::FocusedOctree on the device
//!        (focus/octree_focus.hpp:219-288, focus/octree_focus_mpi.hpp:107-240, 408-481)
class FocusedOctree
{
public:
    FocusedOctree(const Context& c, unsigned bucketSize, float theta)
    {
        check(sphb200_focus_create(c.get(), bucketSize, theta, &f_));
    }
    ~FocusedOctree() { sphb200_focus_destroy(f_); }
    FocusedOctree(const FocusedOctree&)            = delete;
    FocusedOctree& operator=(const FocusedOctree&) = delete;
    //! @brief FocusedOctreeSingleNode::update; keys: sorted device array of all particles the counts shall see
    bool update(const Box& box, const uint64_t* keys, size_t n, uint64_t focusStart, uint64_t focusEnd,
                const std::vector<uint64_t>& mandatoryKeys = {})
    {
        int conv = 0;
        check(sphb200_focus_update(f_, &box, keys, n, focusStart, focusEnd, mandatoryKeys.data(),
                                   int(mandatoryKeys.size()), &conv));
        return conv != 0;
    }
    void updateMinMac(const Box& box, uint64_t focusStart, uint64_t focusEnd)
    {
        check(sphb200_focus_update_min_mac(f_, &box, focusStart, focusEnd));
    }
    bool updateTree(uint64_t focusStart, uint64_t focusEnd, const std::vector<uint64_t>& enforcedKeys = {})
    {
        int conv = 0;
        check(sphb200_focus_update_tree(f_, focusStart, focusEnd, enforcedKeys.data(), int(enforcedKeys.size()), &conv));
        return conv != 0;
    }
    void updateCounts(const uint64_t* keys, size_t n) { check(sphb200_focus_update_counts(f_, keys, n)); }
    void setLeafCounts(const uint32_t* leafCounts) { check(sphb200_focus_set_leaf_counts(f_, leafCounts)); }
    //! @brief octreeViewAcc() / treeLeavesAcc() / leafCountsAcc(): device pointers, valid until the next update
    sphb200_focus_view view() const
    {
        sphb200_focus_view v;
        check(sphb200_focus_get_view(f_, &v));
        return v;
    }

private:
    sphb200_focus* f_{nullptr};
};

//! @brief sph::computeXMass ... computeMomentumEnergy (sph/sph_gpu.hpp:13-82) on [first,last)
struct Hydro
{
    const Context& c;
    int dtype;
    const Box& box;
    sphb200_hydro_const hc;
    const uint32_t* neighbors;

    void computeXMass(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        check(sphb200_xmass(c.get(), dtype, first, last, &box, &hc, neighbors, &d));
    }
    void computeVeDefGradh(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        check(sphb200_ve_def_gradh(c.get(), dtype, first, last, &box, &hc, neighbors, &d));
    }
    void computeEOS(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        check(sphb200_eos(c.get(), dtype, first, last, &hc, &d));
    }
    void computeIadDivvCurlv(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        check(sphb200_iad_divv_curlv(c.get(), dtype, first, last, &box, &hc, neighbors, &d));
    }
    void computeAVswitches(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        check(sphb200_av_switches(c.get(), dtype, first, last, &box, &hc, neighbors, &d));
    }
    double computeMomentumEnergy(uint32_t first, uint32_t last, const sphb200_fields& d) const
    {
        double dt = 0;
        check(sphb200_momentum_energy(c.get(), dtype, first, last, &box, &hc, neighbors, &d, &dt));
        return dt;
    }
};

//! @brief sph::TurbulenceData + driveTurbulence (sph/include/sph/hydro_turb/turbulence_data.hpp:57-183, driver.hpp:101-128)
class Turbulence
{
public:
    explicit Turbulence(const sphb200_turb_config& cfg) { check(sphb200_turb_create(&cfg, &t_)); }
    ~Turbulence() { sphb200_turb_destroy(t_); }
    Turbulence(const Turbulence&)            = delete;
    Turbulence& operator=(const Turbulence&) = delete;
    sphb200_turb* get() const { return t_; }
    int numModes() const { return sphb200_turb_num_modes(t_); }
    //! @brief driveTurbulence on device arrays
    void drive(const Context& c, int dtype, uint32_t first, uint32_t last, const double* x, const double* y,
               const double* z, void* ax, void* ay, void* az, double dt)
    {
        check(sphb200_turb_drive(t_, c.get(), dtype, first, last, x, y, z, ax, ay, az, dt));
    }

private:
    sphb200_turb* t_{nullptr};
};

//! @brief sphexa::computeConservedQuantities, local part (observables/conserved_quantities.hpp:113-159):
//!        {eKin, eInt, egrav, linmom[3], angmom[3], ncsum}
inline std::array<double, 10> conservedQuantities(const Context& c, int dtype, uint32_t first, uint32_t last,
                                                  const sphb200_hydro_const& hc, const sphb200_fields& d)
{
    std::array<double, 10> q{};
    check(sphb200_conserved_quantities(c.get(), dtype, first, last, &hc, &d, q.data()));
    return q;
}

// ---- propagator level -------------------------------------------------------------------------------------------------

//! @brief sphexa::HydroVeProp + cstone::Domain for one rank; owns the device state
//! @brief transport of a multi-rank run: NCCL (one process per GPU) or the single-GPU loopback over P virtual ranks
class Comm
{
public:
    static Comm loopback(int numRanks)
    {
        Comm c;
        check(sphb200_comm_create_loopback(numRanks, &c.comm_));
        return c;
    }
    //! @brief id128: the 128 bytes rank 0 got from uniqueId(), broadcast by the front-end (MPI_Bcast in sphexa.cpp)
    static Comm nccl(int device, int rank, int numRanks, const void* id128)
    {
        Comm c;
        check(sphb200_comm_create_nccl(device, rank, numRanks, id128, &c.comm_));
        return c;
    }
    static std::array<char, 128> uniqueId()
    {
        std::array<char, 128> id{};
        check(sphb200_comm_unique_id(id.data()));
        return id;
    }
    Comm(Comm&& o) noexcept : comm_(std::exchange(o.comm_, nullptr)) {}
    Comm(const Comm&)            = delete;
    Comm& operator=(const Comm&) = delete;
    ~Comm() { sphb200_comm_destroy(comm_); }
    sphb200_comm* get() const { return comm_; }
    int numRanks() const { return sphb200_comm_num_ranks(comm_); }

private:
    Comm() = default;
    sphb200_comm* comm_{nullptr};
};

class HydroVeProp
{
public:
    HydroVeProp(const Context& c, const sphb200_sim_config& cfg)
        : ctx_(c)
    {
        check(sphb200_sim_create(c.get(), &cfg, &sim_));
    }
    //! @brief one rank of a multi-rank run: cstone::Domain(rank, nRanks, bucketSize, bucketSizeFocus, theta, box)
    //!        (domain.hpp:97-113) with the communicator attached; collective calls go through DistHydroVeProp
    HydroVeProp(const Context& c, const sphb200_sim_config& cfg, const Comm& comm)
        : ctx_(c)
    {
        check(sphb200_domain_create(c.get(), &cfg, comm.get(), &sim_));
    }
    sphb200_sim* handle() const { return sim_; }
    ~HydroVeProp() { sphb200_sim_destroy(sim_); }
    HydroVeProp(const HydroVeProp&)            = delete;
    HydroVeProp& operator=(const HydroVeProp&) = delete;

    void upload(size_t n, const sphb200_fields& host) { check(sphb200_sim_upload(sim_, n, &host)); }
    void download(const sphb200_fields& host, uint64_t* keys = nullptr) { check(sphb200_sim_download(sim_, &host, keys)); }

    //! @brief Propagator::sync -> Domain::sync (ve_hydro.hpp:116-130)
    void sync() { check(sphb200_sim_sync(sim_)); }
    //! @brief Propagator::computeForces (ve_hydro.hpp:132-205)
    void computeForces()
    {
        sync();
        check(sphb200_sim_compute_forces(sim_));
    }
    //! @brief Propagator::integrate (ve_hydro.hpp:207-217); globalMinDt < 0: single rank
    void integrate(double globalMinDt = -1.0) { check(sphb200_sim_integrate(sim_, globalMinDt)); }
    //! @brief one step on host-resident state (pinned buffers): transfers staged beside the kernels
    void stepHost(size_t n, const sphb200_fields& in, const sphb200_fields& out)
    {
        check(sphb200_sim_step_host(sim_, n, &in, &out));
    }
    //! @brief TurbVeProp (propagator/turb_ve.hpp:96-118): stir after the hydro forces; nullptr detaches
    void attachTurbulence(Turbulence* t) { check(sphb200_sim_attach_turbulence(sim_, t ? t->get() : nullptr)); }
    //! @brief computeConservedQuantities over the owned particles
    std::array<double, 10> conserved()
    {
        std::array<double, 10> q{};
        check(sphb200_sim_conserved(sim_, q.data()));
        return q;
    }

    //! @brief restart dump of the owned particles (sphexa.cpp:232-246; container: csrc/checkpoint.cu)
    void saveRestart(const std::string& path, uint64_t iteration, uint64_t numParticlesGlobal)
    {
        check(sphb200_sim_save(sim_, path.c_str(), iteration, numParticlesGlobal));
    }
    /*! @brief restart from a single-file dump (file_init.hpp:93-140): reads particles [first, first+count), uploads
     *         them and applies time, time steps, constants, box and turbulence state.  @return stored iteration */
    uint64_t loadRestart(const std::string& path, int dtype, uint64_t first = 0, uint64_t count = ~uint64_t(0))
    {
        sphb200_restart* r = nullptr;
        check(sphb200_restart_open(path.c_str(), &r));
        std::unique_ptr<sphb200_restart, void (*)(sphb200_restart*)> guard(r, sphb200_restart_close);
        sphb200_restart_info info;
        check(sphb200_restart_get_info(r, &info));
        if (count == ~uint64_t(0)) count = info.numParticles - first;
        size_t eb = dtype == SPHB200_F64 ? 8 : 4;
        std::vector<std::vector<char>> buf(14);
        sphb200_fields h{};
        void** realSlots[]  = {&h.h, &h.m, &h.vx, &h.vy, &h.vz, &h.x_m1, &h.y_m1, &h.z_m1, &h.du_m1, &h.alpha};
        double** dblSlots[] = {&h.x, &h.y, &h.z, &h.temp};
        int k = 0;
        for (auto** slot : dblSlots)
        {
            buf[k].resize(std::max<size_t>(count * 8, 8));
            *slot = reinterpret_cast<double*>(buf[k++].data());
        }
        for (auto** slot : realSlots)
        {
            buf[k].resize(std::max<size_t>(count * eb, 8));
            *slot = buf[k++].data();
        }
        check(sphb200_restart_read_fields(r, first, count, dtype, &h));
        check(sphb200_sim_upload(sim_, count, &h));
        // block time steps: the conserved field "rung" (ve_hydro_bdt.hpp:94), absent in dumps of HydroVeProp
        if (sphb200_sim_bdt_rungs(sim_) && sphb200_restart_read_bytes(r, "rung", 0, 0, nullptr) == 0)
        {
            std::vector<uint8_t> rung(std::max<uint64_t>(count, 1));
            check(sphb200_restart_read_bytes(r, "rung", first, count, rung.data()));
            check(sphb200_sim_bdt_upload_rungs(sim_, rung.data(), count));
        }
        ctx_.sync();
        check(sphb200_restart_apply(r, sim_));
        return info.iteration;
    }

    // cstone::Domain getters (domain.hpp:382-429)
    uint32_t startIndex() const { return sphb200_sim_start_index(sim_); }
    uint32_t endIndex() const { return sphb200_sim_end_index(sim_); }
    uint32_t nParticles() const { return endIndex() - startIndex(); }
    uint32_t nParticlesWithHalos() const { return sphb200_sim_n_with_halos(sim_); }
    Box box() const
    {
        sphb200_box b;
        sphb200_sim_get_box(sim_, &b);
        return Box(b.lim[0], b.lim[1], b.lim[2], b.lim[3], b.lim[4], b.lim[5], b.bc[0], b.bc[1], b.bc[2]);
    }
    sphb200_tree_view octreeProperties() const
    {
        sphb200_tree_view t;
        check(sphb200_sim_tree(sim_, &t));
        return t;
    }
    double minDt() const { return sphb200_sim_min_dt(sim_); }
    double ttot() const { return sphb200_sim_ttot(sim_); }
    uint64_t totalNeighbors() { return sphb200_sim_total_neighbors(sim_); }
    //! @brief the "# name: t" lines of ipropagator.hpp:100-128, in ms
    std::array<float, 9> phaseTimes()
    {
        std::array<float, 9> t{};
        check(sphb200_sim_phase_times(sim_, t.data()));
        return t;
    }

protected:
    const Context& ctx_;
    sphb200_sim* sim_{nullptr};
};

/*! @brief sphexa::HydroVeBdtProp (main/src/propagator/ve_hydro_bdt.hpp:51-420): SPH-VE with block time steps
 *
 * Same interface as HydroVeProp; computeForces() runs fullSync or partialSync itself like the reference's sync()
 * (ve_hydro_bdt.hpp:213-218) and evaluates the forces of the groups whose rungs are due, integrate() re-bins the
 * groups (computeRungs), advances the due rungs, drifts the others and updates h of the active groups.
 */
class HydroVeBdtProp : public HydroVeProp
{
public:
    HydroVeBdtProp(const Context& c, const sphb200_sim_config& cfg, int safetySteps = 0, double etaAcc = 0.2,
                   double eps = 0.005)
        : HydroVeProp(c, cfg)
    {
        check(sphb200_sim_bdt_enable(sim_, safetySteps, etaAcc, eps));
    }
    void computeForces() { check(sphb200_sim_bdt_compute_forces(sim_)); }
    void integrate() { check(sphb200_sim_bdt_integrate(sim_)); }
    //! @brief Propagator::isSynced (ve_hydro_bdt.hpp:220): output / observables are only consistent then
    bool isSynced() const { return timestep().isSynced != 0; }
    //! @brief sph::Timestep of the running hierarchy (sph/include/sph/timestep.h:42-66) + group counts
    sphb200_bdt_info timestep() const
    {
        sphb200_bdt_info info;
        check(sphb200_sim_bdt_info(sim_, &info));
        return info;
    }
};

/*! @brief HydroVeProp over the ranks this process drives (one with NCCL, all P in rank order with the loopback):
 *         Domain::sync + computeForces with the halo exchanges in between + integrate with the MIN all-reduce of dt
 *         (main/src/propagator/ve_hydro.hpp:116-217 on P ranks)
 */
class DistHydroVeProp
{
public:
    explicit DistHydroVeProp(std::vector<HydroVeProp*> ranks)
        : ranks_(std::move(ranks))
    {
        for (auto* r : ranks_)
            handles_.push_back(r->handle());
    }
    void sync() { check(sphb200_domain_sync(handles_.data(), int(handles_.size()))); }
    void computeForces() { check(sphb200_domain_compute_forces(handles_.data(), int(handles_.size()))); }
    void integrate() { check(sphb200_domain_integrate(handles_.data(), int(handles_.size()))); }
    void step() { check(sphb200_domain_step(handles_.data(), int(handles_.size()))); }
    //! @brief Domain::exchangeHalos (domain.hpp:375-379); field id = position in sphb200_fields
    void exchangeHalos(const std::vector<int>& fieldIds)
    {
        check(sphb200_domain_exchange_halos(handles_.data(), int(handles_.size()), fieldIds.data(), int(fieldIds.size())));
    }
    std::array<double, 10> conserved()
    {
        std::array<double, 10> q{};
        check(sphb200_domain_conserved(handles_.data(), int(handles_.size()), q.data()));
        return q;
    }
    sphb200_domain_info_t info(size_t localRank) const
    {
        sphb200_domain_info_t i;
        check(sphb200_domain_info(handles_[localRank], &i));
        return i;
    }
    //! @brief one round of the focus-tree update of Domain::sync on every rank (domain.hpp:214-227); after sync()
    bool updateFocus(float theta)
    {
        int conv = 0;
        check(sphb200_domain_update_focus(handles_.data(), int(handles_.size()), theta, &conv));
        return conv != 0;
    }
    sphb200_focus_view focusView(size_t localRank, std::vector<int>* peers = nullptr) const
    {
        sphb200_focus_view v;
        std::vector<int> flags(64, 0);
        check(sphb200_domain_focus_view(handles_[localRank], &v, flags.data()));
        if (peers)
        {
            peers->clear();
            for (int r = 0; r < int(flags.size()); ++r)
                if (flags[r]) peers->push_back(r);
        }
        return v;
    }

private:
    std::vector<HydroVeProp*> ranks_;
    std::vector<sphb200_sim*> handles_;
};

} // namespace sphb200::host
