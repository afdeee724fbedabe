/*! @file forwarding definitions for the reference's `extern` GPU operator templates (operator level, tier 1)
 *
 * The reference declares its SPH-VE GPU operators in sph/include/sph/sph_gpu.hpp:36-57 and defines + explicitly
 * instantiates them in sph/include/sph/hydro_ve/{xmass,ve_def_gradh,iad_divv_curlv,av_switches,momentum_energy}_gpu.cu
 * (archived as sph_gpu, linked into sphexa-cuda).  This translation unit is compiled against those SAME declarations
 * and the reference's own ParticlesData<GpuTag>, and defines the SAME explicit instantiations by forwarding to the
 * C ABI of libsphexa_b200.so.  Linking it in place of the five reference objects gives `sphexa-cuda-b200`: the
 * reference's front-end, initial conditions, cstone::Domain, propagator and observables, unmodified, running this
 * repository's kernels for the pair operators (recipe: build_sphexa_cuda.sh of the test infrastructure).
 *
 * Differences a maintainer should know about:
 *  - the reference re-traverses the octree inside every operator; here computeXMass runs ONE search (with the coupled
 *    h / neighbor-count iteration of xmass_gpu.cu:82-92) and the other four operators reuse its lists.  That is valid
 *    because positions and h do not change between computeXMass and computeMomentumEnergy (ve_hydro.hpp:140-192);
 *  - the GroupView is taken as the particle range [firstBody, lastBody) (global time step).  Block time steps
 *    (`groupDt != nullptr`) are not forwarded by this file: use the tier-2 propagator (sphb200_sim_bdt_*).
 */
#include <stdexcept>
#include <string>

#include <cuda_runtime.h>

#include "cstone/cuda/cuda_utils.cuh"
#include "cstone/sfc/box.hpp"
#include "cstone/traversal/groups.hpp"
#include "cstone/tree/octree.hpp"

#include "sph/sph_gpu.hpp"
#include "sph/particles_data.hpp"

#include "sphexa_b200.h"

namespace
{

using Dataset = sphexa::ParticlesData<cstone::GpuTag>;
using Tc      = sph::SphTypes::CoordinateType;
using KeyType = sph::SphTypes::KeyType;

static_assert(std::is_same_v<Tc, double>, "coordinates are double in both builds");
static_assert(std::is_same_v<sph::SphTypes::HydroType, float> && std::is_same_v<sph::SphTypes::XM1Type, float> &&
                  std::is_same_v<sph::SphTypes::Tmass, float>,
              "the reference's default (mixed) precision maps to SPHB200_MIXED");

void check(int rc)
{
    if (rc != 0) { throw std::runtime_error(std::string("sph-exa_b200: ") + sphb200_last_error()); }
}

struct Bridge
{
    sphb200_ctx* ctx     = nullptr;
    uint32_t* neighbors  = nullptr;
    size_t neighborsSize = 0;

    static Bridge& get()
    {
        static Bridge b;
        // the reference issues everything on the legacy default stream and synchronises after each operator
        if (!b.ctx) { check(sphb200_create(0, nullptr, &b.ctx)); }
        return b;
    }

    uint32_t* lists(size_t numTargets, unsigned ngmax)
    {
        size_t need = sphb200_neighbors_size(numTargets, ngmax);
        if (need > neighborsSize)
        {
            if (neighbors) { cudaFree(neighbors); }
            neighborsSize = need + need / 20;
            if (cudaMalloc(&neighbors, neighborsSize * sizeof(uint32_t)) != cudaSuccess)
            {
                throw std::runtime_error("sph-exa_b200: cannot allocate the neighbor lists");
            }
        }
        return neighbors;
    }
};

sphb200_box toBox(const cstone::Box<Tc>& box)
{
    const double lim[6] = {box.xmin(), box.xmax(), box.ymin(), box.ymax(), box.zmin(), box.zmax()};
    const int bc[3]     = {int(box.boundaryX()), int(box.boundaryY()), int(box.boundaryZ())};
    sphb200_box b;
    sphb200_make_box(lim, bc, &b);
    return b;
}

template<class Vector>
auto* ptrOrNull(Vector& v)
{
    return v.size() ? rawPtr(v) : nullptr;
}

//! @brief device field table of the reference's ParticlesData (whatever is allocated right now: FieldStates aliases
//!        buffers between dependent fields, ve_hydro.hpp:159-191)
sphb200_fields fieldsOf(Dataset& d)
{
    auto& g = d.devData;
    sphb200_fields f{};
    f.x = ptrOrNull(g.x); f.y = ptrOrNull(g.y); f.z = ptrOrNull(g.z);
    f.h = ptrOrNull(g.h); f.m = ptrOrNull(g.m);
    f.vx = ptrOrNull(g.vx); f.vy = ptrOrNull(g.vy); f.vz = ptrOrNull(g.vz);
    f.x_m1 = ptrOrNull(g.x_m1); f.y_m1 = ptrOrNull(g.y_m1); f.z_m1 = ptrOrNull(g.z_m1);
    f.temp = ptrOrNull(g.temp); f.du_m1 = ptrOrNull(g.du_m1); f.alpha = ptrOrNull(g.alpha);
    f.xm = ptrOrNull(g.xm); f.kx = ptrOrNull(g.kx); f.gradh = ptrOrNull(g.gradh);
    f.prho = ptrOrNull(g.prho); f.c = ptrOrNull(g.c);
    f.c11 = ptrOrNull(g.c11); f.c12 = ptrOrNull(g.c12); f.c13 = ptrOrNull(g.c13);
    f.c22 = ptrOrNull(g.c22); f.c23 = ptrOrNull(g.c23); f.c33 = ptrOrNull(g.c33);
    f.divv = ptrOrNull(g.divv); f.curlv = ptrOrNull(g.curlv);
    f.ax = ptrOrNull(g.ax); f.ay = ptrOrNull(g.ay); f.az = ptrOrNull(g.az);
    f.du = ptrOrNull(g.du); f.nc = ptrOrNull(g.nc);
    f.dV11 = ptrOrNull(g.dV11); f.dV12 = ptrOrNull(g.dV12); f.dV13 = ptrOrNull(g.dV13);
    f.dV22 = ptrOrNull(g.dV22); f.dV23 = ptrOrNull(g.dV23); f.dV33 = ptrOrNull(g.dV33);
    f.n = g.x.size();
    return f;
}

sphb200_hydro_const constOf(Dataset& d, int avClean)
{
    sphb200_hydro_const hc{};
    hc.K = d.K; hc.ng0 = d.ng0; hc.ngmax = d.ngmax;
    hc.Kcour = d.Kcour; hc.Krho = d.Krho; hc.gamma = d.gamma; hc.muiConst = d.muiConst;
    hc.alphamin = d.alphamin; hc.alphamax = d.alphamax; hc.decay_constant = d.decay_constant;
    hc.Atmin = d.Atmin; hc.Atmax = d.Atmax; hc.ramp = d.ramp;
    hc.minDt = d.minDt; hc.minDt_m1 = d.minDt_m1;
    hc.wh = rawPtr(d.devData.wh); hc.whd = rawPtr(d.devData.whd);
    hc.avClean = avClean;
    return hc;
}

sphb200_tree_view treeOf(Dataset& d)
{
    const auto& t = d.treeView; // cstone::OctreeNsView<Tc, KeyType>, device pointers (ve_hydro.hpp:129)
    sphb200_tree_view v{};
    v.numLeafNodes   = t.numLeafNodes;
    v.numNodes       = t.numLeafNodes + (t.numLeafNodes - 1) / 7;
    v.prefixes       = t.prefixes;
    v.childOffsets   = t.childOffsets;
    v.internalToLeaf = t.internalToLeaf;
    v.levelRange     = t.levelRange;
    v.leaves         = t.leaves;
    v.layout         = t.layout;
    v.centers        = reinterpret_cast<const double*>(t.centers);
    v.sizes          = reinterpret_cast<const double*>(t.sizes);
    return v;
}

void sync(Bridge& b)
{
    check(sphb200_sync(b.ctx)); // the reference's operators are synchronous (xmass_gpu.cu:113)
}

} // namespace

namespace sph
{
namespace cuda
{

template<class D>
void computeXMass(const GroupView& grp, D& d, const cstone::Box<typename D::RealType>& box)
{
    Bridge& b            = Bridge::get();
    sphb200_box bx       = toBox(box);
    sphb200_fields f     = fieldsOf(d);
    sphb200_hydro_const hc = constOf(d, 0);
    sphb200_tree_view tv = treeOf(d);
    uint32_t first = grp.firstBody, last = grp.lastBody;
    uint32_t* nb = b.lists(last - first, d.ngmax);
    unsigned fails = 0;
    // the search + coupled h / neighbor-count iteration of xmassGpu (xmass_gpu.cu:75-100), once for all five operators
    check(sphb200_find_neighbors(b.ctx, &tv, f.x, f.y, f.z, f.h, SPHB200_MIXED, first, last, &bx, d.ng0, d.ngmax, 1, nb,
                                 f.nc + first, &fails));
    if (fails) { throw std::runtime_error("coupled nc/h-updated failed to converge"); } // xmass_gpu.cu:127
    check(sphb200_xmass(b.ctx, SPHB200_MIXED, first, last, &bx, &hc, nb, &f));
    sync(b);
}

template<class D>
void computeVeDefGradh(const GroupView& grp, D& d, const cstone::Box<typename D::RealType>& box)
{
    Bridge& b              = Bridge::get();
    sphb200_box bx         = toBox(box);
    sphb200_fields f       = fieldsOf(d);
    sphb200_hydro_const hc = constOf(d, 0);
    check(sphb200_ve_def_gradh(b.ctx, SPHB200_MIXED, grp.firstBody, grp.lastBody, &bx, &hc, b.neighbors, &f));
    sync(b);
}

template<class D>
void computeIadDivvCurlv(const GroupView& grp, D& d, const cstone::Box<typename D::RealType>& box)
{
    Bridge& b              = Bridge::get();
    sphb200_box bx         = toBox(box);
    sphb200_fields f       = fieldsOf(d);
    sphb200_hydro_const hc = constOf(d, d.devData.dV11.size() ? 1 : 0);
    check(sphb200_iad_divv_curlv(b.ctx, SPHB200_MIXED, grp.firstBody, grp.lastBody, &bx, &hc, b.neighbors, &f));
    sync(b);
}

template<class D>
void computeAVswitches(const GroupView& grp, D& d, const cstone::Box<typename D::RealType>& box)
{
    Bridge& b              = Bridge::get();
    sphb200_box bx         = toBox(box);
    sphb200_fields f       = fieldsOf(d);
    sphb200_hydro_const hc = constOf(d, 0);
    check(sphb200_av_switches(b.ctx, SPHB200_MIXED, grp.firstBody, grp.lastBody, &bx, &hc, b.neighbors, &f));
    sync(b);
}

template<bool avClean, class D>
void computeMomentumEnergy(const GroupView& grp, float* groupDt, D& d, const cstone::Box<typename D::RealType>& box)
{
    if (groupDt) { throw std::runtime_error("sph-exa_b200: block time steps go through sphb200_sim_bdt_* (tier 2)"); }
    Bridge& b              = Bridge::get();
    sphb200_box bx         = toBox(box);
    sphb200_fields f       = fieldsOf(d);
    sphb200_hydro_const hc = constOf(d, avClean ? 1 : 0);
    double minDtCourant    = 0;
    check(sphb200_momentum_energy(b.ctx, SPHB200_MIXED, grp.firstBody, grp.lastBody, &bx, &hc, b.neighbors, &f,
                                  &minDtCourant));
    d.minDtCourant = minDtCourant; // momentum_energy_gpu.cu:140-142
}

template<class D>
void computeDensity(const GroupView&, D&, const cstone::Box<typename D::RealType>&)
{
    throw std::runtime_error("sph-exa_b200: computeDensity (xmass_gpu.cu:136-166) is not on the forwarded path");
}

// the explicit instantiations of the reference's .cu files (xmass_gpu.cu:130,166; ve_def_gradh_gpu.cu:100;
// iad_divv_curlv_gpu.cu:110; av_switches_gpu.cu:102; momentum_energy_gpu.cu:148-152)
template void computeXMass(const GroupView&, Dataset&, const cstone::Box<Tc>&);
template void computeDensity(const GroupView&, Dataset&, const cstone::Box<Tc>&);
template void computeVeDefGradh(const GroupView&, Dataset&, const cstone::Box<Tc>&);
template void computeIadDivvCurlv(const GroupView&, Dataset&, const cstone::Box<Tc>&);
template void computeAVswitches(const GroupView&, Dataset&, const cstone::Box<Tc>&);
template void computeMomentumEnergy<true>(const GroupView&, float*, Dataset&, const cstone::Box<Tc>&);
template void computeMomentumEnergy<false>(const GroupView&, float*, Dataset&, const cstone::Box<Tc>&);

} // namespace cuda
} // namespace sph
