/* TEST INFRASTRUCTURE ONLY — not part of the product path.
 *
 * Thin extern "C" veneer over the UNMODIFIED reference headers that live under
 * /root/reference (cstone + sph).  It is compiled by oracle/Makefile from the
 * sources where they lie into oracle/_ref/libsphexa_ref.so (git-ignored, travels
 * to the GPU box).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it.  No reference source is
 * copied here: every function below only *calls* reference templates:
 *
 *   keys        cstone::computeSfcKeys              domain/include/cstone/sfc/sfc.hpp:283-291
 *   sort        cstone::sort_by_key                 domain/include/cstone/primitives/gather.hpp:58-90
 *   tree        cstone::computeOctree/updateOctree  domain/include/cstone/tree/csarray.hpp:432-468
 *   counts      cstone::computeNodeCounts           domain/include/cstone/tree/csarray.hpp:203-257
 *   octree      cstone::buildOctreeCpu              domain/include/cstone/tree/octree.hpp:184-214
 *   centers     cstone::nodeFpCenters               domain/include/cstone/focus/source_center.hpp:145-158
 *   neighbors   cstone::findNeighbors               domain/include/cstone/findneighbors.hpp:95-189
 *               sph::findNeighborsSph               sph/include/sph/find_neighbors.hpp:10-45
 *   J-loops     sph::xmassJLoop ... momentumAndEnergyJLoop   sph/include/sph/hydro_ve/ *_kern.hpp
 *   EOS         sph::idealGasEOS                    sph/include/sph/eos.hpp:30-40
 *   integrate   sph::positionUpdate/energyUpdate    sph/include/sph/positions.hpp:54-88
 *   h           sph::updateH                        sph/include/sph/kernels.hpp:26-32
 *   tables      sph::tabulateFunction/kernel_3D_k   sph/include/sph/sph_kernel_tables.hpp
 *
 * The per-particle driver loops (the *_Impl functions in sph/include/sph/hydro_ve/*.hpp)
 * are bound to sph::ParticlesData, which this fork snapshot cannot compile
 * (stray #endif, particles_data.hpp:270); they are 10-line OpenMP loops over the
 * J-loop functions and are re-expressed here exactly as sph/test/ve.cpp:112-233 calls them.
 *
 * T = HydroType: double in "f64" (north-star parity) mode, float in "mixed" (reference default).
 */
#include <cmath>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>
#include <array>
#include <functional>
#include <algorithm>
#include <omp.h>

#include "cstone/sfc/sfc.hpp"
#include "cstone/sfc/box.hpp"
#include "cstone/primitives/gather.hpp"
#include "cstone/tree/csarray.hpp"
#include "cstone/tree/octree.hpp"
#include "cstone/focus/source_center.hpp"
#include "cstone/findneighbors.hpp"

#include "sph/kernels.hpp"
#include "sph/table_lookup.hpp"
#include "sph/sph_kernel_tables.hpp"
#include "sph/eos.hpp"
#include "sph/find_neighbors.hpp"
#include "sph/hydro_ve/xmass_kern.hpp"
#include "sph/hydro_ve/ve_def_gradh_kern.hpp"
#include "sph/hydro_ve/iad_kern.hpp"
#include "sph/hydro_ve/divv_curlv_kern.hpp"
#include "sph/hydro_ve/av_switches_kern.hpp"
#include "sph/hydro_ve/momentum_energy_kern.hpp"
#include "sph/positions.hpp"
#include "cstone/primitives/mpi_wrappers.hpp" // MpiType, against oracle/stub/mpi.h (single rank)
#include "cstone/tree/accel_switch.hpp"
#include "observables/conserved_quantities.hpp"
#include <map>
#include <sstream>
#include <iostream>
#include "sph/hydro_turb/turbulence_data.hpp"
#include "sph/hydro_turb/phases.hpp"
#include "sph/hydro_turb/stirring.hpp"
#include "sph/hydro_turb/driver.hpp"
#include "init/grid.hpp"

using KeyType = uint64_t;
using SfcKey  = cstone::HilbertKey<KeyType>;
using cstone::LocalIndex;
using cstone::TreeNodeIndex;

static cstone::Box<double> makeBox(const double* lim, const int* bc)
{
    return cstone::Box<double>(lim[0], lim[1], lim[2], lim[3], lim[4], lim[5], cstone::BoundaryType(bc[0]),
                               cstone::BoundaryType(bc[1]), cstone::BoundaryType(bc[2]));
}

extern "C"
{

int ref_num_threads() { return omp_get_max_threads(); }
void ref_set_num_threads(int n) { omp_set_num_threads(n); }

/* ---------------- SFC keys ---------------- */
void ref_hilbert_encode(const unsigned* ix, const unsigned* iy, const unsigned* iz, size_t n, uint64_t* keys)
{
    for (size_t i = 0; i < n; ++i)
        keys[i] = cstone::iHilbert<KeyType>(ix[i], iy[i], iz[i]);
}

void ref_hilbert_decode(const uint64_t* keys, size_t n, unsigned* ix, unsigned* iy, unsigned* iz)
{
    for (size_t i = 0; i < n; ++i)
    {
        auto [a, b, c] = cstone::decodeHilbert(keys[i]);
        ix[i] = a; iy[i] = b; iz[i] = c;
    }
}

void ref_compute_sfc_keys(const double* x, const double* y, const double* z, size_t n, const double* lim,
                          const int* bc, uint64_t* keys)
{
    auto box = makeBox(lim, bc);
    cstone::computeSfcKeys(x, y, z, cstone::sfcKindPointer(keys), n, box);
}

/* ---------------- sort ---------------- */
void ref_sort_by_key(uint64_t* keys, size_t n, uint32_t* perm)
{
    std::iota(perm, perm + n, 0u);
    cstone::sort_by_key(keys, keys + n, perm);
}

/* ---------------- cornerstone leaf array ---------------- */
// returns number of leaves or -1 if capacity insufficient; *iterations = number of updateOctree calls
int ref_compute_octree(const uint64_t* keys, size_t n, unsigned bucket, uint64_t* leaves, unsigned* counts,
                       int capacity, int* iterations)
{
    std::vector<KeyType> tree{0, cstone::nodeRange<KeyType>(0)};
    std::vector<unsigned> cnt{unsigned(n)};
    int it = 1;
    while (!cstone::updateOctree(keys, keys + n, bucket, tree, cnt))
        ++it;
    if (iterations) *iterations = it;
    int nl = cstone::nNodes(tree);
    if (nl > capacity) return -1;
    std::copy(tree.begin(), tree.end(), leaves);
    std::copy(cnt.begin(), cnt.end(), counts);
    return nl;
}

// one rebalance+count step; returns new number of leaves (or -1), *converged set
int ref_update_octree(const uint64_t* keys, size_t n, unsigned bucket, uint64_t* leaves, unsigned* counts,
                      int numLeaves, int capacity, int* converged)
{
    std::vector<KeyType> tree(leaves, leaves + numLeaves + 1);
    std::vector<unsigned> cnt(counts, counts + numLeaves);
    bool conv  = cstone::updateOctree(keys, keys + n, bucket, tree, cnt);
    *converged = conv;
    int nl     = cstone::nNodes(tree);
    if (nl > capacity) return -1;
    std::copy(tree.begin(), tree.end(), leaves);
    std::copy(cnt.begin(), cnt.end(), counts);
    return nl;
}

void ref_compute_node_counts(const uint64_t* leaves, int numLeaves, const uint64_t* keys, size_t n, unsigned* counts)
{
    cstone::computeNodeCounts(leaves, counts, numLeaves, keys, keys + n, std::numeric_limits<unsigned>::max(), false);
}

void ref_rebalance_decision(const uint64_t* leaves, const unsigned* counts, int numLeaves, unsigned bucket, int* ops,
                            int* converged)
{
    *converged = cstone::rebalanceDecision(leaves, counts, numLeaves, bucket, ops);
}

/* ---------------- internal octree ---------------- */
void ref_build_octree(const uint64_t* leaves, int numLeaves, uint64_t* prefixes, int* childOffsets, int* parents,
                      int* levelRange, int* internalToLeaf, int* leafToInternal)
{
    int numInternal = (numLeaves - 1) / 7;
    cstone::buildOctreeCpu(leaves, numLeaves, numInternal, prefixes, childOffsets, parents, levelRange, internalToLeaf,
                           leafToInternal);
}

void ref_node_centers(const uint64_t* prefixes, int numNodes, const double* lim, const int* bc, double* centers,
                      double* sizes)
{
    auto box = makeBox(lim, bc);
    cstone::nodeFpCenters<KeyType>({prefixes, size_t(numNodes)}, reinterpret_cast<cstone::Vec3<double>*>(centers),
                                   reinterpret_cast<cstone::Vec3<double>*>(sizes), box);
}

/* ---------------- glass-block assembly (main/src/init/grid.hpp:238-300) ---------------- */
size_t ref_assemble_cuboid(uint64_t keyStart, uint64_t keyEnd, const double* lim, const int* bc, const int* multiplicity,
                           const double* xBlock, const double* yBlock, const double* zBlock, size_t blockSize,
                           double* x, double* y, double* z, size_t capacity)
{
    auto box = makeBox(lim, bc);
    std::vector<double> ox, oy, oz;
    sphexa::assembleCuboid<double>(KeyType(keyStart), KeyType(keyEnd), box,
                                   cstone::Vec3<int>{multiplicity[0], multiplicity[1], multiplicity[2]},
                                   gsl::span<const double>(xBlock, blockSize), gsl::span<const double>(yBlock, blockSize),
                                   gsl::span<const double>(zBlock, blockSize), ox, oy, oz);
    if (ox.size() <= capacity)
    {
        std::copy(ox.begin(), ox.end(), x);
        std::copy(oy.begin(), oy.end(), y);
        std::copy(oz.begin(), oz.end(), z);
    }
    return ox.size();
}

void ref_equidistant_sfc_segment(int rank, int numRanks, int numLeavesPerRank, uint64_t* out)
{
    auto [a, b] = sphexa::equiDistantSfcSegments<KeyType>(rank, numRanks, numLeavesPerRank);
    out[0] = a;
    out[1] = b;
}

int ref_span_sfc_range(uint64_t a, uint64_t b, uint64_t* out)
{
    int n = cstone::spanSfcRange(KeyType(a), KeyType(b));
    cstone::spanSfcRange(KeyType(a), KeyType(b), out);
    out[n] = b;
    return n;
}

size_t ref_cut_sphere(double radius, double* x, double* y, double* z, size_t n)
{
    std::vector<double> vx(x, x + n), vy(y, y + n), vz(z, z + n);
    sphexa::cutSphere(radius, vx, vy, vz);
    std::copy(vx.begin(), vx.end(), x);
    std::copy(vy.begin(), vy.end(), y);
    std::copy(vz.begin(), vz.end(), z);
    return vx.size();
}

} // extern "C"

/* ---------------- neighbor search ---------------- */
struct TreeArgs
{
    int numLeaves;
    const uint64_t* prefixes;
    const int* childOffsets;
    const int* internalToLeaf;
    const int* levelRange;
    const uint64_t* leaves;
    const uint32_t* layout;
    const double* centers;
    const double* sizes;
};

static cstone::OctreeNsView<double, KeyType> nsView(const TreeArgs& t)
{
    cstone::OctreeNsView<double, KeyType> v;
    v.numLeafNodes   = t.numLeaves;
    v.prefixes       = t.prefixes;
    v.childOffsets   = t.childOffsets;
    v.internalToLeaf = t.internalToLeaf;
    v.levelRange     = t.levelRange;
    v.leaves         = t.leaves;
    v.layout         = t.layout;
    v.centers        = reinterpret_cast<const cstone::Vec3<double>*>(t.centers);
    v.sizes          = reinterpret_cast<const cstone::Vec3<double>*>(t.sizes);
    v.searchExtFactor = 1.0f;
    return v;
}

template<class T>
static void findNeighborsT(const double* x, const double* y, const double* z, T* h, uint32_t first, uint32_t last,
                           const double* lim, const int* bc, const TreeArgs* t, unsigned ng0, unsigned ngmax,
                           uint32_t* neighbors, unsigned* nc, int iterateH)
{
    auto box  = makeBox(lim, bc);
    auto view = nsView(*t);
    if (iterateH) { sph::findNeighborsSph(x, y, z, h, first, last, box, view, ng0, ngmax, neighbors, nc); }
    else
    {
#pragma omp parallel for
        for (uint32_t i = first; i < last; ++i)
        {
            nc[i - first] =
                1 + cstone::findNeighbors(i, x, y, z, h, view, box, ngmax, neighbors + size_t(i - first) * ngmax);
        }
    }
}

/* ---------------- SPH-VE drivers (loops as in sph/include/sph/hydro_ve/*.hpp *_Impl) ---------------- */
template<class T>
struct Hydro
{
    static void tables(double sincIndex, T* wh, T* whd, double* K)
    {
        auto a = sph::tabulateFunction<T, sph::lt::kTableSize>(sph::getSphKernel(sph::SphKernelType::sinc_n, sincIndex), 0, 2);
        auto b = sph::tabulateFunction<T, sph::lt::kTableSize>(
            sph::getSphKernelDerivative(sph::SphKernelType::sinc_n, sincIndex), 0, 2);
        std::copy(a.begin(), a.end(), wh);
        std::copy(b.begin(), b.end(), whd);
        *K = sph::kernel_3D_k(sph::getSphKernel(sph::SphKernelType::sinc_n, sincIndex), 2.0);
    }

    static void xmass(uint32_t first, uint32_t last, unsigned ngmax, double K, const cstone::Box<double>& box,
                      const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z,
                      const T* h, const T* m, const T* wh, const T* whd, T* xm)
    {
#pragma omp parallel for schedule(static)
        for (uint32_t i = first; i < last; ++i)
        {
            size_t ni    = i - first;
            unsigned ncC = std::min(nc[ni] - 1, ngmax);
            xm[i]        = sph::xmassJLoop(i, K, box, nb + ngmax * ni, ncC, x, y, z, h, m, wh, whd);
        }
    }

    static void gradh(uint32_t first, uint32_t last, unsigned ngmax, double K, const cstone::Box<double>& box,
                      const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z,
                      const T* h, const T* m, const T* wh, const T* whd, const T* xm, T* kx, T* gradh)
    {
#pragma omp parallel for schedule(static)
        for (uint32_t i = first; i < last; ++i)
        {
            size_t ni    = i - first;
            unsigned ncC = std::min(nc[ni] - 1, ngmax);
            auto [kxi, gradhi] = sph::veDefGradhJLoop(i, K, box, nb + ngmax * ni, ncC, x, y, z, h, m, wh, whd, xm);
            kx[i]    = kxi;
            gradh[i] = gradhi;
        }
    }

    // sph/include/sph/hydro_ve/eos.hpp:52-76 ; muiConst is Tmass (=T here), gamma/temp are double
    static void eos(uint32_t first, uint32_t last, const double* temp, const T* m, const T* kx, const T* xm,
                    const T* gradh, T muiConst, double gamma, T* prho, T* c, T* rho, T* p)
    {
#pragma omp parallel for schedule(static)
        for (uint32_t i = first; i < last; ++i)
        {
            auto rhoi     = kx[i] * m[i] / xm[i];
            auto [pi, ci] = sph::idealGasEOS(temp[i], rhoi, muiConst, gamma);
            prho[i]       = pi / (kx[i] * m[i] * m[i] * gradh[i]);
            c[i]          = ci;
            if (rho) rho[i] = rhoi;
            if (p) p[i] = pi;
        }
    }

    static void iadDivvCurlv(uint32_t first, uint32_t last, unsigned ngmax, double K, const cstone::Box<double>& box,
                             const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z,
                             const T* vx, const T* vy, const T* vz, const T* h, const T* wh, const T* whd, const T* xm,
                             const T* kx, T* c11, T* c12, T* c13, T* c22, T* c23, T* c33, T* divv, T* curlv, T* dV11,
                             T* dV12, T* dV13, T* dV22, T* dV23, T* dV33)
    {
        bool doGradV = dV11 != nullptr;
#pragma omp parallel for
        for (uint32_t i = first; i < last; ++i)
        {
            size_t ni    = i - first;
            unsigned ncC = std::min(nc[ni] - 1, ngmax);
            sph::IADJLoop(i, K, box, nb + ngmax * ni, ncC, x, y, z, h, wh, whd, xm, kx, c11, c12, c13, c22, c23, c33);
            sph::divV_curlVJLoop(i, K, box, nb + ngmax * ni, ncC, x, y, z, vx, vy, vz, h, c11, c12, c13, c22, c23, c33,
                                 wh, whd, kx, xm, divv, curlv, dV11, dV12, dV13, dV22, dV23, dV33, doGradV);
        }
    }

    static void avSwitches(uint32_t first, uint32_t last, unsigned ngmax, double K, const cstone::Box<double>& box,
                           const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z,
                           const T* vx, const T* vy, const T* vz, const T* h, const T* c, const T* c11, const T* c12,
                           const T* c13, const T* c22, const T* c23, const T* c33, const T* wh, const T* whd,
                           const T* kx, const T* xm, const T* divv, double minDt, T alphamin, T alphamax,
                           T decay_constant, T* alpha)
    {
#pragma omp parallel for
        for (uint32_t i = first; i < last; ++i)
        {
            size_t ni    = i - first;
            unsigned ncC = std::min(nc[ni] - 1, ngmax);
            alpha[i] = sph::AVswitchesJLoop(i, K, box, nb + ngmax * ni, ncC, x, y, z, vx, vy, vz, h, c, c11, c12, c13,
                                            c22, c23, c33, wh, whd, kx, xm, divv, minDt, alphamin, alphamax,
                                            decay_constant, alpha[i]);
        }
    }

    template<bool avClean>
    static double momentumEnergy(uint32_t first, uint32_t last, unsigned ngmax, double K,
                                 const cstone::Box<double>& box, const uint32_t* nb, const unsigned* nc,
                                 const double* x, const double* y, const double* z, const T* vx, const T* vy,
                                 const T* vz, const T* h, const T* m, const T* prho, const T* c, const T* c11,
                                 const T* c12, const T* c13, const T* c22, const T* c23, const T* c33, T Atmin, T Atmax,
                                 T ramp, const T* wh, const T* kx, const T* xm, const T* alpha, const T* dV11,
                                 const T* dV12, const T* dV13, const T* dV22, const T* dV23, const T* dV33, T* ax,
                                 T* ay, T* az, double* du, float Kcour)
    {
        T minDt = INFINITY;
#pragma omp parallel for schedule(static) reduction(min : minDt)
        for (uint32_t i = first; i < last; ++i)
        {
            size_t ni    = i - first;
            unsigned ncC = std::min(nc[ni] - 1, ngmax);
            T maxvsignal = 0;
            sph::momentumAndEnergyJLoop<avClean>(i, K, box, nb + ngmax * ni, ncC, x, y, z, vx, vy, vz, h, m, prho,
                                                 (const T*)nullptr, c, c11, c12, c13, c22, c23, c33, Atmin, Atmax, ramp,
                                                 wh, kx, xm, alpha, dV11, dV12, dV13, dV22, dV23, dV33, ax, ay, az, du,
                                                 &maxvsignal);
            T dt_i = sph::tsKCourant(maxvsignal, h[i], c[i], Kcour);
            minDt  = std::min(minDt, dt_i);
        }
        return minDt;
    }

    // sph/include/sph/ts_global.hpp:71-94
    static double rhoTimestep(uint32_t first, uint32_t last, const T* divv, double Krho)
    {
        T maxDivv = -INFINITY;
#pragma omp parallel for reduction(max : maxDivv)
        for (uint32_t i = first; i < last; ++i)
            maxDivv = std::max(divv[i], maxDivv);
        return Krho / std::abs(maxDivv);
    }

    // sph/include/sph/positions.hpp:89-138 (updatePositionsHost + updateTempHost with the decay_t fix)
    // Tv: type of v, x_m1, a (T); du_m1 type Tdm1 = T in reference ParticlesData (XM1Type/HydroType)
    static void positions(uint32_t first, uint32_t last, double dt, double dt_m1, const cstone::Box<double>& box,
                          double* x, double* y, double* z, T* vx, T* vy, T* vz, T* x_m1, T* y_m1, T* z_m1, const T* ax,
                          const T* ay, const T* az, double* temp, const double* du, T* du_m1, const T* h, T muiConst,
                          double gamma)
    {
        using cstone::Vec3;
        bool fbcX   = (box.boundaryX() == cstone::BoundaryType::fixed);
        bool fbcY   = (box.boundaryY() == cstone::BoundaryType::fixed);
        bool fbcZ   = (box.boundaryZ() == cstone::BoundaryType::fixed);
        bool anyFBC = fbcX || fbcY || fbcZ;
        auto fbcCheck = [](double coord, T hh, double top, double bottom, bool fbc)
        { return fbc && (std::abs(top - coord) < T(2) * hh || std::abs(bottom - coord) < T(2) * hh); };
        auto constCv = sph::idealGasCv(muiConst, gamma);
#pragma omp parallel for schedule(static)
        for (uint32_t i = first; i < last; ++i)
        {
            bool skip = false;
            if (anyFBC && vx[i] == T(0) && vy[i] == T(0) && vz[i] == T(0))
            {
                if (fbcCheck(x[i], h[i], box.xmax(), box.xmin(), fbcX) ||
                    fbcCheck(y[i], h[i], box.ymax(), box.ymin(), fbcY) ||
                    fbcCheck(z[i], h[i], box.zmax(), box.zmin(), fbcZ))
                {
                    skip = true;
                }
            }
            if (!skip)
            {
                // same expression sequence as positionUpdate (positions.hpp:77-88) with Vec3<double> operands,
                // as instantiated by updatePositionsHost (positions.hpp:111-115: Vec3<T=Tc> from mixed fields)
                Vec3<double> A{double(ax[i]), double(ay[i]), double(az[i])};
                Vec3<double> X{x[i], y[i], z[i]};
                Vec3<double> X_m1{double(x_m1[i]), double(y_m1[i]), double(z_m1[i])};
                auto [Xnp1, Vnp1, dXnp1] = sph::positionUpdate(dt, dt_m1, X, A, X_m1, box);
                x[i] = Xnp1[0]; y[i] = Xnp1[1]; z[i] = Xnp1[2];
                x_m1[i] = dXnp1[0]; y_m1[i] = dXnp1[1]; z_m1[i] = dXnp1[2];
                vx[i] = Vnp1[0]; vy[i] = Vnp1[1]; vz[i] = Vnp1[2];
            }
            auto cv    = constCv;
            auto u_old = cv * temp[i];
            double dui = du[i];
            double dum = double(du_m1[i]);
            double u_new = sph::energyUpdate(u_old, dt, dt_m1, dui, dum);
            temp[i]  = u_new / cv;
            du_m1[i] = du[i];
        }
    }


    // driftKernel (sph/include/sph/positions_gpu.cu:47-90) on an index list: the block-time-step propagator exists only
    // as CUDA kernels in the reference; the two functions below call the same host-callable building blocks
    // (sph::positionUpdate, sph::energyUpdate) in the kernels' order
    static void drift(const uint32_t* idx, uint32_t count, float dt, float dt_back, const float* dt_m1,
                      const uint8_t* rung, double* x, double* y, double* z, T* vx, T* vy, T* vz, const T* x_m1,
                      const T* y_m1, const T* z_m1, const T* ax, const T* ay, const T* az, double* temp,
                      const double* du, const T* du_m1, T muiConst, double gamma)
    {
        using cstone::Vec3;
        cstone::Box<double> noPbc(0, 1, cstone::BoundaryType::open);
        T constCv = sph::idealGasCv(muiConst, gamma);
        for (uint32_t k = 0; k < count; ++k)
        {
            uint32_t i       = idx[k];
            float dt_m1_rung = rung ? dt_m1[rung[i]] : dt_m1[0];
            Vec3<double> An{double(ax[i]), double(ay[i]), double(az[i])};
            Vec3<double> Xnback{x[i], y[i], z[i]};
            Vec3<double> dXn{double(x_m1[i]), double(y_m1[i]), double(z_m1[i])};
            auto [Xn_recov, ig0, ig1] = sph::positionUpdate(-dt_back, dt_m1_rung, Xnback, An, dXn, noPbc);
            auto [Xnp1, Vnp1, ig2]    = sph::positionUpdate(dt, dt_m1_rung, Xn_recov, An, dXn, noPbc);
            x[i] = Xnp1[0]; y[i] = Xnp1[1]; z[i] = Xnp1[2];
            vx[i] = Vnp1[0]; vy[i] = Vnp1[1]; vz[i] = Vnp1[2];
            T cv         = constCv;
            auto u_recov = sph::energyUpdate(double(temp[i] * cv), -dt_back, dt_m1_rung, du[i], double(du_m1[i]));
            temp[i]      = sph::energyUpdate(u_recov, dt, dt_m1_rung, du[i], double(du_m1[i])) / cv;
        }
    }

    // computePositionsKernel (positions_gpu.cu:118-161) on an index list
    static void positionsRung(const uint32_t* idx, uint32_t count, float dt, const float* dt_m1, const uint8_t* rung,
                              const cstone::Box<double>& box, double* x, double* y, double* z, T* vx, T* vy, T* vz,
                              T* x_m1, T* y_m1, T* z_m1, const T* ax, const T* ay, const T* az, double* temp,
                              const double* du, T* du_m1, const T* h, T muiConst, double gamma)
    {
        using cstone::Vec3;
        bool fbcX   = (box.boundaryX() == cstone::BoundaryType::fixed);
        bool fbcY   = (box.boundaryY() == cstone::BoundaryType::fixed);
        bool fbcZ   = (box.boundaryZ() == cstone::BoundaryType::fixed);
        bool anyFBC = fbcX || fbcY || fbcZ;
        T constCv   = sph::idealGasCv(muiConst, gamma);
        for (uint32_t k = 0; k < count; ++k)
        {
            uint32_t i = idx[k];
            if (anyFBC && vx[i] == T(0) && vy[i] == T(0) && vz[i] == T(0))
            {
                if (sph::fbcCheck(x[i], h[i], box.xmax(), box.xmin(), fbcX) ||
                    sph::fbcCheck(y[i], h[i], box.ymax(), box.ymin(), fbcY) ||
                    sph::fbcCheck(z[i], h[i], box.zmax(), box.zmin(), fbcZ))
                {
                    continue;
                }
            }
            float dt_m1_rung = rung ? dt_m1[rung[i]] : dt_m1[0];
            Vec3<double> A{double(ax[i]), double(ay[i]), double(az[i])};
            Vec3<double> X{x[i], y[i], z[i]};
            Vec3<double> X_m1{double(x_m1[i]), double(y_m1[i]), double(z_m1[i])};
            auto [Xnp1, Vnp1, dXnp1] = sph::positionUpdate(dt, dt_m1_rung, X, A, X_m1, box);
            x[i] = Xnp1[0]; y[i] = Xnp1[1]; z[i] = Xnp1[2];
            x_m1[i] = dXnp1[0]; y_m1[i] = dXnp1[1]; z_m1[i] = dXnp1[2];
            vx[i] = Vnp1[0]; vy[i] = Vnp1[1]; vz[i] = Vnp1[2];
            T cv       = constCv;
            auto u_old = temp[i] * cv;
            temp[i]    = sph::energyUpdate(double(u_old), dt, dt_m1_rung, du[i], double(du_m1[i])) / cv;
            du_m1[i]   = du[i];
        }
    }

    static void updateH(uint32_t first, uint32_t last, unsigned ng0, const unsigned* nc, T* h)
    {
#pragma omp parallel for schedule(static)
        for (uint32_t i = first; i < last; ++i)
            h[i] = sph::updateH(ng0, nc[i], h[i]);
    }
};

#define BOX auto box = makeBox(lim, bc)

#define EXPORT_HYDRO(SUF, T)                                                                                          \
    extern "C" void ref_find_neighbors_##SUF(const double* x, const double* y, const double* z, T* h, uint32_t first, \
                                             uint32_t last, const double* lim, const int* bc, const TreeArgs* t,     \
                                             unsigned ng0, unsigned ngmax, uint32_t* neighbors, unsigned* nc,        \
                                             int iterateH)                                                            \
    {                                                                                                                 \
        findNeighborsT<T>(x, y, z, h, first, last, lim, bc, t, ng0, ngmax, neighbors, nc, iterateH);                 \
    }                                                                                                                 \
    extern "C" void ref_tables_##SUF(double sincIndex, T* wh, T* whd, double* K)                                      \
    {                                                                                                                 \
        Hydro<T>::tables(sincIndex, wh, whd, K);                                                                      \
    }                                                                                                                 \
    extern "C" void ref_xmass_##SUF(uint32_t first, uint32_t last, unsigned ngmax, double K, const double* lim,      \
                                    const int* bc, const uint32_t* nb, const unsigned* nc, const double* x,          \
                                    const double* y, const double* z, const T* h, const T* m, const T* wh,           \
                                    const T* whd, T* xm)                                                              \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::xmass(first, last, ngmax, K, box, nb, nc, x, y, z, h, m, wh, whd, xm);                             \
    }                                                                                                                 \
    extern "C" void ref_ve_def_gradh_##SUF(uint32_t first, uint32_t last, unsigned ngmax, double K,                  \
                                           const double* lim, const int* bc, const uint32_t* nb, const unsigned* nc, \
                                           const double* x, const double* y, const double* z, const T* h,            \
                                           const T* m, const T* wh, const T* whd, const T* xm, T* kx, T* gradh)      \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::gradh(first, last, ngmax, K, box, nb, nc, x, y, z, h, m, wh, whd, xm, kx, gradh);                  \
    }                                                                                                                 \
    extern "C" void ref_eos_##SUF(uint32_t first, uint32_t last, const double* temp, const T* m, const T* kx,        \
                                  const T* xm, const T* gradh, double muiConst, double gamma, T* prho, T* c, T* rho, \
                                  T* p)                                                                               \
    {                                                                                                                 \
        Hydro<T>::eos(first, last, temp, m, kx, xm, gradh, T(muiConst), gamma, prho, c, rho, p);                     \
    }                                                                                                                 \
    extern "C" void ref_iad_divv_curlv_##SUF(                                                                         \
        uint32_t first, uint32_t last, unsigned ngmax, double K, const double* lim, const int* bc,                   \
        const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z, const T* vx,      \
        const T* vy, const T* vz, const T* h, const T* wh, const T* whd, const T* xm, const T* kx, T* c11, T* c12,   \
        T* c13, T* c22, T* c23, T* c33, T* divv, T* curlv, T* dV11, T* dV12, T* dV13, T* dV22, T* dV23, T* dV33)     \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::iadDivvCurlv(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, wh, whd, xm, kx, c11,     \
                               c12, c13, c22, c23, c33, divv, curlv, dV11, dV12, dV13, dV22, dV23, dV33);            \
    }                                                                                                                 \
    extern "C" void ref_av_switches_##SUF(                                                                            \
        uint32_t first, uint32_t last, unsigned ngmax, double K, const double* lim, const int* bc,                   \
        const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z, const T* vx,      \
        const T* vy, const T* vz, const T* h, const T* c, const T* c11, const T* c12, const T* c13, const T* c22,    \
        const T* c23, const T* c33, const T* wh, const T* whd, const T* kx, const T* xm, const T* divv,              \
        double minDt, double alphamin, double alphamax, double decay_constant, T* alpha)                             \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::avSwitches(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, c, c11, c12, c13, c22, c23, \
                             c33, wh, whd, kx, xm, divv, minDt, T(alphamin), T(alphamax), T(decay_constant), alpha); \
    }                                                                                                                 \
    extern "C" double ref_momentum_energy_##SUF(                                                                      \
        uint32_t first, uint32_t last, unsigned ngmax, double K, const double* lim, const int* bc,                   \
        const uint32_t* nb, const unsigned* nc, const double* x, const double* y, const double* z, const T* vx,      \
        const T* vy, const T* vz, const T* h, const T* m, const T* prho, const T* c, const T* c11, const T* c12,     \
        const T* c13, const T* c22, const T* c23, const T* c33, double Atmin, double Atmax, double ramp,             \
        const T* wh, const T* kx, const T* xm, const T* alpha, const T* dV11, const T* dV12, const T* dV13,          \
        const T* dV22, const T* dV23, const T* dV33, T* ax, T* ay, T* az, double* du, float Kcour, int avClean)      \
    {                                                                                                                 \
        BOX;                                                                                                          \
        if (avClean)                                                                                                  \
            return Hydro<T>::momentumEnergy<true>(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, m,     \
                                                  prho, c, c11, c12, c13, c22, c23, c33, T(Atmin), T(Atmax),         \
                                                  T(ramp), wh, kx, xm, alpha, dV11, dV12, dV13, dV22, dV23, dV33,    \
                                                  ax, ay, az, du, Kcour);                                            \
        return Hydro<T>::momentumEnergy<false>(first, last, ngmax, K, box, nb, nc, x, y, z, vx, vy, vz, h, m, prho,  \
                                               c, c11, c12, c13, c22, c23, c33, T(Atmin), T(Atmax), T(ramp), wh, kx, \
                                               xm, alpha, dV11, dV12, dV13, dV22, dV23, dV33, ax, ay, az, du,        \
                                               Kcour);                                                               \
    }                                                                                                                 \
    extern "C" double ref_rho_timestep_##SUF(uint32_t first, uint32_t last, const T* divv, double Krho)              \
    {                                                                                                                 \
        return Hydro<T>::rhoTimestep(first, last, divv, Krho);                                                       \
    }                                                                                                                 \
    extern "C" void ref_positions_##SUF(uint32_t first, uint32_t last, double dt, double dt_m1, const double* lim,   \
                                        const int* bc, double* x, double* y, double* z, T* vx, T* vy, T* vz,         \
                                        T* x_m1, T* y_m1, T* z_m1, const T* ax, const T* ay, const T* az,            \
                                        double* temp, const double* du, T* du_m1, const T* h, double muiConst,       \
                                        double gamma)                                                                 \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::positions(first, last, dt, dt_m1, box, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az, temp,    \
                            du, du_m1, h, T(muiConst), gamma);                                                       \
    }                                                                                                                 \
    extern "C" void ref_drift_##SUF(const uint32_t* idx, uint32_t count, float dt, float dt_back, const float* dt_m1, \
                                    const uint8_t* rung, double* x, double* y, double* z, T* vx, T* vy, T* vz,       \
                                    const T* x_m1, const T* y_m1, const T* z_m1, const T* ax, const T* ay,           \
                                    const T* az, double* temp, const double* du, const T* du_m1, double muiConst,    \
                                    double gamma)                                                                     \
    {                                                                                                                 \
        Hydro<T>::drift(idx, count, dt, dt_back, dt_m1, rung, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az,     \
                        temp, du, du_m1, T(muiConst), gamma);                                                        \
    }                                                                                                                 \
    extern "C" void ref_positions_rung_##SUF(const uint32_t* idx, uint32_t count, float dt, const float* dt_m1,      \
                                             const uint8_t* rung, const double* lim, const int* bc, double* x,       \
                                             double* y, double* z, T* vx, T* vy, T* vz, T* x_m1, T* y_m1, T* z_m1,   \
                                             const T* ax, const T* ay, const T* az, double* temp, const double* du,  \
                                             T* du_m1, const T* h, double muiConst, double gamma)                    \
    {                                                                                                                 \
        BOX;                                                                                                          \
        Hydro<T>::positionsRung(idx, count, dt, dt_m1, rung, box, x, y, z, vx, vy, vz, x_m1, y_m1, z_m1, ax, ay, az, \
                                temp, du, du_m1, h, T(muiConst), gamma);                                             \
    }                                                                                                                 \
    extern "C" void ref_update_h_##SUF(uint32_t first, uint32_t last, unsigned ng0, const unsigned* nc, T* h)        \
    {                                                                                                                 \
        Hydro<T>::updateH(first, last, ng0, nc, h);                                                                   \
    }

EXPORT_HYDRO(f64, double)
EXPORT_HYDRO(f32, float)

extern "C" double ref_sphynx_3D_k(double n) { return sph::sphynx_3D_k(n); }

// ---- observables: sphexa::localConservedQuantities (main/src/observables/conserved_quantities.hpp:48-107)
namespace
{
template<class T>
struct ObsDataset
{
    using AcceleratorType = cstone::CpuTag;
    std::vector<double> x, y, z, temp;
    std::vector<T> vx, vy, vz, m, u, mui;
    double muiConst, gamma;
};

template<class T>
void refConserved(const double* x, const double* y, const double* z, const T* vx, const T* vy, const T* vz, const T* m,
                  const double* temp, size_t n, double mui, double gamma, double* out)
{
    ObsDataset<T> d;
    d.x.assign(x, x + n); d.y.assign(y, y + n); d.z.assign(z, z + n); d.temp.assign(temp, temp + n);
    d.vx.assign(vx, vx + n); d.vy.assign(vy, vy + n); d.vz.assign(vz, vz + n); d.m.assign(m, m + n);
    d.muiConst = mui;
    d.gamma    = gamma;
    auto [eKin, eInt, linmom, angmom] = sphexa::localConservedQuantities(0, n, d);
    out[0] = eKin; out[1] = eInt; out[2] = 0.0;
    for (int k = 0; k < 3; ++k) { out[3 + k] = linmom[k]; out[6 + k] = angmom[k]; }
}
} // namespace

extern "C" void ref_conserved_f64(const double* x, const double* y, const double* z, const double* vx, const double* vy,
                                  const double* vz, const double* m, const double* temp, size_t n, double mui,
                                  double gamma, double* out)
{
    refConserved<double>(x, y, z, vx, vy, vz, m, temp, n, mui, gamma, out);
}
extern "C" void ref_conserved_f32(const double* x, const double* y, const double* z, const float* vx, const float* vy,
                                  const float* vz, const float* m, const double* temp, size_t n, double mui,
                                  double gamma, double* out)
{
    refConserved<float>(x, y, z, vx, vy, vz, m, temp, n, mui, gamma, out);
}

// ---- turbulence stirring: sph::TurbulenceData / updateNoise / computePhases / computeStirring (hydro_turb/*)
namespace
{
using RefTurb = sph::TurbulenceData<double, cstone::CpuTag>;
}

extern "C" void* ref_turb_create(double solWeight, double stMaxModes, double Lbox, double stEnergyPrefac,
                                 double stMachVelocity, double epsilon, double rngSeed, double stSpectForm,
                                 double powerLawExp, double anglesExp)
{
    std::map<std::string, double> c{{"solWeight", solWeight}, {"stMaxModes", stMaxModes}, {"Lbox", Lbox},
                                    {"stEnergyPrefac", stEnergyPrefac}, {"stMachVelocity", stMachVelocity},
                                    {"epsilon", epsilon}, {"rngSeed", rngSeed}, {"stSpectForm", stSpectForm},
                                    {"powerLawExp", powerLawExp}, {"anglesExp", anglesExp}};
    return new RefTurb(c, false);
}
extern "C" void ref_turb_destroy(void* t) { delete static_cast<RefTurb*>(t); }
extern "C" int ref_turb_num_modes(void* t) { return int(static_cast<RefTurb*>(t)->numModes); }
extern "C" void ref_turb_get(void* tp, double* scalars4, double* modes, double* amplitudes, double* phases,
                             double* phasesReal, double* phasesImag)
{
    auto* t     = static_cast<RefTurb*>(tp);
    scalars4[0] = t->variance; scalars4[1] = t->decayTime; scalars4[2] = t->solWeight; scalars4[3] = t->solWeightNorm;
    std::copy(t->modes.begin(), t->modes.end(), modes);
    std::copy(t->amplitudes.begin(), t->amplitudes.end(), amplitudes);
    std::copy(t->phases.begin(), t->phases.end(), phases);
    std::copy(t->phasesReal.begin(), t->phasesReal.end(), phasesReal);
    std::copy(t->phasesImag.begin(), t->phasesImag.end(), phasesImag);
}
//! @brief the CPU branch of driveTurbulence (driver.hpp:101-128) with a one-group view (the CPU
//!        computeGroups never sets numGroups)
template<class Ta>
static void refTurbDrive(void* tp, uint32_t first, uint32_t last, const double* x, const double* y, const double* z,
                         Ta* ax, Ta* ay, Ta* az, double dt)
{
    auto* t = static_cast<RefTurb*>(tp);
    sph::updateNoise(t->phases, t->variance, dt, t->decayTime, t->gen);
    sph::computePhases(t->numModes, t->numDim, t->phases, t->solWeight, t->modes, t->phasesReal, t->phasesImag);
    cstone::LocalIndex gs = first, ge = last;
    cstone::GroupView grp{first, last, 1, &gs, &ge};
    sph::computeStirring(grp, t->numDim, x, y, z, ax, ay, az, t->numModes, t->modes.data(), t->phasesReal.data(),
                         t->phasesImag.data(), t->amplitudes.data(), t->solWeightNorm);
}
//! @brief driveTurbulence (driver.hpp:101-128) with the caller's groups, as TurbVeBdtProp passes its active rungs
template<class Ta>
static void refTurbDriveGroups(void* tp, uint32_t numGroups, const uint32_t* starts, const uint32_t* ends,
                               const double* x, const double* y, const double* z, Ta* ax, Ta* ay, Ta* az, double dt)
{
    auto* t = static_cast<RefTurb*>(tp);
    sph::updateNoise(t->phases, t->variance, dt, t->decayTime, t->gen);
    sph::computePhases(t->numModes, t->numDim, t->phases, t->solWeight, t->modes, t->phasesReal, t->phasesImag);
    cstone::GroupView grp{0, 0, numGroups, starts, ends};
    sph::computeStirring(grp, t->numDim, x, y, z, ax, ay, az, t->numModes, t->modes.data(), t->phasesReal.data(),
                         t->phasesImag.data(), t->amplitudes.data(), t->solWeightNorm);
}
extern "C" void ref_turb_drive_groups_f64(void* t, uint32_t numGroups, const uint32_t* starts, const uint32_t* ends,
                                          const double* x, const double* y, const double* z, double* ax, double* ay,
                                          double* az, double dt)
{
    refTurbDriveGroups<double>(t, numGroups, starts, ends, x, y, z, ax, ay, az, dt);
}
extern "C" void ref_turb_drive_groups_f32(void* t, uint32_t numGroups, const uint32_t* starts, const uint32_t* ends,
                                          const double* x, const double* y, const double* z, float* ax, float* ay,
                                          float* az, double dt)
{
    refTurbDriveGroups<float>(t, numGroups, starts, ends, x, y, z, ax, ay, az, dt);
}
extern "C" void ref_turb_drive_f64(void* t, uint32_t first, uint32_t last, const double* x, const double* y,
                                   const double* z, double* ax, double* ay, double* az, double dt)
{
    refTurbDrive<double>(t, first, last, x, y, z, ax, ay, az, dt);
}
extern "C" void ref_turb_drive_f32(void* t, uint32_t first, uint32_t last, const double* x, const double* y,
                                   const double* z, float* ax, float* ay, float* az, double dt)
{
    refTurbDrive<float>(t, first, last, x, y, z, ax, ay, az, dt);
}

// ---------------------------------------------------------------------------------------------------------------
// multi-rank domain pieces (header-only on the host): assignment, halo discovery, layout
//   makeSfcAssignment / limitBoundaryShifts   domain/include/cstone/domain/domaindecomp.hpp:115-181
//   initialDomainSplits / computeSpanningTree  domaindecomp.hpp:243-256, tree/csarray.hpp:511-534
//   findHalos                                  traversal/collisions.hpp:78-105
//   computeNodeLayout                          domain/layout.hpp:150-164
#include "cstone/domain/domaindecomp.hpp"
#include "cstone/domain/layout.hpp"
#include "cstone/traversal/collisions.hpp"

extern "C" void ref_sfc_assignment(int numRanks, const unsigned* counts, const uint64_t* leaves, int numLeaves,
                                   const uint64_t* oldBounds, uint64_t* bounds, unsigned* rankCounts)
{
    std::vector<unsigned> cnt(counts, counts + numLeaves);
    // uniformBins runs exclusive_scan(begin, end + 1): one readable element past the counts
    cnt.reserve(numLeaves + 1);
    auto assignment = cstone::makeSfcAssignment<KeyType>(numRanks, cnt, leaves);
    if (oldBounds)
    {
        cstone::SfcAssignment<KeyType> old(numRanks);
        for (int r = 0; r <= numRanks; ++r)
            old.set(r, oldBounds[r], 0);
        cstone::limitBoundaryShifts<KeyType>(old, assignment, {leaves, size_t(numLeaves) + 1}, {cnt.data(), cnt.size()});
    }
    for (int r = 0; r <= numRanks; ++r)
        bounds[r] = assignment[r];
    for (int r = 0; r < numRanks; ++r)
        rankCounts[r] = assignment.totalCount(r);
}

extern "C" int ref_initial_global_tree(int numRanks, uint64_t* out, int capacity)
{
    unsigned level = cstone::log8ceil<KeyType>(100 * numRanks);
    auto splits    = cstone::initialDomainSplits<KeyType>(numRanks, level);
    auto tree      = cstone::computeSpanningTree<KeyType>(splits);
    if (int(tree.size()) > capacity) return -int(tree.size());
    std::copy(tree.begin(), tree.end(), out);
    return int(tree.size()) - 1;
}

extern "C" void ref_find_halos(const uint64_t* prefixes, const int* childOffsets, const int* internalToLeaf,
                               const uint64_t* leaves, const float* radii, const double* lim, const int* bc,
                               int firstNode, int lastNode, int* flags)
{
    auto box = makeBox(lim, bc);
    cstone::findHalos(prefixes, childOffsets, internalToLeaf, leaves, radii, box, firstNode, lastNode, flags);
}

//! the radii as Halos::discover prepares them on the host (halos/halos.hpp:176-188)
extern "C" void ref_halo_radii(const double* h, const unsigned* layout, int firstNode, int lastNode, float searchExt,
                               float* radii)
{
    for (int i = firstNode; i < lastNode; ++i)
    {
        radii[i] = 0.0f;
        if (layout[i + 1] > layout[i]) { radii[i] = *std::max_element(h + layout[i], h + layout[i + 1]) * 2 * searchExt; }
    }
}

extern "C" void ref_node_layout(const unsigned* counts, const int* haloFlags, int numLeaves, int firstAssigned,
                                int lastAssigned, unsigned* layout)
{
    cstone::computeNodeLayout({counts, size_t(numLeaves)}, {haloFlags, size_t(numLeaves)}, firstAssigned, lastAssigned,
                              {layout, size_t(numLeaves) + 1});
}

// ---------------------------------------------------------------------------------------------------------------
// TurbulenceMachRMS (main/src/observables/turbulence_mach_rms.hpp:38-83): localMachSquareSum / calculateMachRMS are
// templates on the dataset; they only touch vx, vy, vz, c and numParticlesGlobal, so a four-vector stand-in
// instantiates the reference's own code (no ParticlesData needed)
#include "observables/turbulence_mach_rms.hpp"

template<class T>
struct MachDataset
{
    using AcceleratorType = cstone::CpuTag;
    std::vector<T> vx, vy, vz, c;
    uint64_t numParticlesGlobal;
};

template<class T>
static void refMach(const T* vx, const T* vy, const T* vz, const T* c, size_t n, uint64_t numParticlesGlobal, double* out2)
{
    MachDataset<T> d{{vx, vx + n}, {vy, vy + n}, {vz, vz + n}, {c, c + n}, numParticlesGlobal};
    out2[0] = sphexa::localMachSquareSum(0, n, d);
    out2[1] = sphexa::calculateMachRMS(0, n, d, MPI_COMM_WORLD);
}

extern "C" void ref_mach_f64(const double* vx, const double* vy, const double* vz, const double* c, size_t n,
                             uint64_t numParticlesGlobal, double* out2)
{
    refMach(vx, vy, vz, c, n, numParticlesGlobal, out2);
}
extern "C" void ref_mach_f32(const float* vx, const float* vy, const float* vz, const float* c, size_t n,
                             uint64_t numParticlesGlobal, double* out2)
{
    refMach(vx, vy, vz, c, n, numParticlesGlobal, out2);
}
