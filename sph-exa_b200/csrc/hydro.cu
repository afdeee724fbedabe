// SPH-VE operators: XMass, VeDefGradh, EOS, IAD+divv/curlv, AV switches, momentum+energy,
// time-step reduction, position/energy/smoothing-length update.
//
// One thread per target particle; neighbor indices come from the slab lists written by
// findNeighborsKernel (k-th neighbor of lane l of slab g at nidx[(g*ngmax + k)*32 + l], so a warp
// reads 128 contiguous bytes per k).  Neighbor fields are gathered through the read-only path;
// lists are ascending in j and targets are SFC-consecutive, so a warp's gathers fall into a few
// cache lines.  Pair arithmetic restates the reference J-loops (sph/include/sph/hydro_ve/*_kern.hpp)
// expression by expression with T = double (f64 mode) or float (mixed mode, reference default).
#include "common.cuh"

namespace sphb200
{

constexpr int kHydroBlock = 128;
#ifndef SPHB200_MOMENTUM_MIN_BLOCKS
#define SPHB200_MOMENTUM_MIN_BLOCKS 4
#endif
constexpr int kMomentumMinBlocks = SPHB200_MOMENTUM_MIN_BLOCKS;

template<class T>
struct F
{
    const double *x, *y, *z;
    T *h, *m, *vx, *vy, *vz, *x_m1, *y_m1, *z_m1;
    double* temp;
    T *du_m1, *alpha;
    T *xm, *kx, *gradh, *prho, *c;
    T *c11, *c12, *c13, *c22, *c23, *c33;
    T *divv, *curlv, *ax, *ay, *az;
    double* du;
    uint32_t* nc;
    T *dV11, *dV12, *dV13, *dV22, *dV23, *dV33;
    const uint8_t* active; // GroupView scope (sphb200_set_scope): NULL = every target of [first,last)
    T* dtParticle;         // per-target Courant step of momentumEnergy, NULL = not stored
};

template<class T>
static F<T> typed(const sphb200_fields& f)
{
    F<T> r;
    r.active     = nullptr;
    r.dtParticle = nullptr;
    r.x = f.x; r.y = f.y; r.z = f.z;
#define CP(n) r.n = static_cast<T*>(f.n)
    CP(h); CP(m); CP(vx); CP(vy); CP(vz); CP(x_m1); CP(y_m1); CP(z_m1); CP(du_m1); CP(alpha);
    CP(xm); CP(kx); CP(gradh); CP(prho); CP(c); CP(c11); CP(c12); CP(c13); CP(c22); CP(c23); CP(c33);
    CP(divv); CP(curlv); CP(ax); CP(ay); CP(az); CP(dV11); CP(dV12); CP(dV13); CP(dV22); CP(dV23); CP(dV33);
#undef CP
    r.temp = f.temp; r.du = f.du; r.nc = f.nc;
    return r;
}

//! @brief field table restricted to the active targets of the context's scope
template<class T>
static F<T> scoped(const sphb200_ctx* ctx, const sphb200_fields& f)
{
    F<T> r       = typed<T>(f);
    r.active     = ctx->scopeActive;
    r.dtParticle = static_cast<T*>(ctx->scopeDtParticle);
    return r;
}

struct PbcBox
{
    double lx, ly, lz;
    bool px, py, pz;
    bool fx, fy, fz;
    double lim[6];
};

static PbcBox pbcBox(const sphb200_box& b)
{
    PbcBox p;
    p.lx = b.len[0]; p.ly = b.len[1]; p.lz = b.len[2];
    p.px = b.bc[0] == SPHB200_PERIODIC; p.py = b.bc[1] == SPHB200_PERIODIC; p.pz = b.bc[2] == SPHB200_PERIODIC;
    p.fx = b.bc[0] == SPHB200_FIXED; p.fy = b.bc[1] == SPHB200_FIXED; p.fz = b.bc[2] == SPHB200_FIXED;
    for (int k = 0; k < 6; ++k) p.lim[k] = b.lim[k];
    return p;
}

//! @brief true if the search sphere of a target reaches a periodic face.  Only then can a neighbor be a periodic image
//!        (the image lies outside the box, the sphere inside), i.e. only then does applyPBC ever shift anything; the
//!        1e-6 widening keeps the decision on the safe side of rounding and never changes a value.
template<class T>
__device__ __forceinline__ bool reachesPeriodicFace(const PbcBox& b, double x, double y, double z, T r)
{
    double rr = double(r) * 1.000001;
    return (b.px && (x - rr <= b.lim[0] || x + rr >= b.lim[1])) || (b.py && (y - rr <= b.lim[2] || y + rr >= b.lim[3])) ||
           (b.pz && (z - rr <= b.lim[4] || z + rr >= b.lim[5]));
}

//! @brief legacy minimum-image used inside the J-loops, cstone::applyPBC (sfc/box.hpp:234-255)
template<class T>
__device__ __forceinline__ void applyPBC(const PbcBox& b, T r, T& xx, T& yy, T& zz)
{
    if (b.px && xx > r) xx -= b.lx;
    else if (b.px && xx < -r) xx += b.lx;
    if (b.py && yy > r) yy -= b.ly;
    else if (b.py && yy < -r) yy += b.ly;
    if (b.pz && zz > r) zz -= b.lz;
    else if (b.pz && zz < -r) zz += b.lz;
}

// ---------------------------------------------------------------- explicit contractions
// Sums of products are written with explicit multiply-adds wherever the order of contraction is ambiguous, so that
// every variant of a kernel rounds alike.  The Courant signal velocity can be evaluated in either kernel (sphb200_av_switches_courant lets the step know its
// time step before momentumEnergy starts); explicit fused multiply-adds make both evaluations the same bits.
__device__ __forceinline__ double fmaT(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float fmaT(float a, float b, float c) { return __fmaf_rn(a, b, c); }

template<class T>
__device__ __forceinline__ T dot3(T ax, T ay, T az, T bx, T by, T bz)
{
    return fmaT(az, bz, fmaT(ay, by, ax * bx));
}

//! @brief a b + c d with the contraction spelled out.  Left to the compiler, which of the two products is rounded before
//!        the add depends on the code around the expression, and the L1-table and shared-memory-table variants of a
//!        kernel (same source) came out one bit apart in such sums.
template<class T>
__device__ __forceinline__ T sum2(T a, T b, T c, T d)
{
    return fmaT(c, d, a * b);
}

//! @brief vijsignal of momentum_energy_kern.hpp: 0.5 (c_i + c_j) - 2 w_ij
template<class T>
__device__ __forceinline__ T courantSignal(T ci, T cj, T wij)
{
    return fmaT(T(-2), wij, T(0.5) * (ci + cj));
}

/*! @brief v - idx * dx of sph::lt::lookup (table_lookup.hpp:21).  v is itself a product (dist / h), and whether the
 *         compiler fuses that product or idx * dx into the subtraction was seen to differ between two instances of the
 *         same source line; here v is always rounded first.
 */
template<class T>
__device__ __forceinline__ T tableFraction(T v, int idx)
{
    constexpr T dx = T(2.0) / (kTableSize - 1);
    return fmaT(-T(idx), dx, v);
}

//! @brief tabulated kernel with linear interpolation, sph::lt::lookup (sph/table_lookup.hpp:13-25)
template<class T>
__device__ __forceinline__ T lookup(const T* __restrict__ table, T v)
{
    constexpr int numIntervals = kTableSize - 1;
    constexpr T support        = 2.0;
    constexpr T dx             = support / numIntervals;
    constexpr T invDx          = T(1) / dx;

    int idx = v * invDx;
    if (idx >= numIntervals) return T(0);
    T t0         = __ldg(table + idx);
    T t1         = __ldg(table + idx + 1);
    T derivative = (t1 - t0) * invDx;
    return fmaT(derivative, tableFraction(v, idx), t0);
}

// ---------------------------------------------------------------- gather records
// The J-loops are bound by L1 gather throughput (one tag lookup per touched line, one data cycle per touched
// sector), not by DRAM.  Each operator therefore first packs the neighbor-side fields it reads into 32-byte
// (f64) / 16-byte (mixed) records with one streaming pass, so that a pair costs one 256-bit load per record instead
// of four 64-bit loads that each touch their own lines:
//   P {x, y, z, m} (XMass, gradh, momentum)   Pxm {x, y, z, xm} + VV {vx, vy, vz, xm/kx} (IAD)
//   Pc {x, y, z, c} + VV (AV switches)   R3 {c11, c12, c13, c22}
//   R5 {dV11, dV12, dV13, dV22}   R6 {dV23, dV33, -, -}   (avClean only)
//   momentum only: M1 {vx, vy, vz, 1/h}   M2 {xm, c, prho, rho}   M4 {c23, c33, alpha, m/rho}
// Quantities of the reference J-loops that depend on j alone (V_j = xm_j/kx_j, 1/h_j, rho_j = kx_j m_j/xm_j, m_j/rho_j)
// are evaluated once per particle by the pack kernels with the reference's expression, instead of once per pair:
// same bits, ~100x fewer divisions.
// The tabulated kernel is repacked the same way: entry idx holds {t[idx], t[idx+1]} (and for gradh the derivative
// table's pair as well), so an interpolation is one load.  Values are copies, arithmetic is unchanged.
template<class T>
struct alignas(32) PosRec
{
    double x, y, z;
    T m;
};

template<class T>
struct alignas(4 * sizeof(T)) Rec4
{
    T a, b, c, d;
};

template<class T>
struct alignas(2 * sizeof(T)) Rec2
{
    T a, b;
};

// __ldg has no overloads for user structs: go through the matching built-in vector types
__device__ __forceinline__ Rec4<double> ldRec(const Rec4<double>* __restrict__ p, uint32_t j)
{
    Rec4<double> r;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.a), "=d"(r.b), "=d"(r.c), "=d"(r.d) : "l"(p + j));
    return r;
}
__device__ __forceinline__ Rec4<float> ldRec(const Rec4<float>* __restrict__ p, uint32_t j)
{
    float4 v = __ldg(reinterpret_cast<const float4*>(p) + j);
    return {v.x, v.y, v.z, v.w};
}
__device__ __forceinline__ Rec2<double> ldRec(const Rec2<double>* __restrict__ p, uint32_t j)
{
    double2 v = __ldg(reinterpret_cast<const double2*>(p) + j);
    return {v.x, v.y};
}
__device__ __forceinline__ Rec2<float> ldRec(const Rec2<float>* __restrict__ p, uint32_t j)
{
    float2 v = __ldg(reinterpret_cast<const float2*>(p) + j);
    return {v.x, v.y};
}
__device__ __forceinline__ PosRec<double> ldRec(const PosRec<double>* __restrict__ p, uint32_t j)
{
    PosRec<double> r;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.m) : "l"(p + j));
    return r;
}
__device__ __forceinline__ PosRec<float> ldRec(const PosRec<float>* __restrict__ p, uint32_t j)
{
    PosRec<float> r;
    double w;
    asm("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(w) : "l"(p + j));
    r.m = __int_as_float(__double2loint(w));
    return r;
}
//! @brief first half of a Rec4 (for operators that need only {a, b})
template<class T>
__device__ __forceinline__ Rec2<T> ldHalf(const Rec4<T>* __restrict__ p, uint32_t j)
{
    return ldRec(reinterpret_cast<const Rec2<T>*>(p + j), 0);
}

template<class T>
__global__ void __launch_bounds__(256) packPosKernel(size_t n, const double* __restrict__ x,
                                                     const double* __restrict__ y, const double* __restrict__ z,
                                                     const T* __restrict__ m, PosRec<T>* __restrict__ out,
                                                     const uint8_t* __restrict__ mask)
{
    // mask (sphb200_op_scope::gatherMask): only the particles some active target gathers from are refreshed
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
    PosRec<T> r;
    r.x = x[i]; r.y = y[i]; r.z = z[i]; r.m = m[i];
    out[i] = r;
}

template<class T>
__global__ void __launch_bounds__(256) pack4Kernel(size_t n, const T* __restrict__ a, const T* __restrict__ b,
                                                   const T* __restrict__ c, const T* __restrict__ d,
                                                   Rec4<T>* __restrict__ out, const uint8_t* __restrict__ mask)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
    Rec4<T> r;
    r.a = a ? a[i] : T(0); r.b = b ? b[i] : T(0); r.c = c ? c[i] : T(0); r.d = d ? d[i] : T(0);
    out[i] = r;
}

//! @brief VV = {vx, vy, vz, V = xm/kx}; volume element as in iad_kern.hpp:75 / av_switches_kern.hpp:103
template<class T>
__global__ void __launch_bounds__(256) packVelVolKernel(size_t n, const T* __restrict__ vx, const T* __restrict__ vy,
                                                        const T* __restrict__ vz, const T* __restrict__ xm,
                                                        const T* __restrict__ kx, Rec4<T>* __restrict__ out,
                                                        const uint8_t* __restrict__ mask)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
    out[i] = {vx[i], vy[i], vz[i], xm[i] / kx[i]};
}

//! @brief momentum records M1, M2, M4 with the j-only terms of momentum_energy_kern.hpp:96-97,116-117,165
template<class T>
__global__ void __launch_bounds__(256) packMomentumKernel(size_t n, F<T> f, Rec4<T>* __restrict__ m1,
                                                          Rec4<T>* __restrict__ m2, Rec4<T>* __restrict__ m4,
                                                          const uint8_t* __restrict__ mask)
{
    size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || (mask && !mask[i])) return;
    T m = f.m[i], xm = f.xm[i];
    T rho = f.kx[i] * m / xm;
    m1[i] = {f.vx[i], f.vy[i], f.vz[i], T(1) / f.h[i]};
    m2[i] = {xm, f.c[i], f.prho[i], rho};
    m4[i] = {f.c23[i], f.c33[i], f.alpha[i], m / rho};
}

//! @brief out[idx] = {wh[idx], slope of wh on [idx, idx+1], whd[idx], slope of whd}; the slope is the `derivative` of
//!        sph::lt::lookup (table_lookup.hpp:21), evaluated here once per interval instead of once per pair
template<class T>
__global__ void packTableKernel(const T* __restrict__ wh, const T* __restrict__ whd, Rec2<T>* __restrict__ pairs,
                                Rec4<T>* __restrict__ quads)
{
    constexpr T dx    = T(2.0) / (kTableSize - 1);
    constexpr T invDx = T(1) / dx;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= kTableSize) return;
    int i1   = min(i + 1, kTableSize - 1);
    T slope  = (wh[i1] - wh[i]) * invDx;
    pairs[i] = {wh[i], slope};
    if (whd) quads[i] = {wh[i], slope, whd[i], (whd[i1] - whd[i]) * invDx};
}

//! @brief tabulated kernel with linear interpolation, sph::lt::lookup (sph/table_lookup.hpp:13-25)
template<class T>
__device__ __forceinline__ T lookup(const Rec2<T>* __restrict__ table, T v)
{
    constexpr int numIntervals = kTableSize - 1;
    constexpr T support        = 2.0;
    constexpr T dx             = support / numIntervals;
    constexpr T invDx          = T(1) / dx;

    int idx = v * invDx;
    if (idx >= numIntervals) return T(0);
    Rec2<T> t = ldRec(table, idx); // {t[idx], derivative}
    return fmaT(t.b, tableFraction(v, idx), t.a);
}

//! @brief both tables at the same abscissa (wh and whd), one load
template<class T>
__device__ __forceinline__ void lookup2(const Rec4<T>* __restrict__ table, T v, T& w, T& dw)
{
    constexpr int numIntervals = kTableSize - 1;
    constexpr T support        = 2.0;
    constexpr T dx             = support / numIntervals;
    constexpr T invDx          = T(1) / dx;

    int idx = v * invDx;
    w = dw = T(0);
    if (idx >= numIntervals) return;
    Rec4<T> t = ldRec(table, idx);
    T frac    = tableFraction(v, idx);
    w         = fmaT(t.b, frac, t.a);
    dw        = fmaT(t.d, frac, t.c);
}

//! @brief the packed {t, slope} table in global memory (read through L1)
template<class T>
struct GlobalTab
{
    const Rec2<T>* pairs;
    __device__ __forceinline__ T operator()(T v) const { return lookup(pairs, v); }
};

/*! @brief the plain table resident in shared memory (160 KB in f64): on jittered particle sets the pair abscissae of a
 *         warp are spread over the whole table, so a lookup through L1 costs one tag cycle per lane and mostly misses;
 *         the same 32 lanes read shared memory in a few wavefronts.  Same arithmetic as packTableKernel + lookup: the
 *         slope is the rounded product (t1 - t0) * invDx, then one multiply-add, hence the same bits.
 */
template<class T>
struct SharedTab
{
    const T* t;
    __device__ __forceinline__ T operator()(T v) const
    {
        constexpr int numIntervals = kTableSize - 1;
        constexpr T support        = 2.0;
        constexpr T dx             = support / numIntervals;
        constexpr T invDx          = T(1) / dx;

        int idx = v * invDx;
        if (idx >= numIntervals) return T(0);
        T t0    = t[idx];
        T slope = (t[idx + 1] - t0) * invDx;
        return fmaT(slope, tableFraction(v, idx), t0);
    }
};

constexpr int kTabChunks = 10; // bulk copies per table (16 KB / 8 KB each)

/*! @brief bring the kernel table into shared memory with bulk async copies (TMA, cp.async.bulk) that complete on an
 *         mbarrier; every thread of the CTA returns once the table is readable
 */
template<class T>
__device__ __forceinline__ const T* stageTable(const T* __restrict__ table)
{
    extern __shared__ __align__(128) unsigned char tabSmem[];
    __shared__ __align__(8) unsigned long long tabBar;
    constexpr uint32_t bytes = kTableSize * sizeof(T);
    constexpr uint32_t chunk = bytes / kTabChunks;
    static_assert(bytes % kTabChunks == 0 && chunk % 16 == 0, "bulk copies move multiples of 16 bytes");

    uint32_t bar = uint32_t(__cvta_generic_to_shared(&tabBar));
    uint32_t dst = uint32_t(__cvta_generic_to_shared(tabSmem));
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        const char* src = reinterpret_cast<const char*>(table);
#pragma unroll 1
        for (int c = 0; c < kTabChunks; ++c)
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst + c * chunk), "l"(src + size_t(c) * chunk), "r"(chunk), "r"(bar)
                         : "memory");
    }
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "TAB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
                 "@p bra TAB_DONE;\n"
                 "bra TAB_WAIT;\n"
                 "TAB_DONE:\n"
                 "}" ::"r"(bar)
                 : "memory");
    return reinterpret_cast<const T*>(tabSmem);
}

#define LD(p, j) __ldg((p) + (j))

struct NbList
{
    const uint32_t* base; // entry k at base[k * 32]
    unsigned count;
    // lists are streamed exactly once per operator: keep them out of L1 so the gathered records stay resident
    __device__ __forceinline__ uint32_t operator[](unsigned k) const { return __ldcs(base + size_t(k) * 32); }
};

__device__ __forceinline__ NbList nbList(const uint32_t* nidx, const uint32_t* nc, uint32_t first, uint32_t i,
                                         unsigned ngmax)
{
    uint32_t ni = i - first;
    NbList l;
    l.base  = nidx + (size_t(ni >> 5) * ngmax) * 32 + (ni & 31);
    l.count = min(nc[i] - 1, ngmax);
    return l;
}

//! @brief J-loop with the next index fetched one iteration ahead (hides the list-load latency behind the pair math)
#define FOR_NEIGHBORS(nb, j)                                                                                          \
    uint32_t j##_next = (nb).count ? (nb)[0] : 0u;                                                                    \
    for (unsigned pj = 0, j = j##_next; pj < (nb).count;                                                             \
         ++pj, j = j##_next)                                                                                          \
        if (j##_next = (pj + 1 < (nb).count) ? (nb)[pj + 1] : 0u, true)

// ---------------------------------------------------------------- XMass (xmass_kern.hpp:50-78)
#ifndef SPHB200_XMASS_MIN_BLOCKS
#define SPHB200_XMASS_MIN_BLOCKS 1
#endif
template<bool Pbc, class T, class Tab>
__device__ __forceinline__ void xmassTarget(uint32_t i, uint32_t first, unsigned ngmax, double K, const PbcBox& box,
                                            const uint32_t* __restrict__ nidx, const F<T>& f,
                                            const PosRec<T>* __restrict__ r0, Tab wh)
{
    NbList nb = nbList(nidx, f.nc, first, i, ngmax);

    PosRec<T> pi = r0[i];
    double xi = pi.x, yi = pi.y, zi = pi.z;
    T hi = f.h[i], mi = pi.m;
    T hInv  = 1.0 / hi;
    T h3Inv = hInv * hInv * hInv;

    T rho0i = mi;
    const bool pbcI = Pbc && reachesPeriodicFace(box, xi, yi, zi, T(2) * hi);
    FOR_NEIGHBORS(nb, j)
    {
        PosRec<T> pj_ = ldRec(r0, j);
        T xx = xi - pj_.x, yy = yi - pj_.y, zz = zi - pj_.z;
        if (pbcI) applyPBC(box, T(2) * hi, xx, yy, zz);
        T dist = sqrt(dot3(xx, yy, zz, xx, yy, zz));
        T vloc = dist * hInv;
        T w    = wh(vloc);
        rho0i += w * pj_.m;
    }
    f.xm[i] = T(mi / (rho0i * K * h3Inv));
}

template<bool Pbc, class T>
__global__ void __launch_bounds__(kHydroBlock, SPHB200_XMASS_MIN_BLOCKS)
    xmassKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box, const uint32_t* __restrict__ nidx,
                F<T> f, const PosRec<T>* __restrict__ r0, const Rec2<T>* __restrict__ wh)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last || (f.active && !f.active[i])) return;
    xmassTarget<Pbc, T>(i, first, ngmax, K, box, nidx, f, r0, GlobalTab<T>{wh});
}

// ---- persistent variants with the kernel table in shared memory ---------------------------------------------------
// One CTA per SM holds the whole table; its warps walk the 32-target groups of the neighbor lists with a grid stride
// (adjacent warps take adjacent groups, so the gathered records stay as local in L1 as with 128-target blocks).
// Which variant runs is decided per operator from the first timed launches of each (sphb200_ctx::tabTune): lattices,
// whose warps share a handful of table lines, are faster through L1; jittered sets are faster through shared memory.
#define FOR_TARGET_GROUPS(i, begin, last)                                                                             \
    for (uint32_t g__ = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), i = (begin) + g__ * 32 + (threadIdx.x & 31); \
         g__ < ((last) - (begin) + 31) / 32; g__ += gridDim.x * (blockDim.x >> 5), i = (begin) + g__ * 32 + (threadIdx.x & 31))

constexpr int kXmassSmemBlock = 1024;
template<bool Pbc, class T>
__global__ void __launch_bounds__(kXmassSmemBlock, 1)
    xmassSmemKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                    const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                    const T* __restrict__ table)
{
    SharedTab<T> wh{stageTable(table)};
    FOR_TARGET_GROUPS(i, first, last)
    {
        if (i >= last || (f.active && !f.active[i])) continue;
        xmassTarget<Pbc, T>(i, first, ngmax, K, box, nidx, f, r0, wh);
    }
}

// ---------------------------------------------------------------- VeDefGradh (ve_def_gradh_kern.hpp:43-90)
#ifndef SPHB200_GRADH_MIN_BLOCKS
#define SPHB200_GRADH_MIN_BLOCKS 1
#endif
template<bool Pbc, class T>
__global__ void __launch_bounds__(kHydroBlock, SPHB200_GRADH_MIN_BLOCKS)
    veDefGradhKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                     const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                     const Rec4<T>* __restrict__ whq)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last || (f.active && !f.active[i])) return;
    NbList nb = nbList(nidx, f.nc, first, i, ngmax);

    PosRec<T> pi = r0[i];
    double xi = pi.x, yi = pi.y, zi = pi.z;
    T hi = f.h[i], mi = pi.m, xmassi = f.xm[i];
    T hInv  = T(1) / hi;
    T h3Inv = hInv * hInv * hInv;

    T kxi      = xmassi;
    T whomegai = -T(3) * xmassi;
    T wrho0i   = -T(3) * mi;

    const bool pbcI = Pbc && reachesPeriodicFace(box, xi, yi, zi, T(2) * hi);
    FOR_NEIGHBORS(nb, j)
    {
        PosRec<T> pj_ = ldRec(r0, j);
        T xmassj = LD(f.xm, j);
        T xx = xi - pj_.x, yy = yi - pj_.y, zz = zi - pj_.z;
        if (pbcI) applyPBC(box, T(2) * hi, xx, yy, zz);
        T dist = sqrt(xx * xx + yy * yy + zz * zz);
        T vloc = dist * hInv;
        T w, dw;
        lookup2(whq, vloc, w, dw);
        T dterh = -(T(3) * w + vloc * dw);

        kxi += w * xmassj;
        whomegai += dterh * xmassj;
        wrho0i += dterh * pj_.m;
    }

    // compound assignments exactly as ve_def_gradh_kern.hpp:78-80 (right-hand side evaluated first, K is Tc)
    kxi *= K * h3Inv;
    whomegai *= K * h3Inv * hInv;
    wrho0i *= K * h3Inv * hInv;

    whomegai = whomegai * mi / xmassi + (kxi - K * xmassi * h3Inv) * wrho0i;
    T rhoi   = kxi * mi / xmassi;
    T dhdrho = -hi / (rhoi * T(3));

    f.kx[i]    = kxi;
    f.gradh[i] = T(1) - dhdrho * whomegai;
}

// ---------------------------------------------------------------- EOS (hydro_ve/eos.hpp:52-76, sph/eos.hpp:13-40)
template<class T>
__global__ void eosKernel(uint32_t first, uint32_t last, F<T> f, T muiConst, double gamma)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last) return;
    T rho = f.kx[i] * f.m[i] / f.xm[i];
    // idealGasCv returns in the type of mui (Tmass)
    const T R  = 8.317e7;
    T cv       = T(R / muiConst / (gamma - T(1)));
    double tmp = cv * f.temp[i] * (gamma - 1.0);
    double p   = rho * tmp;
    double c   = sqrt(tmp);
    f.prho[i]  = T(p / (f.kx[i] * f.m[i] * f.m[i] * f.gradh[i]));
    f.c[i]     = T(c);
}

// ---------------------------------------------------------------- IAD + divv/curlv (iad_kern.hpp:43-109, divv_curlv_kern.hpp:44-123)
// One sweep: besides tau_ab = sum_j r_a r_b V_j W the sweep accumulates S_ab = sum_j (v_ji,a xm_j) W r_b; the
// velocity gradient of divv_curlv_kern.hpp, dV_a[b] = sum_j (v_ji,a xm_j) * (-(C_i r)_b W), equals -(C_i S_a)_b
// because C_i does not depend on j.  Same sums in the same order, C_i applied after instead of inside the loop.
#ifndef SPHB200_IAD_MIN_BLOCKS
#define SPHB200_IAD_MIN_BLOCKS 6 // <= 80 registers: 24 warps/SM instead of 16 (measured 6.9 -> 5.75 ms)
#endif
template<bool Pbc, class T, class Tab>
__device__ __forceinline__ void iadTarget(uint32_t i, uint32_t first, unsigned ngmax, double K, const PbcBox& box,
                                          const uint32_t* __restrict__ nidx, const F<T>& f,
                                          const PosRec<T>* __restrict__ r0, const Rec4<T>* __restrict__ r1, Tab wh,
                                          bool doGradV)
{
    NbList nb = nbList(nidx, f.nc, first, i, ngmax);

    PosRec<T> pi = r0[i]; // {x, y, z, xm}
    Rec4<T> vi   = r1[i]; // {vx, vy, vz, V}
    double xi = pi.x, yi = pi.y, zi = pi.z;
    T vxi = vi.a, vyi = vi.b, vzi = vi.c, hi = f.h[i];
    T hiInv = T(1) / hi;

    T tau11 = 0.0, tau12 = 0.0, tau13 = 0.0, tau22 = 0.0, tau23 = 0.0, tau33 = 0.0;
    T Sx[3] = {0, 0, 0}, Sy[3] = {0, 0, 0}, Sz[3] = {0, 0, 0};
    const bool pbcI = Pbc && reachesPeriodicFace(box, xi, yi, zi, T(2) * hi);
    FOR_NEIGHBORS(nb, j)
    {
        PosRec<T> pj_ = ldRec(r0, j);
        Rec4<T> vj    = ldRec(r1, j);
        T xmassj = pj_.m, volj = vj.d; // 4th slots: xm_j and V_j = xm_j / kx_j
        T rx = xi - pj_.x, ry = yi - pj_.y, rz = zi - pj_.z;
        if (pbcI) applyPBC(box, T(2) * hi, rx, ry, rz);
        T dist   = sqrt(dot3(rx, ry, rz, rx, ry, rz));
        T vloc   = dist * hiInv;
        T w      = wh(vloc);
        T volj_w = volj * w;

        tau11 += rx * rx * volj_w;
        tau12 += rx * ry * volj_w;
        tau13 += rx * rz * volj_w;
        tau22 += ry * ry * volj_w;
        tau23 += ry * rz * volj_w;
        tau33 += rz * rz * volj_w;

        T wx = w * rx, wy = w * ry, wz = w * rz;
        T sx = (vj.a - vxi) * xmassj, sy = (vj.b - vyi) * xmassj, sz = (vj.c - vzi) * xmassj;
        Sx[0] += sx * wx; Sx[1] += sx * wy; Sx[2] += sx * wz;
        Sy[0] += sy * wx; Sy[1] += sy * wy; Sy[2] += sy * wz;
        Sz[0] += sz * wx; Sz[1] += sz * wy; Sz[2] += sz * wz;
    }

    auto getExp = [](T val) { return (val == T(0) ? 0 : ilogb(val)); };
    int tauExpSum = getExp(tau11) + getExp(tau12) + getExp(tau13) + getExp(tau22) + getExp(tau23) + getExp(tau33);
    T normalization = ldexp(T(1), -tauExpSum / 6);

    tau11 *= normalization;
    tau12 *= normalization;
    tau13 *= normalization;
    tau22 *= normalization;
    tau23 *= normalization;
    tau33 *= normalization;

    T det = fmaT(-(tau33 * tau12), tau12,
                 fmaT(-(tau22 * tau13), tau13,
                      fmaT(-(tau11 * tau23), tau23, fmaT(T(2) * tau12 * tau23, tau13, tau11 * tau22 * tau33))));
    T factor = normalization * (hi * hi * hi) / (det * K);

    T c11i = sum2(tau22, tau33, -tau23, tau23) * factor;
    T c12i = sum2(tau13, tau23, -tau33, tau12) * factor;
    T c13i = sum2(tau12, tau23, -tau22, tau13) * factor;
    T c22i = sum2(tau11, tau33, -tau13, tau13) * factor;
    T c23i = sum2(tau13, tau12, -tau11, tau23) * factor;
    T c33i = sum2(tau11, tau22, -tau12, tau12) * factor;
    f.c11[i] = c11i; f.c12[i] = c12i; f.c13[i] = c13i; f.c22[i] = c22i; f.c23[i] = c23i; f.c33[i] = c33i;

    // dV_a[b] = -(C_i S_a)_b
    T dVx[3], dVy[3], dVz[3];
    auto applyC = [&](const T* S, T* dV)
    {
        dV[0] = -dot3(c11i, c12i, c13i, S[0], S[1], S[2]);
        dV[1] = -dot3(c12i, c22i, c23i, S[0], S[1], S[2]);
        dV[2] = -dot3(c13i, c23i, c33i, S[0], S[1], S[2]);
    };
    applyC(Sx, dVx);
    applyC(Sy, dVy);
    applyC(Sz, dVz);

    T kxi      = f.kx[i];
    T hiInv3   = hiInv * hiInv * hiInv;
    T norm_kxi = K * hiInv3 / kxi;
    f.divv[i]  = norm_kxi * (dVx[0] + dVy[1] + dVz[2]);
    if (f.curlv)
    {
        T cx = dVz[1] - dVy[2], cy = dVx[2] - dVz[0], cz = dVy[0] - dVx[1];
        f.curlv[i] = norm_kxi * sqrt(dot3(cx, cy, cz, cx, cy, cz));
    }
    if (doGradV)
    {
        f.dV11[i] = norm_kxi * dVx[0];
        f.dV12[i] = norm_kxi * (dVx[1] + dVy[0]);
        f.dV13[i] = norm_kxi * (dVx[2] + dVz[0]);
        f.dV22[i] = norm_kxi * dVy[1];
        f.dV23[i] = norm_kxi * (dVy[2] + dVz[1]);
        f.dV33[i] = norm_kxi * dVz[2];
    }
}

template<bool Pbc, class T>
__global__ void __launch_bounds__(kHydroBlock, SPHB200_IAD_MIN_BLOCKS)
    iadDivvCurlvKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                       const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                       const Rec4<T>* __restrict__ r1, const Rec2<T>* __restrict__ wh, bool doGradV)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last || (f.active && !f.active[i])) return;
    iadTarget<Pbc, T>(i, first, ngmax, K, box, nidx, f, r0, r1, GlobalTab<T>{wh}, doGradV);
}

constexpr int kIadSmemBlock = SPHB200_IAD_MIN_BLOCKS * kHydroBlock; // same registers per thread as the L1 variant
template<bool Pbc, class T>
__global__ void __launch_bounds__(kIadSmemBlock, 1)
    iadDivvCurlvSmemKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                           const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                           const Rec4<T>* __restrict__ r1, const T* __restrict__ table, bool doGradV)
{
    SharedTab<T> wh{stageTable(table)};
    FOR_TARGET_GROUPS(i, first, last)
    {
        if (i >= last || (f.active && !f.active[i])) continue;
        iadTarget<Pbc, T>(i, first, ngmax, K, box, nidx, f, r0, r1, wh, doGradV);
    }
}

// ---------------------------------------------------------------- AV switches (av_switches_kern.hpp:44-137)
// graddivv_b = sum_j factor_j * (-(C_i r)_b K h^-3 W): C_i, K h^-3 are constant in j and applied after the sweep.
#ifndef SPHB200_AV_MIN_BLOCKS
#define SPHB200_AV_MIN_BLOCKS 7 // <= 72 registers: 28 warps/SM (measured 6.6 -> 6.2 ms)
#endif
template<bool Pbc, bool Courant, class T, class Tab>
__device__ __forceinline__ void avSwitchesBody(uint32_t i, uint32_t first, unsigned ngmax, double K, const PbcBox& box,
                                               const uint32_t* __restrict__ nidx, const F<T>& f,
                                               const PosRec<T>* __restrict__ r0, const Rec4<T>* __restrict__ r1,
                                               Tab wh, double dt, T alphamin, T alphamax, T decay_constant,
                                               T& maxvsignal_i)
{
    NbList nb = nbList(nidx, f.nc, first, i, ngmax);

    PosRec<T> pi = r0[i]; // {x, y, z, divv}
    Rec4<T> vi   = r1[i]; // {vx, vy, vz, V}
    double xi = pi.x, yi = pi.y, zi = pi.z;
    T vxi = vi.a, vyi = vi.b, vzi = vi.c, hi = f.h[i];
    T ci = f.c[i];

    T vijsignal_i = T(1.e-40) * ci;
    T hiInv  = T(1) / hi;
    T hiInv3 = hiInv * hiInv * hiInv;
    T divv_i = pi.m;

    T gx = 0.0, gy = 0.0, gz = 0.0;
    const bool pbcI = Pbc && reachesPeriodicFace(box, xi, yi, zi, T(2) * hi);
    FOR_NEIGHBORS(nb, j)
    {
        PosRec<T> pj_ = ldRec(r0, j);
        Rec4<T> vj    = ldRec(r1, j);
        T divv_j      = pj_.m; // 4th slot of the position record
        T rx = xi - pj_.x, ry = yi - pj_.y, rz = zi - pj_.z;
        if (pbcI) applyPBC(box, T(2) * hi, rx, ry, rz);
        T r2_  = dot3(rx, ry, rz, rx, ry, rz);
        T dist = sqrt(r2_);

        T vx_ij = vxi - vj.a;
        T vy_ij = vyi - vj.b;
        T vz_ij = vzi - vj.c;

        T rv           = dot3(rx, ry, rz, vx_ij, vy_ij, vz_ij);
        T vijsignal_ij = 0.0;
        if (Courant)
        {
            // the Courant signal of momentumEnergy needs c_j of every neighbor
            T cj = LD(f.c, j);
            if (rv < T(0)) { vijsignal_ij = ci + cj - T(3) * rv / dist; }
            T vs         = courantSignal(ci, cj, rv / dist);
            maxvsignal_i = (vs > maxvsignal_i) ? vs : maxvsignal_i;
        }
        else if (rv < T(0)) { vijsignal_ij = ci + LD(f.c, j) - T(3) * rv / dist; } // c_j only for approaching pairs
        vijsignal_i = max(vijsignal_i, vijsignal_ij);

        T v1 = dist * hiInv;
        T w  = wh(v1);

        T volj   = vj.d;
        T factor = volj * (divv_i - divv_j) * w;

        gx += factor * rx;
        gy += factor * ry;
        gz += factor * rz;
    }

    T c11i = f.c11[i], c12i = f.c12[i], c13i = f.c13[i], c22i = f.c22[i], c23i = f.c23[i], c33i = f.c33[i];
    T Kh3        = K * hiInv3;
    T graddivv_x = -dot3(c11i, c12i, c13i, gx, gy, gz) * Kh3;
    T graddivv_y = -dot3(c12i, c22i, c23i, gx, gy, gz) * Kh3;
    T graddivv_z = -dot3(c13i, c23i, c33i, gx, gy, gz) * Kh3;
    T graddivv   = sqrt(dot3(graddivv_x, graddivv_y, graddivv_z, graddivv_x, graddivv_y, graddivv_z));

    T alpha_i  = f.alpha[i];
    T alphaloc = 0.0;
    if (divv_i < T(0))
    {
        T a_const = hi * hi * graddivv;
        alphaloc  = alphamax * a_const / fmaT(T(0.05), ci, fmaT(hi, abs(divv_i), a_const));
    }

    if (alphaloc >= alpha_i) { alpha_i = alphaloc; }
    else
    {
        T decay    = hi / (decay_constant * vijsignal_i);
        T alphadot = 0.0;
        if (alphaloc >= alphamin) { alphadot = (alphaloc - alpha_i) / decay; }
        else { alphadot = (alphamin - alpha_i) / decay; }
        alpha_i += alphadot * dt;
    }
    f.alpha[i] = alpha_i;
}

template<bool Pbc, class T>
__global__ void __launch_bounds__(kHydroBlock, SPHB200_AV_MIN_BLOCKS)
    avSwitchesKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                     const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                     const Rec4<T>* __restrict__ r1, const Rec2<T>* __restrict__ wh, double dt, T alphamin, T alphamax,
                     T decay_constant)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last || (f.active && !f.active[i])) return;
    T unused = 0;
    avSwitchesBody<Pbc, false, T>(i, first, ngmax, K, box, nidx, f, r0, r1, GlobalTab<T>{wh}, dt, alphamin, alphamax,
                                  decay_constant, unused);
}

constexpr int kAvSmemBlock = SPHB200_AV_MIN_BLOCKS * kHydroBlock;
template<bool Pbc, class T>
__global__ void __launch_bounds__(kAvSmemBlock, 1)
    avSwitchesSmemKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                         const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                         const Rec4<T>* __restrict__ r1, const T* __restrict__ table, double dt, T alphamin, T alphamax,
                         T decay_constant)
{
    SharedTab<T> wh{stageTable(table)};
    FOR_TARGET_GROUPS(i, first, last)
    {
        if (i >= last || (f.active && !f.active[i])) continue;
        T unused = 0;
        avSwitchesBody<Pbc, false, T>(i, first, ngmax, K, box, nidx, f, r0, r1, wh, dt, alphamin, alphamax,
                                      decay_constant, unused);
    }
}

/*! @brief AVswitches that also evaluates the Courant time step of momentumEnergy (tsKCourant over the same
 *         vijsignal, momentum_energy_kern.hpp / sph/kernels.hpp:10-16): the step then knows dt before its last pair
 *         kernel starts and can integrate and ship particle ranges while that kernel is still running
 */
template<bool Pbc, class T>
__global__ void __launch_bounds__(kHydroBlock, 5)
    avSwitchesCourantKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                            const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                            const Rec4<T>* __restrict__ r1, const Rec2<T>* __restrict__ wh, double dt, T alphamin,
                            T alphamax, T decay_constant, float Kcour, unsigned long long* __restrict__ minDtKey)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    T dt_i     = INFINITY;
    if (i < last && (!f.active || f.active[i]))
    {
        T maxvsignal = 0;
        avSwitchesBody<Pbc, true, T>(i, first, ngmax, K, box, nidx, f, r0, r1, GlobalTab<T>{wh}, dt, alphamin, alphamax,
                                     decay_constant, maxvsignal);
        T v  = maxvsignal > T(0) ? maxvsignal : f.c[i];
        dt_i = T(Kcour * f.h[i] / v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        dt_i = min(dt_i, __shfl_xor_sync(0xffffffffu, dt_i, o));
    if ((threadIdx.x & 31) == 0) atomicMin(minDtKey, orderedKey(double(dt_i)));
}

constexpr int kAvCourantSmemBlock = 5 * kHydroBlock;
template<bool Pbc, class T>
__global__ void __launch_bounds__(kAvCourantSmemBlock, 1)
    avSwitchesCourantSmemKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                                const uint32_t* __restrict__ nidx, F<T> f, const PosRec<T>* __restrict__ r0,
                                const Rec4<T>* __restrict__ r1, const T* __restrict__ table, double dt, T alphamin,
                                T alphamax, T decay_constant, float Kcour, unsigned long long* __restrict__ minDtKey)
{
    SharedTab<T> wh{stageTable(table)};
    T dtMin = INFINITY;
    FOR_TARGET_GROUPS(i, first, last)
    {
        if (i >= last || (f.active && !f.active[i])) continue;
        T maxvsignal = 0;
        avSwitchesBody<Pbc, true, T>(i, first, ngmax, K, box, nidx, f, r0, r1, wh, dt, alphamin, alphamax,
                                     decay_constant, maxvsignal);
        T v   = maxvsignal > T(0) ? maxvsignal : f.c[i];
        dtMin = min(dtMin, T(Kcour * f.h[i] / v));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        dtMin = min(dtMin, __shfl_xor_sync(0xffffffffu, dtMin, o));
    if ((threadIdx.x & 31) == 0) atomicMin(minDtKey, orderedKey(double(dtMin)));
}

// ---------------------------------------------------------------- momentum + energy (momentum_energy_kern.hpp:43-222)
template<class T>
__device__ __forceinline__ T avRvCorrection(T rx, T ry, T rz, T eta_ab, T eta_crit, const T* gi, const T* gj)
{
    // symv: upper-triangular product as written in sph/kernels.hpp:86-94
    T s0 = dot3(gi[0], gi[1], gi[2], rx, ry, rz), s1 = sum2(gi[3], ry, gi[4], rz), s2 = gi[5] * rz;
    T dmy1 = dot3(rx, ry, rz, s0, s1, s2);
    T t0 = dot3(gj[0], gj[1], gj[2], rx, ry, rz), t1 = sum2(gj[3], ry, gj[4], rz), t2 = gj[5] * rz;
    T dmy2 = dot3(rx, ry, rz, t0, t1, t2);
    T dmy3 = T(1);
    if (eta_ab < eta_crit)
    {
        T etaDiff = T(5) * (eta_ab - eta_crit);
        dmy3      = exp(-etaDiff * etaDiff);
    }
    T A_ab   = (dmy2 != T(0)) ? dmy1 / dmy2 : T(0);
    T A_abp1 = T(1) + A_ab;
    T phi_ab = T(0.5) * dmy3 * max(T(0), min(T(1), T(4) * A_ab / (A_abp1 * A_abp1)));
    return -phi_ab * (dmy1 + dmy2);
}

/*! @brief (xm_j / xm_i)^s for a pair inside the Atwood ramp, 0 <= s <= 1
 *
 * Inside the ramp the two volume elements differ by a few ten per cent, so with u = (xm_j - xm_i) / (xm_j + xm_i),
 * log q = 2 atanh(u) and q^s = exp(s log q) are short series (|u| <= 1/4: 11 odd terms; |s log q| <= 0.52: 14 terms,
 * both below 1e-15 relative).  About half the instructions of log + exp, which matters because one lane in this
 * branch holds up its warp.  Pairs outside that range take the library functions.
 */
template<class T>
__device__ __forceinline__ T rampPower(T xmassi, T xmassj, T s)
{
    T u = (xmassj - xmassi) / (xmassj + xmassi);
    if (abs(u) > T(0.25)) return exp(s * log(xmassj / xmassi));
    T u2 = u * u;
    T p  = T(1.0 / 21);
#pragma unroll
    for (int k = 9; k >= 0; --k)
        p = p * u2 + T(1.0 / (2 * k + 1));
    T y = s * (T(2) * u * p); // s log q
    // sum_{n <= 14} y^n / n!
    constexpr double invFact[15] = {1.0, 1.0, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040, 1.0 / 40320,
                                    1.0 / 362880, 1.0 / 3628800, 1.0 / 39916800, 1.0 / 479001600, 1.0 / 6227020800.0,
                                    1.0 / 87178291200.0};
    T e = T(invFact[14]);
#pragma unroll
    for (int n = 13; n >= 0; --n)
        e = e * y + T(invFact[n]);
    return e;
}

template<class T>
struct MomRecs
{
    const PosRec<T>* r0;
    const Rec4<T>*r1, *r2, *r3, *r4, *r5, *r6;
};

//! @brief momentum and energy sums of target i; returns its Courant time step
template<bool Pbc, bool avClean, class T, class Tab>
__device__ __forceinline__ T momentumTarget(uint32_t i, uint32_t first, unsigned ngmax, double K, const PbcBox& box,
                                            const uint32_t* __restrict__ nidx, const F<T>& f, const MomRecs<T>& rec,
                                            Tab wh, T Atmin, T Atmax, T ramp, float Kcour)
{
    NbList nb = nbList(nidx, f.nc, first, i, ngmax);

    PosRec<T> pi = rec.r0[i];
    Rec4<T> vi   = rec.r1[i]; // {v, 1/h}
    Rec4<T> si   = rec.r2[i]; // {xm, c, prho, rho}
    double xi = pi.x, yi = pi.y, zi = pi.z;
    T vxi = vi.a, vyi = vi.b, vzi = vi.c, hi = f.h[i];
    T xmassi = si.a, ci = si.b, prhoi = si.c, rhoi = si.d;
    T alpha_i = f.alpha[i];

    T hiInv  = vi.d;
    T hiInv3 = hiInv * hiInv * hiInv;
    T rhoiInv = T(1) / rhoi;

    T maxvsignali = 0.0;
    T momentum_x = 0.0, momentum_y = 0.0, momentum_z = 0.0, energy = 0.0;
    T a_visc_energy = 0.0;

    T c11i = f.c11[i], c12i = f.c12[i], c13i = f.c13[i], c22i = f.c22[i], c23i = f.c23[i], c33i = f.c33[i];

    T gradV_i[6];
    if (avClean)
    {
        gradV_i[0] = f.dV11[i]; gradV_i[1] = f.dV12[i]; gradV_i[2] = f.dV13[i];
        gradV_i[3] = f.dV22[i]; gradV_i[4] = f.dV23[i]; gradV_i[5] = f.dV33[i];
    }

    T eta_crit = cbrt(T(32) * M_PI / T(3) / T(nb.count + 1));

    const bool pbcI = Pbc && reachesPeriodicFace(box, xi, yi, zi, T(2) * hi);
    FOR_NEIGHBORS(nb, j)
    {
        PosRec<T> pj_ = ldRec(rec.r0, j);
        Rec4<T> vj    = ldRec(rec.r1, j);
        Rec4<T> sj    = ldRec(rec.r2, j);
        Rec4<T> ca    = ldRec(rec.r3, j);
        Rec4<T> cb    = ldRec(rec.r4, j);

        T rx = xi - pj_.x, ry = yi - pj_.y, rz = zi - pj_.z;
        T vxj = vj.a, vyj = vj.b, vzj = vj.c;
        if (pbcI) applyPBC(box, T(2) * hi, rx, ry, rz);

        T r2_  = dot3(rx, ry, rz, rx, ry, rz);
        T dist = sqrt(r2_);

        T vx_ij = vxi - vxj, vy_ij = vyi - vyj, vz_ij = vzi - vzj;

        T hjInv = vj.d;

        T v1 = dist * hiInv;
        T v2 = dist * hjInv;

        T hjInv3 = hjInv * hjInv * hjInv;
        T Wi     = hiInv3 * wh(v1);
        T Wj     = hjInv3 * wh(v2);

        T termA1_i = -dot3(c11i, c12i, c13i, rx, ry, rz) * Wi;
        T termA2_i = -dot3(c12i, c22i, c23i, rx, ry, rz) * Wi;
        T termA3_i = -dot3(c13i, c23i, c33i, rx, ry, rz) * Wi;

        T c11j = ca.a, c12j = ca.b, c13j = ca.c, c22j = ca.d, c23j = cb.a, c33j = cb.b;

        T termA1_j = -dot3(c11j, c12j, c13j, rx, ry, rz) * Wj;
        T termA2_j = -dot3(c12j, c22j, c23j, rx, ry, rz) * Wj;
        T termA3_j = -dot3(c13j, c23j, c33j, rx, ry, rz) * Wj;

        T mj = pj_.m, cj = sj.b, xmassj = sj.a, rhoj = sj.d;

        T rv = dot3(rx, ry, rz, vx_ij, vy_ij, vz_ij);
        if (avClean)
        {
            Rec4<T> ga = ldRec(rec.r5, j);
            Rec2<T> gb = ldHalf(rec.r6, j);
            T gradV_j[6] = {ga.a, ga.b, ga.c, ga.d, gb.a, gb.b};
            rv += avRvCorrection(rx, ry, rz, min(v1, v2), eta_crit, gradV_i, gradV_j);
        }

        T wij = rv / dist;
        // artificial_viscosity (sph/kernels.hpp:70-83), beta = 2
        T viscosity_ij = 0.0;
        if (wij < 0.0)
        {
            T vij_signal = fmaT(T(-2.0), wij, (alpha_i + cb.c) / 4.0 * (ci + cj));
            viscosity_ij = -vij_signal * wij;
        }

        T vijsignal = courantSignal(ci, cj, wij);
        maxvsignali = (vijsignal > maxvsignali) ? vijsignal : maxvsignali;

        // Atwood = |rho_i - rho_j| / (rho_i + rho_j) selects the branch; the quotient itself is only needed inside
        // the ramp, so the two common branches compare against Atmin/Atmax * (rho_i + rho_j) without dividing
        // (the ramp is continuous at both ends, a tie decided the other way changes nothing beyond rounding)
        T a_mom, b_mom;
        T rhoDiff = abs(rhoi - rhoj), rhoSum = rhoi + rhoj;
        if (rhoDiff < Atmin * rhoSum)
        {
            a_mom = xmassi * xmassi;
            b_mom = xmassj * xmassj;
        }
        else if (rhoDiff > Atmax * rhoSum)
        {
            a_mom = xmassi * xmassj;
            b_mom = a_mom;
        }
        else
        {
            // xm_i^(2-s) xm_j^s = xm_i^2 q^s and xm_j^(2-s) xm_i^s = xm_j^2 / q^s with q = xm_j / xm_i: one
            // log + exp instead of four pow.  One lane in the ramp stalls its whole warp for the branch, and on
            // jittered data (turbulence, Noh) some lane is in it in most iterations: 4 pow made the kernel 2.9x
            // slower than on the Sedov lattice.  |s log q| << 1, so the result stays within a few ulp of pow's.
            T Atwood   = rhoDiff / rhoSum;
            T sigma_ij = ramp * (Atwood - Atmin);
            T qs       = rampPower(xmassi, xmassj, sigma_ij);
            a_mom      = xmassi * xmassi * qs;
            b_mom      = xmassj * xmassj / qs;
        }

        T a_visc   = mj * rhoiInv * viscosity_ij; // m_j / rho_i with the reciprocal taken once per target
        T b_visc   = cb.d * viscosity_ij; // (m_j / rho_j) * viscosity_ij
        T a_visc_x = T(0.5) * sum2(a_visc, termA1_i, b_visc, termA1_j);
        T a_visc_y = T(0.5) * sum2(a_visc, termA2_i, b_visc, termA2_j);
        T a_visc_z = T(0.5) * sum2(a_visc, termA3_i, b_visc, termA3_j);
        a_visc_energy += dot3(a_visc_x, a_visc_y, a_visc_z, vx_ij, vy_ij, vz_ij);

        energy += mj * a_mom * dot3(vx_ij, vy_ij, vz_ij, termA1_i, termA2_i, termA3_i);

        T momentum_i = mj * prhoi * a_mom;
        T momentum_j = mj * sj.c * b_mom;
        momentum_x += sum2(momentum_i, termA1_i, momentum_j, termA1_j) + a_visc_x;
        momentum_y += sum2(momentum_i, termA2_i, momentum_j, termA2_j) + a_visc_y;
        momentum_z += sum2(momentum_i, termA3_i, momentum_j, termA3_j) + a_visc_z;
    }

    a_visc_energy = max(T(0), a_visc_energy);
    f.du[i]       = K * fmaT(prhoi, energy, T(0.5) * a_visc_energy);
    f.ax[i]       = -K * momentum_x;
    f.ay[i]       = -K * momentum_y;
    f.az[i]       = -K * momentum_z;

    // tsKCourant (sph/kernels.hpp:10-16)
    T v  = maxvsignali > T(0) ? maxvsignali : ci;
    T dt_i = T(Kcour * hi / v);
    if (f.dtParticle) f.dtParticle[i] = dt_i;
    return dt_i;
}

template<bool Pbc, bool avClean, class T>
__global__ void __launch_bounds__(kHydroBlock, kMomentumMinBlocks)
    momentumEnergyKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                         const uint32_t* __restrict__ nidx, F<T> f, MomRecs<T> rec, const Rec2<T>* __restrict__ wh,
                         T Atmin, T Atmax, T ramp, float Kcour, unsigned long long* __restrict__ minDtKey,
                         uint32_t begin)
{
    // targets [begin, last); the lists are addressed relative to first (begin - first is a multiple of the block size)
    uint32_t i  = begin + blockIdx.x * blockDim.x + threadIdx.x;
    T dt_i      = INFINITY;
    if (i < last && (!f.active || f.active[i]))
        dt_i = momentumTarget<Pbc, avClean, T>(i, first, ngmax, K, box, nidx, f, rec, GlobalTab<T>{wh}, Atmin, Atmax,
                                               ramp, Kcour);

    // block minimum -> one ordered-key atomic per warp
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        dt_i = min(dt_i, __shfl_xor_sync(0xffffffffu, dt_i, o));
    if ((threadIdx.x & 31) == 0) atomicMin(minDtKey, orderedKey(double(dt_i)));
}

constexpr int kMomentumSmemBlock = 32 * kMomentumMinBlocks * (kHydroBlock / 32); // same registers per thread
template<bool Pbc, bool avClean, class T>
__global__ void __launch_bounds__(kMomentumSmemBlock, 1)
    momentumEnergySmemKernel(uint32_t first, uint32_t last, unsigned ngmax, double K, PbcBox box,
                             const uint32_t* __restrict__ nidx, F<T> f, MomRecs<T> rec, const T* __restrict__ table,
                             T Atmin, T Atmax, T ramp, float Kcour, unsigned long long* __restrict__ minDtKey,
                             uint32_t begin)
{
    SharedTab<T> wh{stageTable(table)};
    T dtMin = INFINITY;
    FOR_TARGET_GROUPS(i, begin, last)
    {
        if (i >= last || (f.active && !f.active[i])) continue;
        dtMin = min(dtMin, momentumTarget<Pbc, avClean, T>(i, first, ngmax, K, box, nidx, f, rec, wh, Atmin, Atmax, ramp,
                                                           Kcour));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        dtMin = min(dtMin, __shfl_xor_sync(0xffffffffu, dtMin, o));
    if ((threadIdx.x & 31) == 0) atomicMin(minDtKey, orderedKey(double(dtMin)));
}

// ---------------------------------------------------------------- timestep (ts_global.hpp:71-112)
template<class T>
__global__ void __launch_bounds__(256) maxDivvKernel(const T* __restrict__ divv, uint32_t first, uint32_t last,
                                                     unsigned long long* __restrict__ maxKey)
{
    T m = -INFINITY;
    for (size_t i = size_t(first) + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < last;
         i += size_t(gridDim.x) * blockDim.x)
        m = max(m, divv[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(maxKey, orderedKey(double(m)));
}

__global__ void initDtKeysKernel(unsigned long long* keys)
{
    keys[0] = ~0ull; // min courant
    keys[1] = 0ull;  // max divv
}

__global__ void initCourantKeyKernel(unsigned long long* keys) { keys[0] = ~0ull; }
__global__ void initDivvKeyKernel(unsigned long long* keys) { keys[1] = 0ull; }

//! @brief out[0]=minDtCourant, out[1]=Krho/|maxDivv|, out[2]=min(out0,out1,1.1*minDt)
__global__ void finalizeDtKernel(const unsigned long long* keys, double Krho, double maxDtIncrease, double minDt,
                                 double* out, int isFloat)
{
    double courant = fromOrderedKey(keys[0]);
    double maxDivv = fromOrderedKey(keys[1]);
    double rhoDt;
    if (isFloat) { rhoDt = Krho / double(fabsf(float(maxDivv))); }
    else { rhoDt = Krho / fabs(maxDivv); }
    out[0] = courant;
    out[1] = rhoDt;
    out[2] = fmin(fmin(courant, rhoDt), maxDtIncrease * minDt);
}

// ---------------------------------------------------------------- integrate (positions.hpp:54-138, update_h.hpp:12-24)
template<class T>
__device__ __forceinline__ T updateHT(unsigned ng0, unsigned nc, T h)
{
    const T c0  = T(1023.0);
    const T exp = T(1.0 / 10.0);
    return h * T(0.5) * pow(T(1) + c0 * T(ng0) / T(nc), exp);
}

__device__ __forceinline__ double putInBox1(double X, bool pbc, double lo, double hi, double l)
{
    if (pbc && X > hi) { X -= l; }
    else if (pbc && X < lo) { X += l; }
    return X;
}

template<class T>
__global__ void __launch_bounds__(256)
    positionsKernel(uint32_t first, uint32_t last, double dt, double dt_m1, PbcBox box, F<T> f, T muiConst,
                    double gamma, unsigned ng0, double* __restrict__ xw, double* __restrict__ yw,
                    double* __restrict__ zw)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last) return;

    T hi      = f.h[i];
    bool skip = false;
    bool anyFBC = box.fx || box.fy || box.fz;
    if (anyFBC && f.vx[i] == T(0) && f.vy[i] == T(0) && f.vz[i] == T(0))
    {
        auto fbc = [hi](double coord, double top, double bottom, bool on)
        { return on && (fabs(top - coord) < T(2) * hi || fabs(bottom - coord) < T(2) * hi); };
        skip = fbc(xw[i], box.lim[1], box.lim[0], box.fx) || fbc(yw[i], box.lim[3], box.lim[2], box.fy) ||
               fbc(zw[i], box.lim[5], box.lim[4], box.fz);
    }
    if (!skip)
    {
        double A[3]   = {double(f.ax[i]), double(f.ay[i]), double(f.az[i])};
        double X[3]   = {xw[i], yw[i], zw[i]};
        double dXn[3] = {double(f.x_m1[i]), double(f.y_m1[i]), double(f.z_m1[i])};
        double Xn[3], Vn1[3], dXn1[3];
        const double lo[3] = {box.lim[0], box.lim[2], box.lim[4]}, hiL[3] = {box.lim[1], box.lim[3], box.lim[5]};
        const double len[3] = {box.lx, box.ly, box.lz};
        const bool pb[3]    = {box.px, box.py, box.pz};
#pragma unroll
        for (int d = 0; d < 3; ++d)
        {
            // positionUpdate (positions.hpp:77-88)
            double Vnmhalf = dXn[d] * (1.0 / dt_m1);
            double Vn      = Vnmhalf + 0.5 * dt_m1 * A[d];
            Vn1[d]         = Vn + A[d] * dt;
            dXn1[d]        = (Vn + 0.5 * A[d] * fabs(dt)) * dt;
            Xn[d]          = putInBox1(X[d] + dXn1[d], pb[d], lo[d], hiL[d], len[d]);
        }
        xw[i] = Xn[0]; yw[i] = Xn[1]; zw[i] = Xn[2];
        f.x_m1[i] = T(dXn1[0]); f.y_m1[i] = T(dXn1[1]); f.z_m1[i] = T(dXn1[2]);
        f.vx[i] = T(Vn1[0]); f.vy[i] = T(Vn1[1]); f.vz[i] = T(Vn1[2]);
    }

    // updateTempHost with the intended value semantics (positions.hpp:124-138)
    const T R    = 8.317e7;
    T cv         = T(R / muiConst / (gamma - T(1)));
    double u_old = cv * f.temp[i];
    double du    = f.du[i];
    double dum1  = double(f.du_m1[i]);
    double u_new = u_old + du * dt + 0.5 * (du - dum1) / dt_m1 * fabs(dt) * dt;
    if (u_new < 0.) { u_new = u_old * exp(u_new * dt / u_old); }
    f.temp[i]  = u_new / cv;
    f.du_m1[i] = T(du);

    // updateSmoothingLength (update_h.hpp:12-24)
    f.h[i] = updateHT(ng0, f.nc[i], hi);
}

// ---------------------------------------------------------------- block time steps (positions_gpu.cu:47-180)
struct RungDt
{
    float v[4]; // Timestep::dt_m1 per rung (sph/include/sph/timestep.h:44-52)
};

//! @brief positionUpdate (positions.hpp:77-88) for one axis; open box when !pbc
__device__ __forceinline__ void positionUpdate1(double dt, double dt_m1, double X, double A, double dXn, bool pbc,
                                                double lo, double hi, double len, double& Xn1, double& Vn1,
                                                double& dXn1)
{
    double Vnmhalf = dXn * (1.0 / dt_m1);
    double Vn      = Vnmhalf + 0.5 * dt_m1 * A;
    Vn1            = Vn + A * dt;
    dXn1           = (Vn + 0.5 * A * fabs(dt)) * dt;
    Xn1            = putInBox1(X + dXn1, pbc, lo, hi, len);
}

//! @brief energyUpdate (positions.hpp:54-61)
__device__ __forceinline__ double energyUpdate1(double u_old, double dt, double dt_m1, double du, double du_m1)
{
    // No FMA contraction here: drifting back to an almost-zero energy (Noh: u = 1e-20) leaves a rounding residual
    // whose sign selects the branch below, so the operation sequence has to be the reference's, rounding by rounding
    double second = __dmul_rn(__dmul_rn(__dmul_rn(0.5, du - du_m1) / dt_m1, fabs(dt)), dt);
    double u_new  = __dadd_rn(__dadd_rn(u_old, __dmul_rn(du, dt)), second);
    if (u_new < 0.) { u_new = u_old * exp(u_new * dt / u_old); }
    return u_new;
}

/*! @brief driftKernel (positions_gpu.cu:47-90): drift back by dt_back to where the accelerations were evaluated, then
 *         forward by dt, without touching x_m1 / du_m1 and without folding into the box (the tree is not rebuilt)
 */
template<class T>
__global__ void __launch_bounds__(256)
    driftGroupsKernel(const uint32_t* __restrict__ grpStart, const uint32_t* __restrict__ grpEnd, uint32_t numGroups,
                      float dt, float dt_back, RungDt dt_m1, const uint8_t* __restrict__ rung, F<T> f, T muiConst,
                      double gamma, double* __restrict__ xw, double* __restrict__ yw, double* __restrict__ zw)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    uint32_t i = grpStart[g] + lane;
    if (i >= grpEnd[g]) return;

    float dt_m1_rung = rung ? dt_m1.v[rung[i]] : dt_m1.v[0];
    double A[3]      = {double(f.ax[i]), double(f.ay[i]), double(f.az[i])};
    double X[3]      = {xw[i], yw[i], zw[i]};
    double dXn[3]    = {double(f.x_m1[i]), double(f.y_m1[i]), double(f.z_m1[i])};
    double Xn[3], Vn[3];
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
        double Xrecov, ign0, ign1;
        positionUpdate1(-double(dt_back), dt_m1_rung, X[d], A[d], dXn[d], false, 0, 0, 0, Xrecov, ign0, ign1);
        positionUpdate1(dt, dt_m1_rung, Xrecov, A[d], dXn[d], false, 0, 0, 0, Xn[d], Vn[d], ign1);
    }
    xw[i] = Xn[0]; yw[i] = Xn[1]; zw[i] = Xn[2];
    f.vx[i] = T(Vn[0]); f.vy[i] = T(Vn[1]); f.vz[i] = T(Vn[2]);

    const T R      = 8.317e7;
    T cv           = T(R / muiConst / (gamma - T(1)));
    double du      = f.du[i];
    double dum1    = double(f.du_m1[i]);
    double u_recov = energyUpdate1(__dmul_rn(f.temp[i], double(cv)), -double(dt_back), dt_m1_rung, du, dum1);
    f.temp[i]      = energyUpdate1(u_recov, dt, dt_m1_rung, du, dum1) / cv;
}

//! @brief computePositionsKernel (positions_gpu.cu:118-161): advance the groups' particles by dt from their rung's dt_m1
template<class T>
__global__ void __launch_bounds__(256)
    positionsGroupsKernel(const uint32_t* __restrict__ grpStart, const uint32_t* __restrict__ grpEnd, uint32_t numGroups,
                          float dt, RungDt dt_m1, const uint8_t* __restrict__ rung, PbcBox box, F<T> f, T muiConst,
                          double gamma, double* __restrict__ xw, double* __restrict__ yw, double* __restrict__ zw)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    uint32_t i = grpStart[g] + lane;
    if (i >= grpEnd[g]) return;

    T hi        = f.h[i];
    bool anyFBC = box.fx || box.fy || box.fz;
    if (anyFBC && f.vx[i] == T(0) && f.vy[i] == T(0) && f.vz[i] == T(0))
    {
        auto fbc = [hi](double coord, double top, double bottom, bool on)
        { return on && (fabs(top - coord) < T(2) * hi || fabs(bottom - coord) < T(2) * hi); };
        if (fbc(xw[i], box.lim[1], box.lim[0], box.fx) || fbc(yw[i], box.lim[3], box.lim[2], box.fy) ||
            fbc(zw[i], box.lim[5], box.lim[4], box.fz))
            return;
    }

    float dt_m1_rung = rung ? dt_m1.v[rung[i]] : dt_m1.v[0];
    double A[3]      = {double(f.ax[i]), double(f.ay[i]), double(f.az[i])};
    double X[3]      = {xw[i], yw[i], zw[i]};
    double dXn[3]    = {double(f.x_m1[i]), double(f.y_m1[i]), double(f.z_m1[i])};
    double Xn[3], Vn1[3], dXn1[3];
    const double lo[3] = {box.lim[0], box.lim[2], box.lim[4]}, hiL[3] = {box.lim[1], box.lim[3], box.lim[5]};
    const double len[3] = {box.lx, box.ly, box.lz};
    const bool pb[3]    = {box.px, box.py, box.pz};
#pragma unroll
    for (int d = 0; d < 3; ++d)
        positionUpdate1(dt, dt_m1_rung, X[d], A[d], dXn[d], pb[d], lo[d], hiL[d], len[d], Xn[d], Vn1[d], dXn1[d]);
    xw[i] = Xn[0]; yw[i] = Xn[1]; zw[i] = Xn[2];
    f.x_m1[i] = T(dXn1[0]); f.y_m1[i] = T(dXn1[1]); f.z_m1[i] = T(dXn1[2]);
    f.vx[i] = T(Vn1[0]); f.vy[i] = T(Vn1[1]); f.vz[i] = T(Vn1[2]);

    const T R    = 8.317e7;
    T cv         = T(R / muiConst / (gamma - T(1)));
    double du    = f.du[i];
    double u_old = __dmul_rn(f.temp[i], double(cv));
    f.temp[i]    = energyUpdate1(u_old, dt, dt_m1_rung, du, double(f.du_m1[i])) / cv;
    f.du_m1[i]   = T(du);
}

//! @brief out[i] = updateH(ng0, nc[i], h[i]) for a range (update_h.hpp:12-24), h itself untouched
template<class T>
__global__ void __launch_bounds__(256)
    updateHRangeKernel(uint32_t first, uint32_t last, unsigned ng0, const uint32_t* __restrict__ nc,
                       const T* __restrict__ h, T* __restrict__ out)
{
    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < last) out[i] = updateHT(ng0, nc[i], h[i]);
}

//! @brief updateSmoothingLengthGpuKernel (update_h_gpu.cu:39-50)
template<class T>
__global__ void __launch_bounds__(256)
    updateHGroupsKernel(const uint32_t* __restrict__ grpStart, const uint32_t* __restrict__ grpEnd, uint32_t numGroups,
                        unsigned ng0, const uint32_t* __restrict__ nc, T* __restrict__ h)
{
    uint32_t lane = threadIdx.x & 31;
    uint32_t g    = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (g >= numGroups) return;
    uint32_t i = grpStart[g] + lane;
    if (i >= grpEnd[g]) return;
    h[i] = updateHT(ng0, nc[i], h[i]);
}

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

//! @brief instantiate KERNEL<Pbc, T> for the box at hand (open boxes skip the minimum-image test in the J-loops)
#define LAUNCH_PBC(ctx, box, KERNEL, grid, ...)                                                                       \
    do                                                                                                                \
    {                                                                                                                 \
        PbcBox pb__ = pbcBox(box);                                                                                    \
        if (pb__.px || pb__.py || pb__.pz) SPHB200_LAUNCH(ctx, (KERNEL<true, T>), grid, kHydroBlock, 0, __VA_ARGS__); \
        else SPHB200_LAUNCH(ctx, (KERNEL<false, T>), grid, kHydroBlock, 0, __VA_ARGS__);                              \
    } while (0)

#define DISPATCH(dtype, ...)                                                                                          \
    if ((dtype) == SPHB200_F64) { using T = double; __VA_ARGS__ }                                                     \
    else if ((dtype) == SPHB200_MIXED) { using T = float; __VA_ARGS__ }                                               \
    else throw std::runtime_error("unknown dtype")

static unsigned long long* dtKeys(sphb200_ctx* ctx) { return reinterpret_cast<unsigned long long*>(ctx->devDt + 4); }

// ---- gather-record management ------------------------------------------------------------------------------------
// Records live in grow-only context arenas.  A stand-alone operator call (tier 1) repacks what it reads on every call
// because the caller may have changed the arrays in between; a sphb200_sim (tier 2) sets ctx->packReuse and
// invalidates records itself whenever a member field is rewritten (see invalidate() calls below and in api.cu).
enum RecId { R_POS, R_PXM, R_PDIVV, R_VV, R_CA, R_GA, R_GB, R_MVEL, R_MSTATE, R_MCB, kNumRecs };
static_assert(kNumRecs <= sphb200_ctx::kMaxRecs, "record arenas");

static void beginOp(sphb200_ctx* ctx)
{
    if (!ctx->packReuse) { ctx->packValid = 0; ctx->tabValid = false; }
}
static void invalidate(sphb200_ctx* ctx, std::initializer_list<int> ids)
{
    for (int k : ids)
        ctx->packValid &= ~(1u << k);
}

template<class T>
static const void* ensureRec(sphb200_ctx* ctx, int k, const sphb200_fields& f)
{
    size_t n = f.n;
    if (n == 0) throw std::runtime_error("sphb200_fields.n (array length incl. halos) must be set");
    bool isPos   = k == R_POS || k == R_PXM || k == R_PDIVV;
    size_t bytes = n * (isPos ? sizeof(PosRec<T>) : sizeof(Rec4<T>));
    void* out    = ctx->rec[k].get(bytes);
    if (ctx->packValid & (1u << k)) return out;
    F<T> t   = typed<T>(f);
    int grid = gridFor(n, 256);
    auto p4  = [&](const T* a, const T* b, const T* c, const T* d)
    { SPHB200_LAUNCH(ctx, pack4Kernel<T>, grid, 256, 0, n, a, b, c, d, static_cast<Rec4<T>*>(out), ctx->scopeGather); };
    switch (k)
    {
        case R_POS:
        case R_PXM:
        case R_PDIVV:
        {
            const T* fourth = k == R_POS ? t.m : (k == R_PXM ? t.xm : t.divv);
            SPHB200_LAUNCH(ctx, packPosKernel<T>, grid, 256, 0, n, t.x, t.y, t.z, fourth, static_cast<PosRec<T>*>(out),
                           ctx->scopeGather);
            break;
        }
        case R_VV:
            SPHB200_LAUNCH(ctx, packVelVolKernel<T>, grid, 256, 0, n, t.vx, t.vy, t.vz, t.xm, t.kx,
                           static_cast<Rec4<T>*>(out), ctx->scopeGather);
            break;
        case R_CA: p4(t.c11, t.c12, t.c13, t.c22); break;
        case R_GA: p4(t.dV11, t.dV12, t.dV13, t.dV22); break;
        case R_GB: p4(t.dV23, t.dV33, nullptr, nullptr); break;
        default: throw std::runtime_error("ensureRec: momentum records are packed by packMomentum");
    }
    ctx->packValid |= 1u << k;
    return out;
}

template<class T>
static void ensureTables(sphb200_ctx* ctx, const sphb200_hydro_const* hc, const Rec2<T>*& pairs, const Rec4<T>*& quads)
{
    auto* p = static_cast<Rec2<T>*>(ctx->tabPairs.get(kTableSize * sizeof(Rec2<T>)));
    auto* q = static_cast<Rec4<T>*>(ctx->tabQuads.get(kTableSize * sizeof(Rec4<T>)));
    bool hit = ctx->tabValid && ctx->tabWh == hc->wh && ctx->tabWhd == hc->whd && ctx->tabElem == int(sizeof(T));
    if (!hit)
    {
        if (!hc->wh) throw std::runtime_error("hydro_const.wh must point to the device kernel table");
        SPHB200_LAUNCH(ctx, packTableKernel<T>, gridFor(kTableSize, 256), 256, 0, static_cast<const T*>(hc->wh),
                       static_cast<const T*>(hc->whd), p, q);
        ctx->tabValid = true; ctx->tabWh = hc->wh; ctx->tabWhd = hc->whd; ctx->tabElem = int(sizeof(T));
    }
    pairs = p;
    quads = q;
}

// ---- which table variant runs ------------------------------------------------------------------------------------
// Both variants of an operator produce the same bits, so the choice is a matter of speed only and is made from
// measurements on the data at hand: the first launch of an operator runs (and times, with events on the context's
// stream) the L1 variant, the second the shared-memory variant, and the faster one per target is kept for the next
// kRetuneCalls launches, after which both are timed again (a lattice turns into a disordered set as a run evolves).
// Reading a timing back waits for a launch made one call earlier, which has long finished in a stepping simulation.
enum TabOp { TAB_XMASS, TAB_IAD, TAB_AV, TAB_AVC, TAB_MOMENTUM, TAB_MOMENTUM_RANGE, kNumTabOps };
static_assert(kNumTabOps <= sphb200_ctx::kMaxTabOps, "table tuners");
constexpr int kRetuneCalls       = 256;
constexpr size_t kSmemMinTargets = size_t(kNumSms) * 1024 * 2; // below two waves of a persistent grid: L1 variant

template<class T, class Kernel>
static void allowTableSmem(Kernel kernel)
{
    SPHB200_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kTableSize * sizeof(T))));
}

static int forcedTabVariant()
{
    static int v = []
    {
        const char* e = std::getenv("SPHB200_TABLE");
        if (!e) return -1;
        return std::string(e) == "smem" ? 1 : (std::string(e) == "l1" ? 0 : -1);
    }();
    return v;
}

//! @brief true if the launch that follows shall use the shared-memory table; pair with tabLaunched() after the launch
static bool useSmemTable(sphb200_ctx* ctx, const sphb200_hydro_const* hc, int op, size_t numTargets)
{
    if (reinterpret_cast<uintptr_t>(hc->wh) % 16 != 0) return false; // bulk copies need a 16-byte aligned source
    int forced = ctx->tabForce >= 0 ? ctx->tabForce : forcedTabVariant();
    if (forced >= 0) return forced == 1;
    if (numTargets < kSmemMinTargets) return false;
    // launches under an active-target mask (block time steps, partial substeps) are sparse and short: their timings
    // say nothing about the full launches of the same size, and they are not worth a persistent grid
    if (ctx->scopeActive) return false;

    auto& t = ctx->tabTune[op];
    if (t.pending >= 0)
    {
        float ms   = 0;
        bool valid = t.closed && cudaEventSynchronize(t.ev[1]) == cudaSuccess &&
                     cudaEventElapsedTime(&ms, t.ev[0], t.ev[1]) == cudaSuccess;
        cudaGetLastError();
        if (valid) t.perTarget[t.pending] = double(ms) / double(t.pendingTargets);
        else t.phase = 0; // a launch failed in between: measure again
        if (valid && t.pending == 1)
        {
            t.smem         = t.perTarget[1] < 0.97 * t.perTarget[0];
            t.phase        = 2;
            t.calls        = 0;
            t.tunedTargets = t.pendingTargets;
        }
        t.pending = -1;
    }
    if (t.phase == 2)
    {
        // another launch size is another workload (a context can serve several simulations)
        size_t diff = numTargets > t.tunedTargets ? numTargets - t.tunedTargets : t.tunedTargets - numTargets;
        if (++t.calls < kRetuneCalls && diff <= t.tunedTargets / 64) return t.smem;
        t.phase = 0;
    }
    for (auto& e : t.ev)
        if (!e) SPHB200_CUDA(cudaEventCreate(&e));
    t.pending        = t.phase;
    t.pendingTargets = numTargets;
    t.closed         = false;
    t.phase++;
    SPHB200_CUDA(cudaEventRecord(t.ev[0], ctx->stream));
    return t.pending == 1;
}

static void tabLaunched(sphb200_ctx* ctx, int op)
{
    auto& t = ctx->tabTune[op];
    if (t.pending < 0 || t.closed) return;
    SPHB200_CUDA(cudaEventRecord(t.ev[1], ctx->stream));
    t.closed = true;
}

//! @brief KERNEL<Pbc, T> as one persistent CTA per SM with the table in dynamic shared memory
#define LAUNCH_PBC_SMEM(ctx, box, KERNEL, block, numTargets, ...)                                                     \
    do                                                                                                                \
    {                                                                                                                 \
        PbcBox pb__ = pbcBox(box);                                                                                    \
        int grid__  = persistentGrid(numTargets, block, 1);                                                           \
        if (pb__.px || pb__.py || pb__.pz)                                                                            \
        {                                                                                                             \
            allowTableSmem<T>(KERNEL<true, T>);                                                                       \
            SPHB200_LAUNCH(ctx, (KERNEL<true, T>), grid__, block, kTableSize * sizeof(T), __VA_ARGS__);               \
        }                                                                                                             \
        else                                                                                                          \
        {                                                                                                             \
            allowTableSmem<T>(KERNEL<false, T>);                                                                      \
            SPHB200_LAUNCH(ctx, (KERNEL<false, T>), grid__, block, kTableSize * sizeof(T), __VA_ARGS__);              \
        }                                                                                                             \
    } while (0)

//! @brief the three momentum-only records, always repacked (momentum is their only reader, once per step)
template<class T>
static void packMomentum(sphb200_ctx* ctx, const sphb200_fields& f, MomRecs<T>& rec)
{
    size_t n = f.n;
    if (n == 0) throw std::runtime_error("sphb200_fields.n (array length incl. halos) must be set");
    auto* m1 = static_cast<Rec4<T>*>(ctx->rec[R_MVEL].get(n * sizeof(Rec4<T>)));
    auto* m2 = static_cast<Rec4<T>*>(ctx->rec[R_MSTATE].get(n * sizeof(Rec4<T>)));
    auto* m4 = static_cast<Rec4<T>*>(ctx->rec[R_MCB].get(n * sizeof(Rec4<T>)));
    SPHB200_LAUNCH(ctx, packMomentumKernel<T>, gridFor(n, 256), 256, 0, n, typed<T>(f), m1, m2, m4, ctx->scopeGather);
    rec.r1 = m1;
    rec.r2 = m2;
    rec.r4 = m4;
}

/*! @brief momentumEnergy for the targets [begin, end) of [first, last)
 *
 * pack: (re)build the gather records from the current field arrays.  A caller that integrates particle ranges between
 * launches (sphb200_momentum_energy_range) packs with the first range only: later ranges must still see the
 * pre-integration state of their neighbors, which is what the records hold.
 */
template<class T>
static void launchMomentum(sphb200_ctx* ctx, uint32_t first, uint32_t begin, uint32_t end, bool pack,
                           const sphb200_box* box, const sphb200_hydro_const* hc, const uint32_t* neighbors,
                           const sphb200_fields* f, const Rec2<T>* wh, int tabOp)
{
    MomRecs<T> rec;
    size_t n = f->n;
    if (pack)
    {
        rec.r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_POS, *f));
        rec.r3 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_CA, *f));
        packMomentum<T>(ctx, *f, rec);
    }
    else
    {
        rec.r0 = static_cast<const PosRec<T>*>(ctx->rec[R_POS].get(n * sizeof(PosRec<T>)));
        rec.r3 = static_cast<const Rec4<T>*>(ctx->rec[R_CA].get(n * sizeof(Rec4<T>)));
        rec.r1 = static_cast<const Rec4<T>*>(ctx->rec[R_MVEL].get(n * sizeof(Rec4<T>)));
        rec.r2 = static_cast<const Rec4<T>*>(ctx->rec[R_MSTATE].get(n * sizeof(Rec4<T>)));
        rec.r4 = static_cast<const Rec4<T>*>(ctx->rec[R_MCB].get(n * sizeof(Rec4<T>)));
    }
    rec.r5 = rec.r6 = nullptr;
    if (hc->avClean)
    {
        if (!pack) throw std::runtime_error("momentum_energy_range: avClean is not supported with unpacked ranges");
        rec.r5 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_GA, *f));
        rec.r6 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_GB, *f));
    }
    PbcBox pb = pbcBox(*box);
    bool pbc  = pb.px || pb.py || pb.pz;
    int grid  = gridFor(end - begin, kHydroBlock);
    auto go   = [&](auto kernel)
    {
        SPHB200_LAUNCH(ctx, kernel, grid, kHydroBlock, 0, first, end, hc->ngmax, hc->K, pb, neighbors, scoped<T>(ctx, *f),
                       rec, wh, T(hc->Atmin), T(hc->Atmax), T(hc->ramp), float(hc->Kcour), dtKeys(ctx), begin);
    };
    auto goSmem = [&](auto kernel)
    {
        allowTableSmem<T>(kernel);
        SPHB200_LAUNCH(ctx, kernel, persistentGrid(end - begin, kMomentumSmemBlock, 1), kMomentumSmemBlock,
                       kTableSize * sizeof(T), first, end, hc->ngmax, hc->K, pb, neighbors, scoped<T>(ctx, *f), rec,
                       static_cast<const T*>(hc->wh), T(hc->Atmin), T(hc->Atmax), T(hc->ramp), float(hc->Kcour),
                       dtKeys(ctx), begin);
    };
    if (useSmemTable(ctx, hc, tabOp, end - begin))
    {
        if (pbc && hc->avClean) goSmem(momentumEnergySmemKernel<true, true, T>);
        else if (pbc) goSmem(momentumEnergySmemKernel<true, false, T>);
        else if (hc->avClean) goSmem(momentumEnergySmemKernel<false, true, T>);
        else goSmem(momentumEnergySmemKernel<false, false, T>);
    }
    else if (pbc && hc->avClean) go(momentumEnergyKernel<true, true, T>);
    else if (pbc) go(momentumEnergyKernel<true, false, T>);
    else if (hc->avClean) go(momentumEnergyKernel<false, true, T>);
    else go(momentumEnergyKernel<false, false, T>);
    tabLaunched(ctx, tabOp);
}

extern "C" void sphb200_invalidate_records(sphb200_ctx* ctx) { ctx->packValid = 0; }

extern "C" int sphb200_set_table_variant(sphb200_ctx* ctx, int variant)
{
    SPHB200_TRY
    if (variant < -1 || variant > 1) throw std::runtime_error("set_table_variant: -1 (timed), 0 (L1) or 1 (shared memory)");
    ctx->tabForce = variant;
    SPHB200_CATCH
}

extern "C" int sphb200_xmass(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, const sphb200_box* box,
                             const sphb200_hydro_const* hc, const uint32_t* neighbors, const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    beginOp(ctx);
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             auto* r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_POS, *f));
             if (useSmemTable(ctx, hc, TAB_XMASS, last - first))
                 LAUNCH_PBC_SMEM(ctx, *box, xmassSmemKernel, kXmassSmemBlock, last - first, first, last, hc->ngmax,
                                 hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, static_cast<const T*>(hc->wh));
             else
                 LAUNCH_PBC(ctx, *box, xmassKernel, gridFor(last - first, kHydroBlock), first, last, hc->ngmax, hc->K,
                            pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, wh);
             tabLaunched(ctx, TAB_XMASS););
    invalidate(ctx, {R_PXM, R_VV}); // xm rewritten
    SPHB200_CATCH
}

extern "C" int sphb200_ve_def_gradh(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, const sphb200_box* box,
                                    const sphb200_hydro_const* hc, const uint32_t* neighbors, const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    beginOp(ctx);
    if (!hc->whd) throw std::runtime_error("ve_def_gradh: hydro_const.whd must be set");
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             auto* r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_POS, *f));
             LAUNCH_PBC(ctx, *box, veDefGradhKernel, gridFor(last - first, kHydroBlock), first, last, hc->ngmax, hc->K,
                        pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, whq););
    invalidate(ctx, {R_VV}); // kx rewritten
    SPHB200_CATCH
}

extern "C" int sphb200_eos(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, const sphb200_hydro_const* hc,
                           const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, eosKernel<T>, gridFor(last - first, 256), 256, 0, first, last, typed<T>(*f),
                                   T(hc->muiConst), hc->gamma););
    SPHB200_CATCH
}

extern "C" int sphb200_iad_divv_curlv(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                      const sphb200_box* box, const sphb200_hydro_const* hc, const uint32_t* neighbors,
                                      const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    beginOp(ctx);
    bool doGradV = f->dV11 != nullptr;
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             auto* r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_PXM, *f));
             auto* r1 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_VV, *f));
             if (useSmemTable(ctx, hc, TAB_IAD, last - first))
                 LAUNCH_PBC_SMEM(ctx, *box, iadDivvCurlvSmemKernel, kIadSmemBlock, last - first, first, last, hc->ngmax,
                                 hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1,
                                 static_cast<const T*>(hc->wh), doGradV);
             else
                 LAUNCH_PBC(ctx, *box, iadDivvCurlvKernel, gridFor(last - first, kHydroBlock), first, last, hc->ngmax,
                            hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1, wh, doGradV);
             tabLaunched(ctx, TAB_IAD););
    invalidate(ctx, {R_PDIVV, R_CA, R_GA, R_GB}); // divv and c_ab rewritten
    SPHB200_CATCH
}

extern "C" int sphb200_av_switches(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, const sphb200_box* box,
                                   const sphb200_hydro_const* hc, const uint32_t* neighbors, const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    beginOp(ctx);
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             auto* r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_PDIVV, *f));
             auto* r1 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_VV, *f));
             if (useSmemTable(ctx, hc, TAB_AV, last - first))
                 LAUNCH_PBC_SMEM(ctx, *box, avSwitchesSmemKernel, kAvSmemBlock, last - first, first, last, hc->ngmax,
                                 hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1,
                                 static_cast<const T*>(hc->wh), hc->minDt, T(hc->alphamin), T(hc->alphamax),
                                 T(hc->decay_constant));
             else
                 LAUNCH_PBC(ctx, *box, avSwitchesKernel, gridFor(last - first, kHydroBlock), first, last, hc->ngmax,
                            hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1, wh, hc->minDt, T(hc->alphamin),
                            T(hc->alphamax), T(hc->decay_constant));
             tabLaunched(ctx, TAB_AV););
    SPHB200_CATCH
}

extern "C" int sphb200_momentum_energy(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                       const sphb200_box* box, const sphb200_hydro_const* hc,
                                       const uint32_t* neighbors, const sphb200_fields* f, double* minDtCourant_h)
{
    SPHB200_TRY
    SPHB200_LAUNCH(ctx, initCourantKeyKernel, 1, 1, 0, dtKeys(ctx));
    if (last > first)
    {
        beginOp(ctx);
        if (hc->avClean && !f->dV11) throw std::runtime_error("momentum_energy: avClean requires dV11..dV33");
        DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
                 launchMomentum<T>(ctx, first, first, last, true, box, hc, neighbors, f, wh, TAB_MOMENTUM););
    }
    if (minDtCourant_h)
    {
        unsigned long long key;
        SPHB200_CUDA(cudaMemcpyAsync(&key, dtKeys(ctx), sizeof(key), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        *minDtCourant_h = fromOrderedKey(key);
    }
    SPHB200_CATCH
}

extern "C" int sphb200_momentum_energy_range(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, uint32_t begin,
                                             uint32_t end, int pack, const sphb200_box* box,
                                             const sphb200_hydro_const* hc, const uint32_t* neighbors,
                                             const sphb200_fields* f)
{
    SPHB200_TRY
    if (begin < first || end > last || (begin - first) % kHydroBlock != 0)
        throw std::runtime_error("momentum_energy_range: [begin,end) must lie in [first,last), begin - first a multiple of 128");
    if (end <= begin) return 0;
    beginOp(ctx);
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             launchMomentum<T>(ctx, first, begin, end, pack != 0, box, hc, neighbors, f, wh, TAB_MOMENTUM_RANGE););
    SPHB200_CATCH
}

extern "C" int sphb200_av_switches_courant(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                           const sphb200_box* box, const sphb200_hydro_const* hc,
                                           const uint32_t* neighbors, const sphb200_fields* f)
{
    SPHB200_TRY
    if (hc->avClean) throw std::runtime_error("av_switches_courant: not available with avClean (rv needs grad v)");
    SPHB200_LAUNCH(ctx, initCourantKeyKernel, 1, 1, 0, dtKeys(ctx));
    if (last <= first) return 0;
    beginOp(ctx);
    DISPATCH(dtype, const Rec2<T>* wh; const Rec4<T>* whq; ensureTables<T>(ctx, hc, wh, whq);
             auto* r0 = static_cast<const PosRec<T>*>(ensureRec<T>(ctx, R_PDIVV, *f));
             auto* r1 = static_cast<const Rec4<T>*>(ensureRec<T>(ctx, R_VV, *f));
             if (useSmemTable(ctx, hc, TAB_AVC, last - first))
                 LAUNCH_PBC_SMEM(ctx, *box, avSwitchesCourantSmemKernel, kAvCourantSmemBlock, last - first, first, last,
                                 hc->ngmax, hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1,
                                 static_cast<const T*>(hc->wh), hc->minDt, T(hc->alphamin), T(hc->alphamax),
                                 T(hc->decay_constant), float(hc->Kcour), dtKeys(ctx));
             else
                 LAUNCH_PBC(ctx, *box, avSwitchesCourantKernel, gridFor(last - first, kHydroBlock), first, last,
                            hc->ngmax, hc->K, pbcBox(*box), neighbors, scoped<T>(ctx, *f), r0, r1, wh, hc->minDt,
                            T(hc->alphamin), T(hc->alphamax), T(hc->decay_constant), float(hc->Kcour), dtKeys(ctx));
             tabLaunched(ctx, TAB_AVC););
    SPHB200_CATCH
}

extern "C" int sphb200_timestep(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                const sphb200_hydro_const* hc, const sphb200_fields* f, double* out_h)
{
    SPHB200_TRY
    SPHB200_LAUNCH(ctx, initDivvKeyKernel, 1, 1, 0, dtKeys(ctx));
    if (last > first)
    {
        int grid = persistentGrid(last - first, 256, 8);
        DISPATCH(dtype, SPHB200_LAUNCH(ctx, maxDivvKernel<T>, grid, 256, 0, static_cast<const T*>(f->divv), first, last,
                                       dtKeys(ctx) + 1););
    }
    SPHB200_LAUNCH(ctx, finalizeDtKernel, 1, 1, 0, dtKeys(ctx), hc->Krho, 1.1, hc->minDt, ctx->devDt,
                   int(dtype == SPHB200_MIXED));
    if (out_h)
    {
        SPHB200_CUDA(cudaMemcpyAsync(ctx->hostDt, ctx->devDt, 3 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
        out_h[0] = ctx->hostDt[0];
        out_h[1] = ctx->hostDt[1];
        out_h[2] = ctx->hostDt[2];
    }
    SPHB200_CATCH
}

extern "C" double* sphb200_timestep_device_ptr(sphb200_ctx* ctx) { return ctx->devDt; }

extern "C" int sphb200_positions_update_h(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                          const sphb200_box* box, const sphb200_hydro_const* hc, double dt,
                                          double dt_m1, const sphb200_fields* f)
{
    SPHB200_TRY
    if (last <= first) return 0;
    ctx->packValid = 0;
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, positionsKernel<T>, gridFor(last - first, 256), 256, 0, first, last, dt, dt_m1,
                                   pbcBox(*box), typed<T>(*f), T(hc->muiConst), hc->gamma, hc->ng0, f->x, f->y, f->z););
    SPHB200_CATCH
}

// ---------------------------------------------------------------- block time steps: group operators
extern "C" int sphb200_drift_positions(sphb200_ctx* ctx, int dtype, const sphb200_group_view* grp, float dt,
                                       float dt_back, const float* dt_m1_h, const uint8_t* rung,
                                       const sphb200_hydro_const* hc, const sphb200_fields* f)
{
    SPHB200_TRY
    if (grp->numGroups == 0) return 0;
    RungDt r;
    for (int k = 0; k < 4; ++k)
        r.v[k] = dt_m1_h[k];
    int grid = gridFor(size_t(grp->numGroups) * 32, 256);
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, driftGroupsKernel<T>, grid, 256, 0, grp->groupStart, grp->groupEnd,
                                   grp->numGroups, dt, dt_back, r, rung, typed<T>(*f), T(hc->muiConst), hc->gamma, f->x,
                                   f->y, f->z););
    ctx->packValid = 0; // positions, velocities changed
    SPHB200_CATCH
}

extern "C" int sphb200_compute_positions_groups(sphb200_ctx* ctx, int dtype, const sphb200_group_view* grp, float dt,
                                                const float* dt_m1_h, const uint8_t* rung, const sphb200_box* box,
                                                const sphb200_hydro_const* hc, const sphb200_fields* f)
{
    SPHB200_TRY
    if (grp->numGroups == 0) return 0;
    RungDt r;
    for (int k = 0; k < 4; ++k)
        r.v[k] = dt_m1_h[k];
    int grid = gridFor(size_t(grp->numGroups) * 32, 256);
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, positionsGroupsKernel<T>, grid, 256, 0, grp->groupStart, grp->groupEnd,
                                   grp->numGroups, dt, r, rung, pbcBox(*box), typed<T>(*f), T(hc->muiConst), hc->gamma,
                                   f->x, f->y, f->z););
    ctx->packValid = 0;
    SPHB200_CATCH
}

extern "C" int sphb200_update_h_groups(sphb200_ctx* ctx, int dtype, const sphb200_group_view* grp, unsigned ng0,
                                       const uint32_t* nc, void* h)
{
    SPHB200_TRY
    if (grp->numGroups == 0) return 0;
    int grid = gridFor(size_t(grp->numGroups) * 32, 256);
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, updateHGroupsKernel<T>, grid, 256, 0, grp->groupStart, grp->groupEnd,
                                   grp->numGroups, ng0, nc, static_cast<T*>(h)););
    ctx->packValid = 0;
    SPHB200_CATCH
}

extern "C" int sphb200_update_h_into(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, unsigned ng0,
                                     const uint32_t* nc, const void* h, void* hNew)
{
    SPHB200_TRY
    if (last <= first) return 0;
    DISPATCH(dtype, SPHB200_LAUNCH(ctx, updateHRangeKernel<T>, gridFor(last - first, 256), 256, 0, first, last, ng0, nc,
                                   static_cast<const T*>(h), static_cast<T*>(hNew)););
    SPHB200_CATCH
}
