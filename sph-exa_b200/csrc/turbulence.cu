// Turbulence stirring: Ornstein-Uhlenbeck driven Fourier modes added to the accelerations.
//
// Reference: sph/include/sph/hydro_turb/{stirring.hpp:43-122, stirring_gpu.cu:43-93, driver.hpp:79-130,
// phases.hpp:46-72, create_modes.hpp:57-244, turbulence_data.hpp:57-183}, driven from TurbVeProp::computeForces
// (main/src/propagator/turb_ve.hpp:113-118) right after the hydro forces.
//
// Split like the reference: the mode table, the OU phase sequence (std::mt19937 + std::normal_distribution: the stream is
// a property of the host library, so it stays on the host) and the projection onto real/imaginary phases are host code;
// the per-particle sum over modes is the kernel.  Two kernels:
//  * stirringKernel      -- the operator as the reference states it: sin/cos of k_m . x per mode and axis
//  * stirringTabledKernel -- used by sphb200_turb: the wave numbers of a mode table are a handful of distinct values per
//    axis (2 pi i / L, i = 0..3 and their negatives for the band/parabola spectra), so sin/cos are evaluated once per
//    distinct value and particle (9 sincos instead of 6 x 112) and looked up per mode.  Same arguments, same functions,
//    identical results.
#include <algorithm>
#include <cmath>
#include <memory>
#include <random>
#include <sstream>

#include "common.cuh"

namespace sphb200
{

constexpr int kStirBlock   = 128;
constexpr int kMaxUnique   = 8;  // distinct wave numbers per axis handled by the tabled kernel
constexpr int kMaxModesSm  = 2048;

//! @brief mode table in shared memory: per mode kx ky kz | pRx pRy pRz | pIx pIy pIz | amplitude
template<class Ta>
__global__ void __launch_bounds__(kStirBlock)
    stirringKernel(uint32_t first, uint32_t last, const double* __restrict__ x, const double* __restrict__ y,
                   const double* __restrict__ z, Ta* __restrict__ ax, Ta* __restrict__ ay, Ta* __restrict__ az,
                   int numModes, const double* __restrict__ modes, const double* __restrict__ phaseReal,
                   const double* __restrict__ phaseImag, const double* __restrict__ amplitudes, double solWeightNorm,
                   const uint8_t* __restrict__ activeMask)
{
    extern __shared__ double sm[];
    double* sModes = sm;
    double* sRe    = sm + 3 * numModes;
    double* sIm    = sm + 6 * numModes;
    double* sAmp   = sm + 9 * numModes;
    for (int k = threadIdx.x; k < 3 * numModes; k += blockDim.x)
    {
        sModes[k] = modes[k];
        sRe[k]    = phaseReal[k];
        sIm[k]    = phaseImag[k];
    }
    for (int k = threadIdx.x; k < numModes; k += blockDim.x)
        sAmp[k] = amplitudes[k];
    __syncthreads();

    uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last || (activeMask && !activeMask[i])) return; // GroupView scope (TurbVeBdtProp stirs the active rungs)
    double xi = x[i], yi = y[i], zi = z[i];
    // stirParticle (stirring.hpp:43-81): accumulators in the type of the accelerations
    Ta tx = 0, ty = 0, tz = 0;
    for (int m = 0; m < numModes; ++m)
    {
        double sk, ck, sj, cj, si, ci;
        sincos(sModes[3 * m + 2] * zi, &sk, &ck);
        sincos(sModes[3 * m + 1] * yi, &sj, &cj);
        sincos(sModes[3 * m] * xi, &si, &ci);
        double re = (ci * cj - si * sj) * ck - (si * cj + ci * sj) * sk;
        double im = ci * (cj * sk + sj * ck) + si * (cj * ck - sj * sk);
        double a  = sAmp[m];
        tx += a * (sRe[3 * m] * re - sIm[3 * m] * im);
        ty += a * (sRe[3 * m + 1] * re - sIm[3 * m + 1] * im);
        tz += a * (sRe[3 * m + 2] * re - sIm[3 * m + 2] * im);
    }
    ax[i] += solWeightNorm * tx;
    ay[i] += solWeightNorm * ty;
    az[i] += solWeightNorm * tz;
}

struct UniqueK
{
    double k[3][kMaxUnique]; // distinct wave numbers per axis
    int n[3];
};

//! @brief as above with sin/cos per distinct wave number; idx[3*m + axis] selects the table entry of mode m
template<class Ta>
__global__ void __launch_bounds__(kStirBlock)
    stirringTabledKernel(uint32_t first, uint32_t last, const double* __restrict__ x, const double* __restrict__ y,
                         const double* __restrict__ z, Ta* __restrict__ ax, Ta* __restrict__ ay, Ta* __restrict__ az,
                         int numModes, UniqueK uk, const unsigned char* __restrict__ idx,
                         const double* __restrict__ phaseReal, const double* __restrict__ phaseImag,
                         const double* __restrict__ amplitudes, double solWeightNorm,
                         const uint8_t* __restrict__ activeMask)
{
    extern __shared__ double sm[];
    double* sRe  = sm;
    double* sIm  = sm + 3 * numModes;
    double* sAmp = sm + 6 * numModes;
    // per-thread sin/cos table, entry-major so that a warp reads consecutive words: [axis][u][sin|cos][thread]
    double* trig = sm + 7 * numModes;
    unsigned char* sIdx = reinterpret_cast<unsigned char*>(trig + 3 * kMaxUnique * 2 * kStirBlock);
    for (int k = threadIdx.x; k < 3 * numModes; k += blockDim.x)
    {
        sRe[k]  = phaseReal[k];
        sIm[k]  = phaseImag[k];
        sIdx[k] = idx[k];
    }
    for (int k = threadIdx.x; k < numModes; k += blockDim.x)
        sAmp[k] = amplitudes[k];

    uint32_t i   = first + blockIdx.x * blockDim.x + threadIdx.x;
    bool active  = i < last && (!activeMask || activeMask[i]);
    double p[3]  = {0, 0, 0};
    if (active)
    {
        p[0] = x[i];
        p[1] = y[i];
        p[2] = z[i];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a)
        for (int u = 0; u < uk.n[a]; ++u)
        {
            double s, c;
            sincos(uk.k[a][u] * p[a], &s, &c);
            trig[((a * kMaxUnique + u) * 2) * kStirBlock + threadIdx.x]     = s;
            trig[((a * kMaxUnique + u) * 2 + 1) * kStirBlock + threadIdx.x] = c;
        }
    __syncthreads();
    if (!active) return;

    Ta tx = 0, ty = 0, tz = 0;
    for (int m = 0; m < numModes; ++m)
    {
        const double* ti = trig + ((0 * kMaxUnique + sIdx[3 * m]) * 2) * kStirBlock + threadIdx.x;
        const double* tj = trig + ((1 * kMaxUnique + sIdx[3 * m + 1]) * 2) * kStirBlock + threadIdx.x;
        const double* tk = trig + ((2 * kMaxUnique + sIdx[3 * m + 2]) * 2) * kStirBlock + threadIdx.x;
        double si = ti[0], ci = ti[kStirBlock], sj = tj[0], cj = tj[kStirBlock], sk = tk[0], ck = tk[kStirBlock];
        double re = (ci * cj - si * sj) * ck - (si * cj + ci * sj) * sk;
        double im = ci * (cj * sk + sj * ck) + si * (cj * ck - sj * sk);
        double a  = sAmp[m];
        tx += a * (sRe[3 * m] * re - sIm[3 * m] * im);
        ty += a * (sRe[3 * m + 1] * re - sIm[3 * m + 1] * im);
        tz += a * (sRe[3 * m + 2] * re - sIm[3 * m + 2] * im);
    }
    ax[i] += solWeightNorm * tx;
    ay[i] += solWeightNorm * ty;
    az[i] += solWeightNorm * tz;
}

template<class Ta>
static void launchStirring(sphb200_ctx* ctx, uint32_t first, uint32_t last, const double* x, const double* y,
                           const double* z, void* ax, void* ay, void* az, int numModes, const double* modes,
                           const double* pr, const double* pi, const double* amp, double solWeightNorm)
{
    if (numModes > kMaxModesSm) throw std::runtime_error("stirring: too many modes for the shared-memory table");
    size_t smem = size_t(numModes) * 10 * sizeof(double);
    auto kern   = stirringKernel<Ta>;
    if (smem > 48 * 1024)
        SPHB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    kern<<<gridFor(last - first, kStirBlock), kStirBlock, smem, ctx->stream>>>(
        first, last, x, y, z, static_cast<Ta*>(ax), static_cast<Ta*>(ay), static_cast<Ta*>(az), numModes, modes, pr, pi,
        amp, solWeightNorm, ctx->scopeActive);
    ctx->launches++;
    SPHB200_CUDA(cudaGetLastError());
}

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_stirring(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last, const double* x,
                                const double* y, const double* z, void* ax, void* ay, void* az, int numModes,
                                const double* modes, const double* phaseReal, const double* phaseImag,
                                const double* amplitudes, double solWeightNorm)
{
    try
    {
        if (last <= first || numModes <= 0) return 0;
        if (dtype == SPHB200_F64)
            launchStirring<double>(ctx, first, last, x, y, z, ax, ay, az, numModes, modes, phaseReal, phaseImag,
                                   amplitudes, solWeightNorm);
        else if (dtype == SPHB200_MIXED)
            launchStirring<float>(ctx, first, last, x, y, z, ax, ay, az, numModes, modes, phaseReal, phaseImag,
                                  amplitudes, solWeightNorm);
        else throw std::runtime_error("unknown dtype");
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

// ---- host side: mode table, OU sequence, phase projection -------------------------------------------------------------
struct sphb200_turb
{
    sphb200_turb_config cfg;
    int numDim = 3;
    double variance = 0, decayTime = 0, solWeight = 0, solWeightNorm = 0;
    int numModes = 0;
    std::vector<double> modes, amplitudes, phases, phasesReal, phasesImag;
    std::mt19937 gen;
    // device copies, made by the first call that launches a kernel (ensureDevice): creating the driver and advancing
    // its Ornstein-Uhlenbeck sequence are host work and need no device
    double *dModes = nullptr, *dAmp = nullptr, *dRe = nullptr, *dIm = nullptr;
    unsigned char* dIdx = nullptr;
    std::vector<unsigned char> idxHost;
    bool onDevice = false, phasesStale = true;
    UniqueK uk;
    bool tabled = false;
};

namespace
{

struct Mode
{
    double k[3], amplitude;
};

//! @brief wave vectors and amplitudes of the driving modes (create_modes.hpp:57-244); spectForm 0 band, 1 parabola,
//!        2 power law with randomly sampled angles (consumes the generator)
std::vector<Mode> makeModes(double L, size_t maxModes, double kMin, double kMax, int ndim, int spectForm,
                            double powerLawExp, double anglesExp, std::mt19937& gen)
{
    const double twopi = 2.0 * M_PI;
    std::vector<Mode> out;
    double kc = (spectForm == 1) ? 0.5 * (kMin + kMax) : kMin;
    auto room = [&]() { return double(out.size()) + std::pow(2, ndim - 1) <= double(maxModes); };

    if (spectForm != 2)
    {
        // full sampling of the non-negative octant; the sign variants of ky, kz follow each accepted vector
        double parabola = -4.0 / ((kMax - kMin) * (kMax - kMin));
        int nx = 256, ny = ndim > 1 ? 256 : 0, nz = ndim > 2 ? 256 : 0;
        for (int ix = 0; ix <= nx; ++ix)
        {
            double kx = twopi * ix / L;
            for (int iy = 0; iy <= ny; ++iy)
            {
                double ky = twopi * iy / L;
                for (int iz = 0; iz <= nz; ++iz)
                {
                    double kz = twopi * iz / L;
                    double k  = std::sqrt(kx * kx + ky * ky + kz * kz);
                    if (!(k >= kMin && k <= kMax)) continue;
                    if (!room()) break;
                    double amp = 1.0;
                    if (spectForm == 1) amp = std::abs(parabola * (k - kc) * (k - kc) + 1.0);
                    amp = 2.0 * std::sqrt(amp) * std::pow(kc / k, 0.5 * (ndim - 1));
                    out.push_back({{kx, ky, kz}, amp});
                    if (ndim > 1) out.push_back({{kx, -ky, kz}, amp});
                    if (ndim > 2)
                    {
                        out.push_back({{kx, ky, -kz}, amp});
                        out.push_back({{kx, -ky, -kz}, amp});
                    }
                }
            }
        }
        return out;
    }

    std::uniform_real_distribution<double> uni(0, 1);
    int ikLo = std::max(1, int(kMin * L / twopi + 0.5));
    int ikHi = int(kMax * L / twopi + 0.5);
    for (int ik = ikLo; ik <= ikHi; ++ik)
    {
        size_t nang = size_t(std::pow(2, ndim) * std::ceil(std::pow(ik, anglesExp)));
        for (size_t a = 1; a <= nang; ++a)
        {
            double phi = twopi * uni(gen);
            if (ndim == 1) phi = (phi < twopi / 2) ? 0.0 : twopi / 2.0;
            double theta = twopi / 4.0;
            if (ndim > 2) theta = std::acos(1.0 - 2.0 * uni(gen));
            double rad = ik + uni(gen) - 0.5;
            double kx  = twopi * std::round(rad * std::sin(theta) * std::cos(phi)) / L;
            double ky  = (ndim > 1) ? twopi * std::round(rad * std::sin(theta) * std::sin(phi)) / L : 0.0;
            double kz  = (ndim > 2) ? twopi * std::round(rad * std::cos(theta)) / L : 0.0;
            double k   = std::sqrt(kx * kx + ky * ky + kz * kz);
            if (!(k >= kMin && k <= kMax)) continue;
            if (!room()) break;
            double amp = std::pow(k / kc, powerLawExp);
            amp = std::sqrt(amp * (std::pow(ik, ndim - 1) * 4.0 * std::sqrt(3.0) / nang)) *
                  std::pow(kc / k, (ndim - 1) / 2.0);
            out.push_back({{kx, ky, kz}, amp});
        }
    }
    return out;
}

template<class V>
double* toDevice(const V& v)
{
    double* d = nullptr;
    SPHB200_CUDA(cudaMalloc(&d, std::max<size_t>(1, v.size()) * sizeof(double)));
    SPHB200_CUDA(cudaMemcpy(d, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice));
    return d;
}

//! @brief OU step x <- f x + sigma sqrt(1 - f^2) z (driver.hpp:79-92) and the solenoidal/compressive projection
//!        (phases.hpp:46-72)
void advancePhases(sphb200_turb* t, double dt)
{
    double f = std::exp(-dt / t->decayTime);
    double g = std::sqrt(1.0 - f * f);
    std::normal_distribution<double> unit(0, 1);
    for (double& p : t->phases)
    {
        double zn = unit(t->gen);
        p         = p * f + t->variance * g * zn;
    }
}

void projectPhases(sphb200_turb* t)
{
    int nd = t->numDim;
    for (int i = 0; i < t->numModes; ++i)
    {
        const double* k  = &t->modes[3 * i];
        const double* ou = &t->phases[6 * i];
        double ka = 0, kb = 0, kk = 0;
        for (int j = 0; j < nd; ++j)
        {
            kk = kk + k[j] * k[j];
            ka = ka + k[j] * ou[2 * j + 1];
            kb = kb + k[j] * ou[2 * j];
        }
        for (int j = 0; j < nd; ++j)
        {
            double diva = k[j] * ka / kk, divb = k[j] * kb / kk;
            double curla = ou[2 * j] - divb, curlb = ou[2 * j + 1] - diva;
            t->phasesReal[3 * i + j] = t->solWeight * curla + (1.0 - t->solWeight) * divb;
            t->phasesImag[3 * i + j] = t->solWeight * curlb + (1.0 - t->solWeight) * diva;
        }
    }
}

} // namespace

extern "C" int sphb200_turb_create(const sphb200_turb_config* c, sphb200_turb** out)
{
    try
    {
        std::unique_ptr<sphb200_turb> owner(new sphb200_turb);
        sphb200_turb* t = owner.get();
        t->cfg       = *c;
        t->solWeight = c->solWeight;
        t->gen.seed(size_t(c->rngSeed));
        // TurbulenceData::initModes (turbulence_data.hpp:150-183)
        const double twopi = 2.0 * M_PI;
        double energy  = c->stEnergyPrefac * std::pow(c->stMachVelocity, 3) / c->Lbox;
        double stirMin = (1.0 - c->epsilon) * twopi / c->Lbox;
        double stirMax = (3.0 + c->epsilon) * twopi / c->Lbox;
        t->decayTime   = c->Lbox / (2.0 * c->stMachVelocity);
        t->variance    = std::sqrt(energy / t->decayTime);
        double nd      = double(t->numDim);
        t->solWeightNorm =
            std::sqrt(3.0) * std::sqrt(3.0 / nd) / std::sqrt(1.0 - 2.0 * t->solWeight + nd * t->solWeight * t->solWeight);

        auto modes  = makeModes(c->Lbox, size_t(c->stMaxModes), stirMin, stirMax, t->numDim, c->stSpectForm, c->powerLawExp,
                                c->anglesExp, t->gen);
        t->numModes = int(modes.size());
        t->modes.resize(3 * modes.size());
        t->amplitudes.resize(modes.size());
        for (size_t m = 0; m < modes.size(); ++m)
        {
            for (int a = 0; a < 3; ++a)
                t->modes[3 * m + a] = modes[m].k[a];
            t->amplitudes[m] = modes[m].amplitude;
        }
        t->phases.resize(6 * modes.size());
        t->phasesReal.assign(3 * modes.size(), 0.0);
        t->phasesImag.assign(3 * modes.size(), 0.0);
        // initial OU state: gaussian with standard deviation `variance` (turbulence_data.hpp:180-182)
        std::normal_distribution<double> dist(0, t->variance);
        for (double& p : t->phases)
            p = dist(t->gen);

        // distinct wave numbers per axis for the tabled kernel
        std::vector<unsigned char> idx(3 * modes.size());
        t->tabled = true;
        for (int a = 0; a < 3; ++a)
        {
            t->uk.n[a] = 0;
            for (size_t m = 0; m < modes.size() && t->tabled; ++m)
            {
                double k = modes[m].k[a];
                int u    = 0;
                while (u < t->uk.n[a] && t->uk.k[a][u] != k)
                    ++u;
                if (u == t->uk.n[a])
                {
                    if (u == kMaxUnique) { t->tabled = false; break; }
                    t->uk.k[a][t->uk.n[a]++] = k;
                }
                idx[3 * m + a] = (unsigned char)u;
            }
        }
        t->idxHost = std::move(idx);
        *out       = owner.release();
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

//! @brief mode table and phases on the current device (once), phases refreshed if the host advanced them since
static void ensureDevice(sphb200_turb* t, cudaStream_t stream)
{
    if (!t->onDevice)
    {
        t->dModes = toDevice(t->modes);
        t->dAmp   = toDevice(t->amplitudes);
        t->dRe    = toDevice(t->phasesReal);
        t->dIm    = toDevice(t->phasesImag);
        SPHB200_CUDA(cudaMalloc(&t->dIdx, std::max<size_t>(1, t->idxHost.size())));
        SPHB200_CUDA(cudaMemcpy(t->dIdx, t->idxHost.data(), t->idxHost.size(), cudaMemcpyHostToDevice));
        t->onDevice    = true;
        t->phasesStale = false;
    }
    if (t->phasesStale)
    {
        size_t bytes = t->phasesReal.size() * sizeof(double);
        // pageable source: the copies complete (w.r.t. the host vectors) before cudaMemcpyAsync returns
        SPHB200_CUDA(cudaMemcpyAsync(t->dRe, t->phasesReal.data(), bytes, cudaMemcpyHostToDevice, stream));
        SPHB200_CUDA(cudaMemcpyAsync(t->dIm, t->phasesImag.data(), bytes, cudaMemcpyHostToDevice, stream));
        t->phasesStale = false;
    }
}

extern "C" void sphb200_turb_destroy(sphb200_turb* t)
{
    if (!t) return;
    cudaFree(t->dModes); cudaFree(t->dAmp); cudaFree(t->dRe); cudaFree(t->dIm); cudaFree(t->dIdx);
    delete t;
}

extern "C" int sphb200_turb_num_modes(const sphb200_turb* t) { return t->numModes; }

extern "C" int sphb200_turb_get(const sphb200_turb* t, double* scalars4, double* modes, double* amplitudes,
                                double* phases, double* phasesReal, double* phasesImag)
{
    auto cp = [](double* dst, const std::vector<double>& v) { if (dst) std::copy(v.begin(), v.end(), dst); };
    if (scalars4)
    {
        scalars4[0] = t->variance; scalars4[1] = t->decayTime; scalars4[2] = t->solWeight; scalars4[3] = t->solWeightNorm;
    }
    cp(modes, t->modes); cp(amplitudes, t->amplitudes); cp(phases, t->phases); cp(phasesReal, t->phasesReal);
    cp(phasesImag, t->phasesImag);
    return 0;
}

//! @brief restart: overwrite the OU state (TurbulenceData::loadOrStore); rngState = text form of the engine or NULL
extern "C" int sphb200_turb_set_state(sphb200_turb* t, const double* phases, const char* rngState)
{
    try
    {
        if (phases) std::copy(phases, phases + t->phases.size(), t->phases.begin());
        if (rngState)
        {
            std::stringstream s;
            s << rngState;
            s >> t->gen;
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_turb_rng_state(const sphb200_turb* t, char* buf, size_t bufSize)
{
    std::stringstream s;
    s << t->gen;
    std::string str = s.str();
    if (buf && bufSize > str.size()) std::copy(str.begin(), str.end() + 1 - 1, buf), buf[str.size()] = 0;
    return int(str.size() + 1);
}

/*! @brief driveTurbulence (driver.hpp:101-128): advance the OU phases by dt, project, upload, add the stirring
 *         accelerations to ax, ay, az of the particles [first, last) */
extern "C" int sphb200_turb_advance(sphb200_turb* t, sphb200_ctx* ctx, double dt)
{
    try
    {
        advancePhases(t, dt);
        projectPhases(t);
        t->phasesStale = true;
        if (ctx) ensureDevice(t, ctx->stream); // ctx == NULL: host only, the next stir uploads
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_turb_stir(sphb200_turb* t, sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                 const double* x, const double* y, const double* z, void* ax, void* ay, void* az)
{
    try
    {
        if (last <= first || t->numModes == 0) return 0;
        ensureDevice(t, ctx->stream);
        if (!t->tabled || t->numModes > kMaxModesSm)
            return sphb200_stirring(ctx, dtype, first, last, x, y, z, ax, ay, az, t->numModes, t->dModes, t->dRe,
                                    t->dIm, t->dAmp, t->solWeightNorm);
        size_t smem = size_t(t->numModes) * 7 * sizeof(double) + size_t(3) * kMaxUnique * 2 * kStirBlock * sizeof(double) +
                      size_t(3) * t->numModes;
        int grid = gridFor(last - first, kStirBlock);
        if (dtype == SPHB200_F64)
        {
            auto kern = stirringTabledKernel<double>;
            SPHB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            kern<<<grid, kStirBlock, smem, ctx->stream>>>(first, last, x, y, z, static_cast<double*>(ax),
                                                         static_cast<double*>(ay), static_cast<double*>(az), t->numModes,
                                                         t->uk, t->dIdx, t->dRe, t->dIm, t->dAmp, t->solWeightNorm,
                                                         ctx->scopeActive);
        }
        else
        {
            auto kern = stirringTabledKernel<float>;
            SPHB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
            kern<<<grid, kStirBlock, smem, ctx->stream>>>(first, last, x, y, z, static_cast<float*>(ax),
                                                         static_cast<float*>(ay), static_cast<float*>(az), t->numModes,
                                                         t->uk, t->dIdx, t->dRe, t->dIm, t->dAmp, t->solWeightNorm,
                                                         ctx->scopeActive);
        }
        ctx->launches++;
        SPHB200_CUDA(cudaGetLastError());
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" int sphb200_turb_drive(sphb200_turb* t, sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                  const double* x, const double* y, const double* z, void* ax, void* ay, void* az,
                                  double dt)
{
    if (int rc = sphb200_turb_advance(t, ctx, dt)) return rc;
    return sphb200_turb_stir(t, ctx, dtype, first, last, x, y, z, ax, ay, az);
}
