// Conserved-quantity observables on the device.
//
// sphexa::computeConservedQuantities (main/src/observables/conserved_quantities.hpp:50-180, conserved_gpu.cu) runs
// every step: kinetic and internal energy, linear and angular momentum, total neighbor count.  Here: ONE grid-stride
// pass over the owned particles with per-thread fp64 accumulators, a fixed-shape block reduction and a second
// single-block pass over the per-block partials in index order, so the result is reproducible run to run and
// independent of scheduling.  The 10 values stay on the device for an in-stream all-reduce (multi-rank) and are
// copied to the host only when asked for.
#include "common.cuh"

namespace sphb200
{

constexpr int kObsThreads = 256;
constexpr int kObsValues  = 10; // eKin, eInt, egrav, linmom[3], angmom[3], ncsum  (conserved_quantities.hpp:150-159)

template<class T>
__global__ void __launch_bounds__(kObsThreads)
    conservedPartialsKernel(uint32_t first, uint32_t last, const double* __restrict__ x, const double* __restrict__ y,
                            const double* __restrict__ z, const T* __restrict__ vx, const T* __restrict__ vy,
                            const T* __restrict__ vz, const T* __restrict__ m, const double* __restrict__ temp,
                            const uint32_t* __restrict__ nc, double cv, double* __restrict__ partials)
{
    double acc[kObsValues] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (size_t i = size_t(first) + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < last;
         i += size_t(gridDim.x) * blockDim.x)
    {
        // conserved_quantities.hpp:71-79,97-103: X, V in double, mass in its storage type
        double X[3] = {x[i], y[i], z[i]};
        double V[3] = {double(vx[i]), double(vy[i]), double(vz[i])};
        T mi        = m[i];
        acc[0] += mi * (V[0] * V[0] + V[1] * V[1] + V[2] * V[2]);
        acc[1] += cv * temp[i] * mi;
        acc[3] += mi * V[0];
        acc[4] += mi * V[1];
        acc[5] += mi * V[2];
        acc[6] += mi * (X[1] * V[2] - X[2] * V[1]);
        acc[7] += mi * (X[2] * V[0] - X[0] * V[2]);
        acc[8] += mi * (X[0] * V[1] - X[1] * V[0]);
        if (nc) acc[9] += double(nc[i]);
    }
    __shared__ double sh[kObsThreads / 32][kObsValues];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kObsValues; ++k)
    {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
            v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) sh[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < kObsValues)
    {
        double v = 0;
        for (int w = 0; w < kObsThreads / 32; ++w)
            v += sh[w][threadIdx.x];
        partials[size_t(blockIdx.x) * kObsValues + threadIdx.x] = v;
    }
}

//! @brief sum of the per-block partials in block order; out[0] gets the 1/2 of the kinetic energy
__global__ void conservedFinalKernel(const double* __restrict__ partials, int numBlocks, double* __restrict__ out)
{
    int k = threadIdx.x;
    if (k >= kObsValues) return;
    double v = 0;
    for (int b = 0; b < numBlocks; ++b)
        v += partials[size_t(b) * kObsValues + k];
    out[k] = (k == 0) ? 0.5 * v : v;
}

/*! @brief localMachSquareSum (main/src/observables/turbulence_mach_rms.hpp:38-56, machSquareSumGpu gpu_reductions.cu:
 *         110-122): sum of |v|^2 / c^2 over the owned particles, same fixed-order reduction as above (slot 0)
 */
template<class T>
__global__ void __launch_bounds__(kObsThreads)
    machPartialsKernel(uint32_t first, uint32_t last, const T* __restrict__ vx, const T* __restrict__ vy,
                       const T* __restrict__ vz, const T* __restrict__ c, double* __restrict__ partials)
{
    double acc = 0;
    for (size_t i = size_t(first) + size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < last;
         i += size_t(gridDim.x) * blockDim.x)
    {
        double V[3] = {double(vx[i]), double(vy[i]), double(vz[i])};
        acc += (V[0] * V[0] + V[1] * V[1] + V[2] * V[2]) / (c[i] * c[i]);
    }
    __shared__ double sh[kObsThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) sh[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double v = 0;
        for (int w = 0; w < kObsThreads / 32; ++w)
            v += sh[w];
        partials[blockIdx.x] = v;
    }
}

__global__ void machFinalKernel(const double* __restrict__ partials, int numBlocks, double* __restrict__ out)
{
    double v = 0;
    for (int b = 0; b < numBlocks; ++b)
        v += partials[b];
    out[0] = v;
}

} // namespace sphb200

using namespace sphb200;

extern "C" int sphb200_conserved_quantities(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                            const sphb200_hydro_const* hc, const sphb200_fields* f, double* out_h)
{
    try
    {
        int grid         = last > first ? persistentGrid(last - first, kObsThreads, 4) : 1;
        auto* partials   = static_cast<double*>(ctx->obsTmp.get((size_t(grid) + 2) * kObsValues * sizeof(double)));
        double* out      = partials + size_t(grid) * kObsValues;
        // sph::idealGasCv (sph/eos.hpp:13-21) in double, as conserved_quantities.hpp:64
        const double R = 8.317e7;
        double cv      = R / hc->muiConst / (hc->gamma - 1.0);
        if (dtype == SPHB200_F64)
            SPHB200_LAUNCH(ctx, conservedPartialsKernel<double>, grid, kObsThreads, 0, first, last, f->x, f->y, f->z,
                           static_cast<const double*>(f->vx), static_cast<const double*>(f->vy),
                           static_cast<const double*>(f->vz), static_cast<const double*>(f->m), f->temp, f->nc, cv,
                           partials);
        else if (dtype == SPHB200_MIXED)
            SPHB200_LAUNCH(ctx, conservedPartialsKernel<float>, grid, kObsThreads, 0, first, last, f->x, f->y, f->z,
                           static_cast<const float*>(f->vx), static_cast<const float*>(f->vy),
                           static_cast<const float*>(f->vz), static_cast<const float*>(f->m), f->temp, f->nc, cv,
                           partials);
        else throw std::runtime_error("unknown dtype");
        SPHB200_LAUNCH(ctx, conservedFinalKernel, 1, 32, 0, partials, grid, out);
        ctx->devObs = out;
        if (out_h)
        {
            SPHB200_CUDA(cudaMemcpyAsync(ctx->hostDt + 8, out, kObsValues * sizeof(double), cudaMemcpyDeviceToHost,
                                         ctx->stream));
            SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
            for (int k = 0; k < kObsValues; ++k)
                out_h[k] = ctx->hostDt[8 + k];
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" double* sphb200_conserved_device_ptr(sphb200_ctx* ctx) { return ctx->devObs; }

extern "C" int sphb200_mach_square_sum(sphb200_ctx* ctx, int dtype, uint32_t first, uint32_t last,
                                       const sphb200_fields* f, double* out_h)
{
    try
    {
        int grid       = last > first ? persistentGrid(last - first, kObsThreads, 4) : 1;
        auto* partials = static_cast<double*>(ctx->machTmp.get((size_t(grid) + 2) * sizeof(double)));
        double* out      = partials + grid;
        if (dtype == SPHB200_F64)
            SPHB200_LAUNCH(ctx, machPartialsKernel<double>, grid, kObsThreads, 0, first, last,
                           static_cast<const double*>(f->vx), static_cast<const double*>(f->vy),
                           static_cast<const double*>(f->vz), static_cast<const double*>(f->c), partials);
        else if (dtype == SPHB200_MIXED)
            SPHB200_LAUNCH(ctx, machPartialsKernel<float>, grid, kObsThreads, 0, first, last,
                           static_cast<const float*>(f->vx), static_cast<const float*>(f->vy),
                           static_cast<const float*>(f->vz), static_cast<const float*>(f->c), partials);
        else throw std::runtime_error("unknown dtype");
        SPHB200_LAUNCH(ctx, machFinalKernel, 1, 1, 0, partials, grid, out);
        ctx->devMach = out;
        if (out_h)
        {
            SPHB200_CUDA(cudaMemcpyAsync(ctx->hostDt + 20, out, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
            SPHB200_CUDA(cudaStreamSynchronize(ctx->stream));
            *out_h = ctx->hostDt[20];
        }
        return 0;
    }
    catch (std::exception& e)
    {
        setError(e.what());
        return 1;
    }
}

extern "C" double* sphb200_mach_device_ptr(sphb200_ctx* ctx) { return ctx->devMach; }
