// Compile/link check of the C++ host shim (and, on a GPU box, a 2-step Sedov 12^3 run through it).
#include <cmath>
#include <cstdio>
#include <string>
#include <vector>
#include "sphexa_b200.hpp"

using namespace sphb200::host;

int main(int argc, char** argv)
{
    bool run = argc > 1;
    if (!run)
    {
        std::printf("link ok\n");
        return 0;
    }
    try
    {
        const int side = 12;
        const size_t n = size_t(side) * side * side;
        std::vector<double> x(n), y(n), z(n), temp(n), h(n), m(n, 1.0 / n), zero(n, 0.0), alpha(n, 0.05);
        double step = 1.0 / side, hInit = 0.5 * std::cbrt(3.0 / (4 * M_PI) * 100 / n);
        double cv = 8.317e7 / 10.0 / (5.0 / 3.0 - 1.0), ener0 = 1.0 / std::pow(M_PI, 1.5) / 1e-3;
        for (size_t i = 0; i < n; ++i)
        {
            x[i] = -0.5 + 0.5 * step + (i % side) * step;
            y[i] = -0.5 + 0.5 * step + ((i / side) % side) * step;
            z[i] = -0.5 + 0.5 * step + (i / (side * side)) * step;
            h[i] = hInit;
            temp[i] = (ener0 * std::exp(-(x[i] * x[i] + y[i] * y[i] + z[i] * z[i]) / 0.01) + 1e-8) / cv;
        }
        // tabulated sinc^6 kernel (sph_kernel_tables.hpp): the caller owns the tables like ParticlesData does
        std::vector<double> wh(20000), whd(20000);
        auto sinc = [](double v) { double p = M_PI_2 * v; return v == 0.0 ? 1.0 : std::sin(p) / p; };
        for (int i = 0; i < 20000; ++i)
        {
            double v = i * (2.0 / 19999);
            wh[i]    = std::pow(sinc(v), 6.0);
            double p = M_PI_2 * v;
            whd[i]   = v == 0.0 ? 0.0 : 6.0 * std::pow(sinc(v), 5.0) * sinc(v) * M_PI_2 * (std::cos(p) / std::sin(p) - 1.0 / p);
        }
        Context ctx(0);
        sphb200_sim_config cfg{};
        cfg.dtype = SPHB200_F64; cfg.rank = 0; cfg.numRanks = 1; cfg.bucketSize = 64; cfg.bucketSizeFocus = 64;
        cfg.capacity = n;
        cfg.box = Box(-0.5, 0.5, SPHB200_PERIODIC);
        cfg.hc = sphb200_hydro_const{0.7904495894323034, 100, 150, 0.2, 0.06, 5.0 / 3.0, 10.0, 0.05, 1.0, 0.2,
                                     0.1, 0.2, 10.0, 1e-6, 1e-6, wh.data(), whd.data(), 0};
        HydroVeProp prop(ctx, cfg);
        sphb200_fields f{};
        f.x = x.data(); f.y = y.data(); f.z = z.data(); f.h = h.data(); f.m = m.data();
        f.vx = zero.data(); f.vy = zero.data(); f.vz = zero.data(); f.x_m1 = zero.data(); f.y_m1 = zero.data();
        f.z_m1 = zero.data(); f.temp = temp.data(); f.du_m1 = zero.data(); f.alpha = alpha.data();
        prop.upload(n, f);
        for (int s = 0; s < 2; ++s)
        {
            prop.computeForces();
            prop.integrate();
        }
        std::printf("shim ok: n=%u leaves=%d neighbors=%llu minDt=%.6e\n", prop.nParticles(),
                    prop.octreeProperties().numLeafNodes, (unsigned long long)prop.totalNeighbors(), prop.minDt());
        // restart: dump, reload into a second propagator, one more step on both, same time step and neighbor sum
        {
            std::string path = std::string(argc > 2 ? argv[2] : "/tmp") + "/sphb200_shim_restart.bin";
            prop.saveRestart(path, 2, n);
            HydroVeProp again(ctx, cfg);
            uint64_t it = again.loadRestart(path, SPHB200_F64);
            prop.computeForces();
            prop.integrate();
            again.computeForces();
            again.integrate();
            bool same = it == 2 && again.minDt() == prop.minDt() && again.ttot() == prop.ttot() &&
                        again.totalNeighbors() == prop.totalNeighbors();
            std::remove(path.c_str());
            if (!same)
            {
                std::fprintf(stderr, "restart mismatch: it=%llu dt %.17g vs %.17g\n", (unsigned long long)it,
                             again.minDt(), prop.minDt());
                return 3;
            }
            std::printf("restart ok\n");
        }
        // block time steps: one hierarchy of HydroVeBdtProp (full sync, then partial syncs until the next full one)
        {
            HydroVeBdtProp bdt(ctx, cfg);
            bdt.upload(n, f);
            int substeps = 0, partial = 0;
            do
            {
                partial += !bdt.isSynced();
                bdt.computeForces();
                bdt.integrate();
                ++substeps;
            } while (!bdt.isSynced() && substeps < 64);
            auto ts = bdt.timestep();
            std::printf("bdt ok: rungs=%d substeps=%d partial=%d groups=%u nextDt=%.6e\n", ts.numRungs, substeps, partial,
                        ts.numGroups, double(ts.nextDt));
            if (substeps != (1 << (ts.numRungs - 1))) return 4;
        }
        // two virtual ranks on this GPU (loopback transport): the same problem through the multi-rank Domain; the time
        // step and the neighbor total must equal the single-rank propagator's after the same number of steps
        {
            Comm comm = Comm::loopback(2);
            Context c0(0, nullptr), c1(0, nullptr); // one context per virtual rank, one stream
            sphb200_sim_config cr = cfg;
            cr.numRanks = 2;
            cr.capacity = 2 * n;
            cr.rank     = 0;
            HydroVeProp r0(c0, cr, comm);
            cr.rank = 1;
            HydroVeProp r1(c1, cr, comm);
            auto slice = [&](size_t first) {
                sphb200_fields g = f;
                g.x += first; g.y += first; g.z += first; g.temp += first;
                for (void** p : {&g.h, &g.m, &g.vx, &g.vy, &g.vz, &g.x_m1, &g.y_m1, &g.z_m1, &g.du_m1, &g.alpha})
                    *p = static_cast<double*>(*p) + first;
                return g;
            };
            r0.upload(n / 2, slice(0));
            r1.upload(n - n / 2, slice(n / 2));
            HydroVeProp one(ctx, cfg);
            one.upload(n, f);
            DistHydroVeProp dist({&r0, &r1});
            for (int s = 0; s < 3; ++s)
            {
                dist.step();
                one.computeForces();
                one.integrate();
            }
            auto i0 = dist.info(0), i1 = dist.info(1);
            bool same = r0.minDt() == one.minDt() && r1.minDt() == one.minDt() &&
                        r0.totalNeighbors() + r1.totalNeighbors() == one.totalNeighbors() &&
                        (i0.endIndex - i0.startIndex) + (i1.endIndex - i1.startIndex) == n;
            std::printf("two virtual ranks: owned %u + %u, halos %u + %u, dt %.6e\n", i0.endIndex - i0.startIndex,
                        i1.endIndex - i1.startIndex, i0.nWithHalos - (i0.endIndex - i0.startIndex),
                        i1.nWithHalos - (i1.endIndex - i1.startIndex), r0.minDt());
            if (!same)
            {
                std::fprintf(stderr, "2 virtual ranks differ from 1 rank: dt %.17g vs %.17g\n", r0.minDt(), one.minDt());
                return 5;
            }
            // the reference-style focus tree of both ranks (FocusedOctree::converge on the first call)
            dist.sync();
            bool conv = dist.updateFocus(1.0f);
            std::vector<int> peers;
            auto v0 = dist.focusView(0, &peers), v1 = dist.focusView(1);
            std::printf("focus trees: %d + %d leaves, rank 0 has %zu peer(s), converged %d\n", v0.numLeaves, v1.numLeaves,
                        peers.size(), int(conv));
            if (!conv || peers.size() != 1 || peers[0] != 1 || v0.numLeaves < 8 || !v0.counts || !v0.macs)
            {
                std::fprintf(stderr, "focus tree update failed\n");
                return 6;
            }
            std::printf("domain ok\n");
        }
        // error path: precondition violation -> std::runtime_error
        bool threw = false;
        try
        {
            prop.upload(n + 1, f);
        }
        catch (const std::runtime_error& e)
        {
            threw = std::string(e.what()).find("capacity") != std::string::npos;
        }
        return threw ? 0 : 2;
    }
    catch (const std::exception& e)
    {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
}
