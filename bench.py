#!/usr/bin/env python
"""bench.py — SPH particle-steps/sec of the SPH-VE full time step (domain sync + neighbors + all hydro
kernels + integrate) on synthetic Sedov ICs, BASELINE.json's metric.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--side S] [--dtype f64|mixed] [--impl reference]

N = 1: Sedov 200^3 (8.0 M particles, BASELINE.json configs[1]) on one B200.
N > 1: launched by torchrun, one rank per GPU; the particle set is partitioned by SFC key ranges
       by the C++ multi-rank domain (sphb200_domain_*, sph-exa_b200/domain.py); weak scaling: 200^3 particles per GPU
       (global lattice side = round(200 * N^(1/3)), made even).
--impl reference: the reference's own `sphexa` CPU binary (baseline/_ref/sphexa, built from its sources where they
       lie) on all host cores of this box, Sedov at --ref-side (200).
One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "SPH particle-steps/sec (Sedov, SPH-VE full timestep: domain sync + neighbors + hydro + integrate)"
UNIT = "particle-steps/s"

# algorithmic bytes per particle per launch (compulsory field traffic), f64 / mixed
ALGO_BYTES = {
    "domain::sync": (464, 384), "FindNeighbors": (32 + 4, 28 + 4), "XMass": (60, 44),
    "Normalization & Gradh": (64, 44), "EquationOfState": (56, 32), "IadVelocityDivCurl": (136, 80),
    "AVswitches": (152, 88), "MomentumAndEnergy": (184, 108), "Timestep+UpdateQuantities": (204, 140),  # f64: 108 B read + 96 B written (positionUpdate rebuilds v from dx / dt_m1)
}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_peaks():
    """DFMA rate and record-gather cost measured on a B200 of this pool by profiles/microbench/peaks.cu"""
    p = os.path.join(ROOT, "profiles", "peaks_r02.json")
    if not os.path.exists(p):
        return None
    with open(p) as f:
        return json.load(f)


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            parts = [p.strip() for p in r.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


SPHEXA_BIN = os.path.join(ROOT, "baseline", "_ref", "sphexa")


SPHEXA_CUDA_BIN = os.path.join(ROOT, "baseline", "_ref", "sphexa-cuda")


def reference_gpu(side, phase_ms_ours, ms_ours, mixed_ours=None, iterations=10):
    """BASELINE.md section 3b, the second on-box baseline: the reference's own GPU build (`sphexa-cuda`, its generic CUDA
    kernels recompiled for sm_100a by oracle/build_sphexa_cuda.sh) on the same workload on this GPU.  Test
    infrastructure, never linked into the product.  The reference has one precision (mixed: float hydro fields)."""
    if not os.path.exists(SPHEXA_CUDA_BIN):
        return {"unavailable": "baseline/_ref/sphexa-cuda was not built (oracle/build_sphexa_cuda.sh needs /root/reference)"}
    its, phases_all, nb, _ = run_sphexa(side, iterations, binary=SPHEXA_CUDA_BIN, all_phases=True)
    skip = 3
    timed = sorted(its[skip:])
    med = timed[len(timed) // 2]
    names = ["domain::sync", "FindNeighbors", "XMass", "Normalization & Gradh", "EquationOfState",
             "IadVelocityDivCurl", "AVswitches", "MomentumAndEnergy"]
    ref_ms = {}
    for k in names:
        v = sorted(p.get(k, 0.0) for p in phases_all[skip:])
        ref_ms[k] = v[len(v) // 2] * 1e3
    n = side ** 3
    out = {"program": f"baseline/_ref/sphexa-cuda --init sedov -n {side} -s {iterations} --ascii (reference GPU build, "
                      f"sm_100a, mixed precision), iterations {skip + 1}..{iterations}, median",
           "ms_per_step": med * 1e3, "particle_steps_per_s": n / med,
           "ms_per_step_min_max": [timed[0] * 1e3, timed[-1] * 1e3],
           "phase_ms": ref_ms,
           "ours_fp64": {"ms_per_step": ms_ours, "speedup_vs_reference_gpu": med * 1e3 / ms_ours,
                         "phase_speedup": {k: ref_ms[k] / phase_ms_ours[k] for k in names
                                           if phase_ms_ours.get(k) and ref_ms.get(k)}},
           "note": "the reference traverses the tree inside every pair kernel (FindNeighbors is only its first "
                   "traversal); ours searches once and stores the lists"}
    if mixed_ours:
        out["ours_mixed"] = {"ms_per_step": mixed_ours, "speedup_vs_reference_gpu": med * 1e3 / mixed_ours}
    # the same reference front-end (its Domain, propagator, initial conditions) with the five SPH-VE pair operators
    # forwarded to this library (sph-exa_b200/host/b200_forward.cu): the operator-level drop-in, timed by the
    # reference's own timers
    dropin = os.path.join(ROOT, "baseline", "_ref", "sphexa-cuda-b200")
    if os.path.exists(dropin):
        its2, ph2, _, _ = run_sphexa(side, iterations, binary=dropin, all_phases=True)
        t2 = sorted(its2[skip:])
        med2 = t2[len(t2) // 2]
        d_ms = {}
        for k in names:
            v = sorted(p.get(k, 0.0) for p in ph2[skip:])
            d_ms[k] = v[len(v) // 2] * 1e3
        out["drop_in"] = {"program": "baseline/_ref/sphexa-cuda-b200: reference front-end + cstone::Domain + HydroVeProp, "
                                     "pair operators, focus-tree decisions and Cornerstone keys / sort / octree / halo-search / gather forwarded to libsphexa_b200.so (mixed precision)",
                          "ms_per_step": med2 * 1e3, "particle_steps_per_s": n / med2, "phase_ms": d_ms,
                          "speedup_vs_reference_gpu": med / med2,
                          "note": "XMass includes the one neighbor search all five operators share"}
    return out


def run_sphexa(side, iterations, timeout=1500, binary=None, all_phases=False):
    """the reference's own CPU front-end (main/src/sphexa/sphexa.cpp, built by oracle/build_sphexa.sh into
    baseline/_ref/sphexa): `sphexa --init sedov -n <side> -s <iterations> --ascii` on all host threads.  Returns
    (seconds per iteration [list], phase seconds of the last iteration, neighbors per particle, threads)."""
    import re
    import tempfile
    threads = host_threads()
    env = dict(os.environ)
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm gets every core of the box
    env.update(OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="close", OMP_PLACES="cores")
    with tempfile.TemporaryDirectory() as tmp:
        res = subprocess.run([binary or SPHEXA_BIN, "--init", "sedov", "-n", str(side), "-s", str(iterations), "--ascii"],
                             cwd=tmp,
                             env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    if res.returncode != 0:
        raise RuntimeError(f"sphexa exited with {res.returncode}: {res.stdout[-400:]}")
    it_s, phases, cur, nb, every = [], {}, {}, None, []
    for line in res.stdout.splitlines():
        m = re.match(r"# ([^:]+(?:::[^:]+)?): ([0-9.eE+-]+)s", line)
        if m and not line.startswith("# Total execution"):
            cur[m.group(1)] = cur.get(m.group(1), 0.0) + float(m.group(2))
        m = re.match(r"=== Total time for iteration\((\d+)\) ([0-9.eE+-]+)s", line)
        if m:
            it_s.append(float(m.group(2)))
            every.append(cur)
            phases, cur = cur, {}
        m = re.search(r"Avg neighbor count per particle: ([0-9.]+)", line)
        if m:
            nb = float(m.group(1))
    if len(it_s) != iterations:
        raise RuntimeError(f"sphexa printed {len(it_s)} iterations, expected {iterations}")
    return it_s, (every if all_phases else phases), nb, threads


def reference_arm(args):
    """the reference's CPU implementation of the path on this box's host cores, all threads: the reference's own
    `sphexa` binary on the arm's configuration (Sedov, --side^3 particles, full step).  No repo kernel, no repo module
    and no Python inside the timed region: the times are the ones the reference prints per iteration."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    side = args.ref_side
    n = side ** 3
    if os.path.exists(SPHEXA_BIN):
        its, phases, nb, cores = run_sphexa(side, args.warmup + args.steps)
        timed = its[args.warmup:]
        dt = sum(timed)
        value = n * len(timed) / dt
        how = (f"baseline/_ref/sphexa --init sedov -n {side} -s {args.warmup + args.steps} --ascii (the reference's own "
               f"front-end and CPU operators, g++ -O3 -fopenmp, OMP_PROC_BIND=close OMP_PLACES=cores), iterations "
               f"{args.warmup + 1}..{args.warmup + args.steps} as printed by the program; the reference has one "
               f"precision, mixed (HydroType = float, sph/types.hpp:39-46): its fp64 equivalent would be slower")
        phase_ms = {k: v * 1e3 for k, v in phases.items()}
        nb_ms = phase_ms.get("FindNeighbors")
        program = "sphexa (reference front-end)"
    else:
        # the binary did not travel: the reference's operators behind the veneer of oracle/ref_pipeline.py
        import numpy as np
        os.environ.setdefault("OMP_PROC_BIND", "close")
        os.environ.setdefault("OMP_PLACES", "cores")
        from oracle import ref, ic_ref
        from oracle.ref_pipeline import RefSim
        ref.set_num_threads(host_threads())
        f, box, s = ic_ref.sedov(side)
        sim = RefSim(f, box, s, real=np.float64)
        for _ in range(args.warmup):
            sim.step()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            sim.step()
        dt = time.perf_counter() - t0
        value = n * args.steps / dt
        cores = ref.num_threads()
        phase_ms = {k: v * 1e3 for k, v in sim.times.items()}
        nb_ms = phase_ms.get("FindNeighbors")
        how = f"oracle/_ref operators sequenced by oracle/ref_pipeline.py, fp64, Sedov {side}^3"
        program = "oracle/_ref veneer"
    same = side == args.side
    sample = (f"Sedov {side}^3 ({n} particles)" + ("" if same else f" instead of {args.side}^3") + f": {how}")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "mixed (reference default)" if program.startswith("sphexa") else "f64",
        "data": "synthetic",
        "config": {"workload": f"Sedov blast wave {side}^3 ({n} particles) SPH-VE, full step, reference CPU build on "
                               f"{cores} host threads", "l2": "n/a (CPU)", "program": program,
                   "same_size_as_gpu_arm": same},
        "find_neighbors_ms": nb_ms, "phase_ms": phase_ms,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline(dtype):
    """bounded CPU sample for the default run's JSON line (rank 0, N=1): the reference's own binary on 100^3 particles
    (about 10-20 s of CPU work), plus BASELINE.json configs[0] literally (`sphexa --init sedov -n 50 -s 10`)"""
    if not os.path.exists(SPHEXA_BIN):
        raise RuntimeError("baseline/_ref/sphexa is missing (oracle/build_sphexa.sh builds it where /root/reference exists)")
    its, phases, nb, cores = run_sphexa(100, 4)
    n = 100 ** 3
    value = n * 3 / sum(its[1:])
    its0, _, _, _ = run_sphexa(50, 10)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "iterations 2..4 of `baseline/_ref/sphexa --init sedov -n 100 -s 4 --ascii` (1.0 M particles; the "
                      "reference's own front-end and CPU operators, mixed precision, g++ -O3 -fopenmp, all host "
                      "threads, OMP_PROC_BIND=close OMP_PLACES=cores)",
            "find_neighbors_ms": phases.get("FindNeighbors", 0.0) * 1e3,
            "phase_ms": {k: v * 1e3 for k, v in phases.items()},
            "configs0": {"command": "sphexa --init sedov -n 50 -s 10 --ascii", "particles": 125000,
                         "total_s": sum(its0), "iteration_s": its0,
                         "particle_steps_per_s": 125000 * 9 / sum(its0[1:])}}


NCU_SUMMARY = "ncu_full_r02c_summary.csv"  # replaced by this round's capture when it exists
NCU_KERNEL = {"MomentumAndEnergy": "momentumEnergyKernel", "AVswitches": "avSwitchesKernel",
              "IadVelocityDivCurl": "iadDivvCurlvKernel", "FindNeighbors": "findNeighborsKernel",
              "XMass": "xmassKernel", "Normalization & Gradh": "veDefGradhKernel"}


def ncu_traffic(phase, args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same workload (profiles/ncu_full_r01h_summary.csv: Sedov 200^3, f64).
    None for any other workload: the capture does not transfer."""
    import csv
    path = os.path.join(ROOT, "profiles", NCU_SUMMARY)
    if args.side != 200 or args.dtype != "f64" or phase not in NCU_KERNEL or not os.path.exists(path):
        return None, None
    with open(path) as fh:
        rows = list(csv.reader(fh))
    hdr = rows[0]
    ir, iw = hdr.index("dram_read [Gbyte]"), hdr.index("dram_write [Gbyte]")
    for r in rows[1:]:
        if NCU_KERNEL[phase] in r[0]:
            return (float(r[ir]) + float(r[iw])) * 1e9, "profiles/" + NCU_SUMMARY
    return None, None


def ncu_column(phase, args, column):
    """one column of the committed ncu summary for the dominant kernel (None for other workloads)"""
    import csv
    path = os.path.join(ROOT, "profiles", NCU_SUMMARY)
    if args.side != 200 or args.dtype != "f64" or phase not in NCU_KERNEL or not os.path.exists(path):
        return None
    with open(path) as fh:
        rows = list(csv.reader(fh))
    if column not in rows[0]:
        return None
    for r in rows[1:]:
        if NCU_KERNEL[phase] in r[0]:
            return float(r[rows[0].index(column)])
    return None


def other_config(pkg, ctx, S, torch, case, side, what, dtype, dtype_name, steps=5, warmup=3):
    """a few steps of one of the other BASELINE.json configs on one GPU: step time, phase times, neighbor-count
    spread, leaf-array changes per step (the octree rebalancing the KH config is there to exercise)"""
    import numpy as np
    f, box, s = pkg.ic.CASES[case](side, dtype_name)
    n = f["x"].size
    gs = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    turb = None
    if case == "turbulence":
        from sph_exa_b200.turbulence import Turbulence
        turb = Turbulence()
        gs.attachTurbulence(turb)
    gs.upload(f)
    for _ in range(warmup):
        gs.step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        gs.step()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    phase, changed = {}, []
    prev = gs.leaves()
    for _ in range(3):
        gs.step()
        for k, v in gs.phaseTimes().items():
            phase[k] = phase.get(k, 0.0) + v / 3
        cur = gs.leaves()
        changed.append(int(np.setxor1d(prev, cur).size))
        prev = cur
    nc = gs.download(names=["nc"])["nc"].astype(np.int64) - 1
    out = {"workload": what, "particles": int(n), "ms_per_step": ms, "particle_steps_per_s": n / (ms * 1e-3),
           "phase_ms": phase, "leaves": int(gs.numLeaves), "leaf_keys_changed_per_step": changed,
           "neighbors": {"min": int(nc.min()), "mean": float(nc.mean()), "max": int(nc.max()),
                         "std": float(nc.std())}, "box": gs.boxLimits()}
    gs.close()
    if turb is not None:
        turb.close()
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--side", type=int, default=200, help="lattice side per GPU (200 -> 8.0 M particles)")
    ap.add_argument("--dtype", default="f64", choices=["f64", "mixed"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ref-side", type=int, default=200, help="lattice side of the CPU reference arm (the arm's own 200)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-bdt", action="store_true", help="skip the block-time-step (ve-bdt) hierarchy report")
    ap.add_argument("--no-configs", action="store_true", help="skip the Noh / turbulence / KH reports")
    ap.add_argument("--no-reference-gpu", action="store_true", help="skip the run of the reference's own GPU build")
    ap.add_argument("--case", default="sedov", choices=["sedov", "turbulence", "noh"],
                    help="--gpus > 1 only: turbulence = BASELINE.json configs[3] (stirred, weak scaling); the headline "
                         "metric is the default")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else max(args.warmup, 1)

    if args.impl == "reference":
        return reference_arm(args)

    import numpy as np
    import torch
    import sphexa_b200_loader
    pkg = sphexa_b200_loader.load()
    from sph_exa_b200 import sim as S, binding as B

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 or world > 1:
        from sph_exa_b200 import domain as D
        return D.bench_main(args, pkg, METRIC, UNIT, ALGO_BYTES, load_peaks, ClockSampler)

    torch.cuda.set_device(local_rank)
    # the host-resident step copies through pinned buffers: allocate them on the GPU's own NUMA node (restored before
    # the CPU baselines, which get every core)
    all_cpus = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    from sph_exa_b200.domain import bind_to_gpu_numa_node
    bind_to_gpu_numa_node(local_rank)
    dtype = B.F64 if args.dtype == "f64" else B.MIXED
    f, box, s = pkg.ic.sedov(args.side, args.dtype)
    n = f["x"].size
    ctx = pkg.Context(local_rank)
    gs = S.HydroVeProp(ctx, box, s, capacity=n, dtype=dtype)
    gs.upload(f)

    def sync():
        torch.cuda.synchronize()

    # ---------------- device-resident throughput ("value")
    for _ in range(args.warmup):
        gs.step()
    sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = ctx.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    phase_ms = {}
    e0.record()
    for _ in range(args.steps):
        gs.step()
    e1.record()
    sync()
    total_ms = e0.elapsed_time(e1)
    launches = ctx.launches - launches0
    clocks = sampler.stop()
    # per-phase device times (cudaEvents recorded by the library on its launch stream), averaged over extra steps
    reps = 3
    for _ in range(reps):
        gs.step()
        for k, v in gs.phaseTimes().items():
            phase_ms[k] = phase_ms.get(k, 0.0) + v / reps
    ms_per_step = total_ms / args.steps
    value = n * args.steps / (total_ms * 1e-3)
    total_nb = gs.totalNeighbors()
    num_leaves = gs.numLeaves

    # ---------------- roofline of the dominant kernel
    # The pair kernels are gathers: one 32-byte record per (pair, record type) through the L1 data stage, which is what
    # saturates (ncu: l1tex data-stage wavefronts 90-95 % of peak, DRAM 3-16 %), then the fp64 pipe.  Their roof is
    # therefore the MEASURED gather rate and the MEASURED DFMA rate of this GPU (profiles/microbench/peaks.cu ->
    # profiles/peaks_r02.json), not HBM; the HBM roofline is quoted for the HBM-bound kernels (hbm_kernels below) and for
    # the step as a whole.
    peak, peak_src = load_peaks()
    mp = measured_peaks()
    kernels = {k: v for k, v in phase_ms.items() if k not in ("domain::sync",)}
    dom = max(kernels, key=kernels.get)
    col = 0 if args.dtype == "f64" else 1
    algo = ALGO_BYTES[dom][col] * n
    achieved_hbm = algo / (phase_ms[dom] * 1e-3) / 1e9
    traffic, traffic_src = ncu_traffic(dom, args)
    pair_flop = {"XMass": 22, "Normalization & Gradh": 36, "IadVelocityDivCurl": 96, "AVswitches": 58,
                 "MomentumAndEnergy": 150}
    # 32-byte (f64) / 16-byte (mixed) records gathered per neighbor pair
    pair_records = {"XMass": 1, "Normalization & Gradh": 1, "IadVelocityDivCurl": 2, "AVswitches": 2,
                    "MomentumAndEnergy": 5}
    pairs = float(total_nb) - n  # nc counts the particle itself
    if dom in pair_records and mp is not None and args.dtype == "f64":
        rec_per_s = pair_records[dom] * pairs / (phase_ms[dom] * 1e-3)
        cyc = mp["gather_32B_record"]["lattice_neighbor_lists"]["ldg256_aos"]
        sm_hz = (clocks.get("sm_mhz") or mp["sm_clock_max_mhz"]) * 1e6
        rec_peak = mp["sms"] * sm_hz / cyc * 32
        tf = pair_flop[dom] * pairs / (phase_ms[dom] * 1e-3) / 1e12
        roofline = {"bound": "l1_gather", "kernel": dom, "achieved": rec_per_s / 1e9, "peak": rec_peak / 1e9,
                    "unit": "G records/s (32-byte gather records through the L1 data stage)",
                    "frac": rec_per_s / rec_peak,
                    "peak_source": f"measured: profiles/peaks_r02.json, {cyc} SM cycles per warp-level 32-byte record "
                                   f"gather over real neighbor lists, {mp['sms']} SMs at the sampled clock",
                    "traffic": traffic, "traffic_source": traffic_src,
                    "fp64": {"achieved_tflops": tf, "peak_tflops": mp["dfma_tflops"], "frac": tf / mp["dfma_tflops"],
                             "peak_source": "measured DFMA rate (profiles/peaks_r02.json)",
                             "flop_per_pair": pair_flop[dom]},
                    "hbm": {"achieved": achieved_hbm, "peak": peak, "unit": "GB/s", "frac": achieved_hbm / peak,
                            "peak_source": peak_src, "algorithmic_bytes_per_launch": algo,
                            "note": "algorithmic field bytes only: this kernel is not HBM-bound"},
                    "l1_wavefront_pct_of_peak_ncu": ncu_column(dom, args, "l1_data_wavefronts_pct [%]"),
                    "fp64_pipe_pct_ncu": ncu_column(dom, args, "fp64_pipe_pct [%]"),
                    "kernel_ms": phase_ms[dom], "share_of_step": phase_ms[dom] / sum(phase_ms.values())}
    else:
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved_hbm, "peak": peak, "unit": "GB/s",
                    "frac": achieved_hbm / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": algo, "kernel_ms": phase_ms[dom],
                    "share_of_step": phase_ms[dom] / sum(phase_ms.values())}
    # whole-step algorithmic traffic (1360 B f64 / 920 B mixed per particle-step) against the HBM peak
    step_bytes = (1360 if args.dtype == "f64" else 920) * n
    roofline["step_achieved_gbs"] = step_bytes / (ms_per_step * 1e-3) / 1e9
    roofline["step_frac"] = roofline["step_achieved_gbs"] / peak

    # ---------------- the HBM-bound kernels of the step, one by one (north star: >= 70 % of the HBM roofline on the
    # gather kernels): the operators of Domain::sync replayed through the tier-1 ABI on the current particle state,
    # CUDA events on the launching stream, inputs (>= 192 MB each) larger than L2
    hbm_kernels = None
    try:
        from sph_exa_b200 import ops
        real_t = torch.float64 if args.dtype == "f64" else torch.float32
        state = gs.download()
        dev = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in state.items()}
        bx = pkg.make_box(*box)
        ev = lambda: torch.cuda.Event(enable_timing=True)

        def timed(fn, reps=5):
            fn()
            sync()
            a, b = ev(), ev()
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            sync()
            return a.elapsed_time(b) / reps

        keys = ops.computeSfcKeysGpu(ctx, dev["x"], dev["y"], dev["z"], bx)
        t_keys = timed(lambda: ops.computeSfcKeysGpu(ctx, dev["x"], dev["y"], dev["z"], bx, keys))
        # the state of a running simulation: SFC-sorted one step ago, every particle moved by v dt since (what the
        # sort and the gather of the next Domain::sync see)
        keys0 = keys.clone()
        work = keys0.clone()
        perm = ops.sortByKeyGpu(ctx, work)

        def sort_once():
            work.copy_(keys0)
            ops.sortByKeyGpu(ctx, work)

        t_copy = timed(lambda: work.copy_(keys0))
        t_sort = timed(sort_once) - t_copy
        srcs = [dev[k] for k in B.CONSERVED]
        dsts = [torch.empty_like(a) for a in srcs]
        t_gather = timed(lambda: ops.gatherGpu(ctx, perm, srcs, dsts))
        # halo pack / unpack (gatherRanges / scatterRanges, halos/gather_halos_gpu.cu:41-117): the x,y,z,h,m exchange of
        # a 200^3 rank among 8, i.e. ~10 % of the particles in ~4000 ranges of ~200 particles, 5 fields
        eb = 8 if args.dtype == "f64" else 4
        n_ranges, per_range = 4096, 200
        total_h = n_ranges * per_range
        begins = (torch.arange(n_ranges, device="cuda", dtype=torch.int64) * (n // n_ranges)).int()
        scan = (torch.arange(n_ranges, device="cuda", dtype=torch.int64) * per_range).int()
        hsrc = [dev[k] for k in ("x", "y", "z", "h", "m")]
        packed = torch.empty(ops.packedBytes(total_h, hsrc), dtype=torch.uint8, device="cuda")
        # one launch moves 0.8 M particles only: repeat over shifted ranges so that the sources do not sit in L2
        shifts = [begins + 250 * j for j in range(8)]
        t_pack = timed(lambda: [ops.packRanges(ctx, scan, b_, total_h, hsrc, packed) for b_ in shifts]) / len(shifts)
        t_unpack = timed(lambda: [ops.unpackRanges(ctx, scan, b_, total_h, hsrc, packed) for b_ in shifts]) / len(shifts)
        halo_bpp = 2 * (3 * 8 + 2 * eb)  # read + write per halo particle
        # the observable reduction that runs every step (conserved_quantities.hpp:113-180): 9 values per particle
        t_cons = timed(lambda: gs.conserved())
        cons_bpp = 4 * 8 + 4 * eb + 4  # x, y, z, temp (f64) + m, vx, vy, vz + nc
        col = 0 if args.dtype == "f64" else 1
        rows = {"sfcKeys": (t_keys, 32), "radixSort (8 passes)": (t_sort, 204),
                "haloPack (5 fields, 4096 ranges)": (t_pack * n / total_h, halo_bpp),
                "haloUnpack (5 fields, 4096 ranges)": (t_unpack * n / total_h, halo_bpp),
                "conservedQuantities (reduction)": (t_cons, cons_bpp),
                "gatherFields (14 fields)": (t_gather, (228, 148)[col]),
                "eos": (phase_ms["EquationOfState"], ALGO_BYTES["EquationOfState"][col]),
                "positions+updateH": (phase_ms["Timestep+UpdateQuantities"], ALGO_BYTES["Timestep+UpdateQuantities"][col])}
        hbm_kernels = {k: {"ms": ms, "algorithmic_bytes_per_particle": bpp, "achieved_gbs": bpp * n / (ms * 1e-3) / 1e9,
                           "frac": bpp * n / (ms * 1e-3) / 1e9 / peak} for k, (ms, bpp) in rows.items()}
        del dev, srcs, dsts, keys0, work, perm, keys
        torch.cuda.empty_cache()
    except Exception as ex:  # diagnostics only: never let them take the headline down
        hbm_kernels = {"error": str(ex)}

    # ---------------- end to end through the C ABI with host buffers
    e2e = None
    if not args.no_e2e:
        real = np.float64 if args.dtype == "f64" else np.float32
        host_in, host_out = {}, {}
        state = gs.download()
        for k in B.CONSERVED:
            want = np.float64 if k in B.ALWAYS_F64 else real
            tin = torch.empty(n, dtype=torch.float64 if want == np.float64 else torch.float32).pin_memory()
            tout = torch.empty_like(tin).pin_memory()
            tin.numpy()[:] = state[k]
            host_in[k], host_out[k] = tin, tout
        fin, fout = B.Fields(), B.Fields()
        for k in B.FIELD_NAMES:
            setattr(fin, k, host_in[k].data_ptr() if k in host_in else None)
            setattr(fout, k, host_out[k].data_ptr() if k in host_out else None)
        h2d = sum(t.numel() * t.element_size() for t in host_in.values())
        d2h = sum(t.numel() * t.element_size() for t in host_out.values())
        e2e_steps = max(2, min(args.steps, 5))
        e2e_warm = 3  # untimed: first-touch of the pinned buffers, the timed choice of the kernel-table variants
        for it in range(e2e_warm + e2e_steps):
            if it == e2e_warm:
                sync()
                t0 = time.perf_counter()
            # H2D of this step's inputs from pinned memory, the step, D2H of the new state (synchronises); the
            # library stages the copies beside the kernels (sphb200_sim_step_host)
            gs.step_host(fin, fout, n)
            for k in host_in:            # next step's input = this step's output
                host_in[k], host_out[k] = host_out[k], host_in[k]
            for k in B.FIELD_NAMES:
                setattr(fin, k, host_in[k].data_ptr() if k in host_in else None)
                setattr(fout, k, host_out[k].data_ptr() if k in host_out else None)
        sync()
        dt = time.perf_counter() - t0
        e2e = {"value": n * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": dt / e2e_steps * 1e3, "steps": e2e_steps}

    if all_cpus is not None:
        os.sched_setaffinity(0, all_cpus)
    cpu = None
    if not args.no_cpu_baseline:
        try:
            cpu = cpu_baseline(args.dtype)
        except Exception as ex:  # the checker library is absent: report it, never substitute
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": f"unavailable: {ex}"}

    # ---------------- block time steps (HydroVeBdtProp): three hierarchies on the same IC, the last one
    # reported beside the headline (the headline itself stays the global-time-step propagator BASELINE.json names); the
    # first two are warm-up (the first full syncs also time the two kernel-table variants of the pair operators)
    bdt = None
    if not args.no_bdt:
        gs.close()
        gs = S.HydroVeBdtProp(ctx, box, s, capacity=n, dtype=dtype)
        gs.upload(f)
        sub = []
        hierarchies = 0
        while hierarchies < 3 and len(sub) < 60:
            full = gs.isSynced()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            gs.step()
            b.record()
            sync()
            info = gs.timestep()
            sub.append({"full_sync": full, "ms": a.elapsed_time(b), "active_groups": info.numActiveGroups,
                        "groups": info.numGroups, "rungs": info.numRungs, "dt": info.nextDt})
            hierarchies += bool(info.isSynced)
        last = [x for x in sub if x["rungs"] == sub[-1]["rungs"]][-(1 << (sub[-1]["rungs"] - 1)):]
        part = [x for x in last if not x["full_sync"]]
        hier_ms = sum(x["ms"] for x in last)
        bdt = {"propagator": "ve-bdt", "rungs": last[-1]["rungs"], "substeps_per_hierarchy": len(last),
               "rung_ranges": list(gs.timestep().rungRanges), "groups": last[-1]["groups"],
               "full_sync_substep_ms": next((x["ms"] for x in last if x["full_sync"]), None),
               "partial_substep_ms": (sum(x["ms"] for x in part) / len(part)) if part else None,
               "active_group_fraction_partial": (sum(x["active_groups"] for x in part) / len(part) / last[-1]["groups"])
               if part else None,
               "hierarchy_ms": hier_ms,
               # the same simulated time with the global-time-step propagator: one full step per substep
               "global_dt_equivalent_ms": len(last) * ms_per_step,
               "speedup_vs_global_dt": len(last) * ms_per_step / hier_ms if hier_ms > 0 else None}

    # ---------------- the other BASELINE.json configs at their per-GPU size (parity-test cases in tests/; here only how
    # the same path behaves on them: open box, stirred turbulence, density contrast).  Not bench lines.
    other = None
    if not args.no_configs:
        gs.close()
        gs = None
        torch.cuda.empty_cache()
        other = {}
        for name, side, what in (("noh", 150, "Noh implosion, per-GPU share of configs[2] (300^3 lattice cut to the "
                                              "sphere over 8 GPUs), open box: bounding-box reduction every step"),
                                 ("turbulence", 250, "configs[3] per GPU: 250^3, gamma 1.001, turbulence-ve "
                                                     "(global dt) with the stirring driver attached"),
                                 ("kh", 160, "configs[4]: Kelvin-Helmholtz slab, 2:1 density contrast, alpha = 1")):
            try:
                other[name] = other_config(pkg, ctx, S, torch, name, side, what, dtype, args.dtype)
            except Exception as ex:  # diagnostics only
                other[name] = {"error": str(ex)}
        if args.dtype == "f64":
            # the reference's default build is mixed precision (float fields, double coordinates and energy)
            try:
                other["sedov_mixed_precision"] = other_config(
                    pkg, ctx, S, torch, "sedov", args.side, "the headline workload in the reference's default mixed "
                    "precision (HydroType = float): 16-byte gather records", B.MIXED, "mixed")
            except Exception as ex:
                other["sedov_mixed_precision"] = {"error": str(ex)}

    ref_gpu = None
    if not args.no_reference_gpu:
        try:
            if gs is not None:
                gs.close()
                gs = None
            torch.cuda.empty_cache()
            mixed = (other or {}).get("sedov_mixed_precision", {}).get("ms_per_step") if isinstance(other, dict) else None
            ref_gpu = reference_gpu(args.side, phase_ms, ms_per_step, mixed)
        except Exception as ex:  # diagnostics only
            ref_gpu = {"error": str(ex)[-300:]}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic",
        "config": {"workload": f"Sedov blast wave {args.side}^3 ({n} particles) SPH-VE on 1 B200, full step",
                   "l2": "inputs larger than L2 (particle state >= 2.5 GB, neighbor lists 4.8 GB)",
                   "ng0": 100, "ngmax": 150, "bucket_focus": 64},
        "find_neighbors_ms": phase_ms.get("FindNeighbors"),
        "phase_ms": phase_ms, "total_neighbors": total_nb, "leaves": num_leaves,
        "clocks": clocks, "gpu_launches": launches, "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
        "hbm_kernels": hbm_kernels,
        "hbm_fracs": ({k: round(v["frac"], 3) for k, v in hbm_kernels.items() if isinstance(v, dict) and "frac" in v}
                      if isinstance(hbm_kernels, dict) else None),
        "block_time_steps": bdt, "other_configs": other, "reference_gpu": ref_gpu,
    }
    print(json.dumps(line), flush=True)
    if gs is not None:
        gs.close()
    ctx.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
