"""The reference's own GPU application against this library's propagators, iteration by iteration.

`baseline/_ref/sphexa-cuda` is the unmodified reference (main/src/sphexa/sphexa.cpp with its initial conditions,
cstone::Domain, propagators and observables, built for sm_100a by oracle/build_sphexa_cuda.sh).  Its iteration log
(`### Check ###` lines, main/src/propagator/ipropagator.hpp:100-127) carries, per iteration: the number of global and
focus tree leaves, the neighbor total, simulated time and time step, total / internal / kinetic energy.  The same
problem is stepped through `sphb200_sim_*` in the reference's mixed precision and must produce the same log.

 * `--prop ve`: BASELINE.json configs[0] literally (`--init sedov -n 50 -s 10`).
 * `--prop ve-bdt`: the block-time-step propagator, which the reference runs on GPUs only.  This is the one check of
   the SEQUENCING of HydroVeBdtProp (ve_hydro_bdt.hpp:280-420) against the reference itself: every substep's time step
   depends on which rungs were drifted, advanced and re-ranked before it.
"""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "sphexa-cuda")
NUM = r"([-+0-9.eE]+|nan|inf)"


def reference_log(side, steps, prop, tmp, extra=()):
    res = subprocess.run([REF, "--init", "sedov", "--prop", prop, "-n", str(side), "-s", str(steps), "--ascii", *extra],
                         cwd=tmp, capture_output=True, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    out = res.stdout
    rows = []
    trees = re.findall(r"### Check ### Global Tree Nodes: (\d+), Particles: (\d+), Halos: (\d+)", out)
    nbs = re.findall(r"### Check ### Total Neighbors: (\d+),", out)
    times = re.findall(rf"### Check ### Total time: {NUM}, current time-step: {NUM}", out)
    ener = re.findall(rf"### Check ### Total energy: {NUM}, \(internal: {NUM}, kinetic: {NUM},", out)
    focus = re.findall(r"### Check ### Focus Tree Nodes: (\d+),", out)
    # a block-time-step run only stops where its hierarchy is synced: it may take a few iterations more than asked for
    assert len(trees) == len(nbs) == len(times) == len(ener) == len(focus) >= steps, out[-3000:]
    for i in range(len(trees)):
        rows.append(dict(global_leaves=int(trees[i][0]), particles=int(trees[i][1]), neighbors=int(nbs[i]),
                         t=float(times[i][0]), dt=float(times[i][1]), etot=float(ener[i][0]), eint=float(ener[i][1]),
                         ecin=float(ener[i][2]), focus_leaves=int(focus[i])))
    # "# New block-TS 4 rungs, R0: 82 (4.1%) R1: 92 (4.6%) R2: 38 (1.9%) R3: 1788 (89.4%) All: 2000 (100%)"
    hierarchies = []
    for m in re.finditer(r"# New block-TS (\d+) rungs,(.*?) All: (\d+)", out):
        sizes = [int(v) for v in re.findall(r"R\d+: (\d+)", m.group(2))]
        hierarchies.append(dict(rungs=int(m.group(1)), sizes=sizes, groups=int(m.group(3))))
    return rows, hierarchies


def close(a, b, rel, floor=0.0):
    return abs(a - b) <= rel * max(abs(a), abs(b)) + floor


@pytest.mark.parametrize("av_clean", [False, True])
def test_ve_run_equals_the_reference_application(pkg, ctx, tmp_path, av_clean):
    if not os.path.exists(REF):
        pytest.skip("baseline/_ref/sphexa-cuda not built (oracle/build_sphexa_cuda.sh needs /root/reference)")
    from sph_exa_b200 import sim as S
    side, steps = 50, 10
    ref, _ = reference_log(side, steps, "ve", tmp_path, ("--avclean",) if av_clean else ())
    f, box, s = pkg.ic.sedov(side, "mixed")
    gs = S.HydroVeProp(ctx, box, s, capacity=f["x"].size, dtype=pkg.MIXED, av_clean=av_clean)
    gs.upload(f)
    t = 0.0
    for i, r in enumerate(ref):
        gs.computeForces()
        c = gs.conserved()
        gs.integrate()
        assert r["particles"] == f["x"].size
        assert gs.numGlobalLeaves == r["global_leaves"], i
        assert gs.numLeaves == r["focus_leaves"], i
        assert gs.totalNeighbors() == r["neighbors"], i
        # the log prints 6 significant digits of float-precision sums
        assert close(gs.minDt, r["dt"], 2e-5), (i, gs.minDt, r["dt"])
        assert close(t, r["t"], 2e-5, 1e-12), (i, t, r["t"])
        assert close(c["etot"], r["etot"], 2e-5) and close(c["eint"], r["eint"], 2e-5), (i, c, r)
        assert close(c["ecin"], r["ecin"], 2e-4, 1e-12), (i, c["ecin"], r["ecin"])
        t += gs.minDt
    gs.close()


def test_ve_bdt_run_equals_the_reference_application(pkg, ctx, tmp_path):
    if not os.path.exists(REF):
        pytest.skip("baseline/_ref/sphexa-cuda not built (oracle/build_sphexa_cuda.sh needs /root/reference)")
    from sph_exa_b200 import sim as S
    side, steps = 40, 30
    ref, hierarchies = reference_log(side, steps, "ve-bdt", tmp_path)
    f, box, s = pkg.ic.sedov(side, "mixed")
    # ParticlesData defaults (sph/particles_data.hpp:111-113): eps 0.005, etaAcc 0.2
    gs = S.HydroVeBdtProp(ctx, box, s, capacity=f["x"].size, dtype=pkg.MIXED, eta_acc=0.2, eps=0.005)
    gs.upload(f)
    t, rungs_seen, synced_checks = 0.0, set(), 0
    for i, r in enumerate(ref):
        synced = gs.isSynced()
        gs.computeForces()
        if synced:
            # observables are only evaluated where the hierarchy is synced (sphexa.cpp:153-156)
            c = gs.conserved()
            assert close(c["etot"], r["etot"], 2e-5) and close(c["eint"], r["eint"], 2e-5), (i, c, r)
            assert close(c["ecin"], r["ecin"], 5e-4, 1e-12), (i, c["ecin"], r["ecin"])
            assert gs.numLeaves == r["focus_leaves"], i
            synced_checks += 1
        gs.integrate()
        ts = gs.timestep()
        if synced:
            # a new hierarchy: number of rungs and groups per rung as the reference prints them
            h = hierarchies.pop(0)
            sizes = [int(ts.rungRanges[k + 1]) - int(ts.rungRanges[k]) for k in range(4)]
            assert (ts.numRungs, sizes, int(ts.numGroups)) == (h["rungs"], h["sizes"], h["groups"]), (i, h, sizes)
        rungs_seen.add(ts.numRungs)
        assert close(ts.nextDt, r["dt"], 5e-5), (i, ts.nextDt, r["dt"], ts.numRungs, ts.substep)
        assert close(t, r["t"], 5e-5, 1e-12), (i, t, r["t"])
        t += float(ts.nextDt)
    assert synced_checks >= 3 and max(rungs_seen) >= 2, (synced_checks, rungs_seen)
    gs.close()
