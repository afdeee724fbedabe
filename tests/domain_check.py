"""NCCL parity driver of the C++ multi-rank domain (not a pytest file; run under torchrun on a box with >= 2 GPUs):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
      tests/domain_check.py --case sedov --side 40 --steps 3

Every rank advances its SFC partition through sphb200_domain_* with the NCCL transport and compares its owned
particles with the matching key range of the reference's single-rank CPU run (oracle/_ref, RefSim): keys and neighbor
counts exactly, fields <= 1e-10.  Also: the rank's MAC-coarsened focus tree == cstone::FocusedOctreeSingleNode on the
global particle set, and sphb200_domain_step_host == upload + step + download.
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import sphexa_b200_loader  # noqa: E402

FIELDS = ["x", "y", "z", "h", "m", "vx", "vy", "vz", "temp", "alpha", "xm", "kx", "gradh", "prho", "c", "c11", "c22",
          "c33", "divv", "curlv", "ax", "ay", "az", "du"]
VEC = {"ax": ("ax", "ay", "az"), "ay": ("ax", "ay", "az"), "az": ("ax", "ay", "az"), "vx": ("vx", "vy", "vz"),
       "vy": ("vx", "vy", "vz"), "vz": ("vx", "vy", "vz")}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="sedov")
    ap.add_argument("--side", type=int, default=40)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    pkg = sphexa_b200_loader.load()
    from sph_exa_b200 import domain as D, binding as B
    from oracle.ref_pipeline import RefSim

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    f, box, s = pkg.ic.CASES[args.case](args.side)
    n = f["x"].size
    s = dict(s)
    s["numParticlesGlobal"] = n
    first, last = (n * rank) // world, (n * (rank + 1)) // world
    ctx = pkg.Context(local)
    comm = D.nccl_comm(rank, world, local)
    cap = int(3.0 * n / world) + 4096
    sim = D.DomainRank(ctx, box, s, cap, comm, rank, bucket_size=max(64, n // (100 * world)))
    sim.upload({k: v[first:last] for k, v in f.items()})
    group = D.DistHydroVe([sim])
    rs = RefSim(f, box, s)
    ok = True
    for step in range(args.steps):
        group.computeForces()
        rs.computeForces()
        out = sim.download(names=FIELDS + ["nc"], with_keys=True)
        a = int(np.searchsorted(rs.keys, out["keys"][0], side="left")) if out["keys"].size else 0
        b = a + out["keys"].size
        if not np.array_equal(out["keys"], rs.keys[a:b]):
            ok = False
            print(f"[rank {rank}] step {step}: keys differ", flush=True)
        if not np.array_equal(out["nc"], rs.nc[a:b]):
            ok = False
            print(f"[rank {rank}] step {step}: neighbor counts differ", flush=True)
        for k in FIELDS:
            scale = max(max(np.abs(rs.f[g]).max() for g in VEC.get(k, (k,))), 1e-300)
            err = np.abs(out[k] - rs.f[k][a:b]).max() / scale if out[k].size else 0.0
            if not err < 1e-10:
                ok = False
                print(f"[rank {rank}] step {step}: {k} rel err {err:.3e}", flush=True)
        group.integrate()
        rs.integrate()
        if abs(sim.minDt - rs.minDt) > 1e-12 * rs.minDt:
            ok = False
            print(f"[rank {rank}] step {step}: dt {sim.minDt} vs {rs.minDt}", flush=True)
    cnt = torch.tensor([sim.endIndex - sim.startIndex], dtype=torch.int64, device="cuda")
    dist.all_reduce(cnt)
    if int(cnt.item()) != n:
        ok = False
        print(f"[rank {rank}] particles lost: {int(cnt.item())} != {n}", flush=True)

    # the reference-style focus tree of this rank (sphb200_domain_update_focus: all-gather + all-reduce over NCCL)
    # == cstone::FocusedOctreeSingleNode converged on the global particle set for this rank's segment and peers
    from oracle import ref
    from sph_exa_b200 import focus as F
    theta = 1.0
    group.sync()
    if not group.updateFocus(theta):
        ok = False
        print(f"[rank {rank}] focus trees not converged after the first update", flush=True)
    leaves, counts, peers = sim.focusTree()
    parts = [None] * world
    dist.all_gather_object(parts, sim.download(names=["x"], with_keys=True)["keys"])
    gkeys = np.concatenate(parts)
    info = sim.info()
    bounds = np.array(info.assignment[:world + 1], np.uint64)
    gbox = (sim.boxLimits(), list(box[1]))
    want_peers = ref.find_peers_mac(rank, bounds, sim.globalLeaves(), gbox, F.invThetaMinMac(theta))
    mand = np.array([bounds[p + d] for p in peers for d in (0, 1)], np.uint64)
    theirs = ref.FocusedOctreeSingleNode(64, theta)
    for _ in range(40):
        if theirs.update(gbox, gkeys, bounds[rank], bounds[rank + 1], mand):
            break
    wl, wc = theirs.leaves_counts()
    if peers != want_peers or not np.array_equal(leaves, wl) or not np.array_equal(counts, wc):
        ok = False
        print(f"[rank {rank}] focus tree differs from the reference: {leaves.size - 1} vs {wl.size - 1} leaves, peers "
              f"{peers} vs {want_peers}", flush=True)
    elif rank == 0:
        print(f"focus tree of rank 0: {leaves.size - 1} leaves (guided tree: {info.numFocusLeaves}), peers {peers}",
              flush=True)

    # host-resident step == upload + step + download
    state = sim.download()
    n_h = state["x"].size
    hin = {k: torch.empty(cap, dtype=torch.float64).pin_memory() for k in B.CONSERVED}
    hout = {k: torch.empty(cap, dtype=torch.float64).pin_memory() for k in B.CONSERVED}
    for k in B.CONSERVED:
        hin[k].numpy()[:n_h] = state[k]
    fi, fo = B.Fields(), B.Fields()
    for k in B.FIELD_NAMES:
        setattr(fi, k, hin[k].data_ptr() if k in hin else None)
        setattr(fo, k, hout[k].data_ptr() if k in hout else None)
    n_out = D.step_host(group, n_h, fi, fo, cap)
    rs.step()
    keys_ref = rs.keys  # order of the step just taken
    got = {k: hout[k].numpy()[:n_out].copy() for k in ("x", "vx", "temp", "h")}
    now = sim.download(names=["x", "vx", "temp", "h"])
    for k in got:
        if not np.array_equal(got[k], now[k]):
            ok = False
            print(f"[rank {rank}] step_host: {k} differs from the device state", flush=True)
    flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        info = sim.info()
        print(f"ranks {world}, case {args.case} {args.side}^3: focus leaves {info.numFocusLeaves}, halo boxes "
              f"{info.numSurfaceBoxes}, halos {info.nWithHalos - (info.endIndex - info.startIndex)}", flush=True)
        print("DOMAIN_CHECK PASS" if int(flag.item()) == 1 else "DOMAIN_CHECK FAIL", flush=True)
    group.close()
    comm.close()
    dist.destroy_process_group()
    return 0 if int(flag.item()) == 1 else 1


if __name__ == "__main__":
    sys.exit(main())
