import os, sys, numpy as np, torch, ctypes as C
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
import sphexa_b200_loader
pkg = sphexa_b200_loader.load()
from sph_exa_b200.binding import lib, check
ctx = pkg.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
g = torch.Generator(device="cuda"); g.manual_seed(1)
def run(keys, label):
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    work = keys.clone()
    def once():
        work.copy_(keys)
        check(lib.sphb200_sort_by_key(ctx.h, C.c_void_p(work.data_ptr()), C.c_void_p(perm.data_ptr()), C.c_size_t(n)))
    once(); torch.cuda.synchronize()
    ref = torch.sort(keys.view(torch.int64), stable=True)
    ok = bool(torch.equal(work.view(torch.int64), ref.values)) and bool(torch.equal(perm.long(), ref.indices))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # time the copy alone, then copy + sort
    e0.record()
    for _ in range(10): work.copy_(keys)
    e1.record(); torch.cuda.synchronize(); tc = e0.elapsed_time(e1) / 10
    e0.record()
    for _ in range(10): once()
    e1.record(); torch.cuda.synchronize(); ts = e0.elapsed_time(e1) / 10 - tc
    print(f"{label}: n={n} sort {ts:.3f} ms = {204 * n / ts / 1e6:.0f} GB/s algorithmic, correct={ok}, mode={os.environ.get('SPHB200_SORT', 'chained')}", flush=True)
rand = torch.randint(0, 2**62, (n,), generator=g, device="cuda", dtype=torch.int64)
run(rand, "random 62-bit keys")
srt = torch.sort(rand).values
jit = srt + torch.randint(-2**40, 2**40, (n,), generator=g, device="cuda", dtype=torch.int64)
run(jit.clamp_(min=0), "nearly sorted keys")
run(torch.full((n,), 12345, device="cuda", dtype=torch.int64), "all keys equal")
