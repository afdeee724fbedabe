import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import sphexa_b200_loader as L
pkg=L.load()
from sph_exa_b200 import ops
from conftest import to_dev,to_host
ctx=pkg.Context(0)
n=64
for bits in (16,24,32,40,48,56,63):
    rng=np.random.Generator(np.random.MT19937(n+bits))
    keys=rng.integers(0,1<<bits,n,dtype=np.uint64)
    dk=to_dev(keys.copy()); perm=to_host(ops.sortByKeyGpu(ctx,dk),np.uint32); got=to_host(dk,np.uint64)
    want=np.sort(keys,kind='stable')
    print("bits",bits,np.array_equal(got,want))
for byte in range(8):
    rng=np.random.Generator(np.random.MT19937(byte))
    keys=rng.integers(0,256,n,dtype=np.uint64)<<np.uint64(8*byte)
    keys &= np.uint64((1<<63)-1)
    dk=to_dev(keys.copy()); perm=to_host(ops.sortByKeyGpu(ctx,dk),np.uint32); got=to_host(dk,np.uint64)
    want=np.sort(keys,kind='stable')
    print("byte",byte,np.array_equal(got,want), np.array_equal(perm,np.argsort(keys,kind='stable')))
    if not np.array_equal(got,want): print([hex(v) for v in got[:8]], [hex(v) for v in want[:8]])
