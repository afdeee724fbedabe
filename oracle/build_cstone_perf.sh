#!/bin/bash
# TEST INFRASTRUCTURE: builds the reference's own GPU performance drivers (domain/test/performance/neighbor_driver.cu,
# hilbert.cu, octree.cu: self-checking, single process, no MPI) for sm_100a from the sources where they lie, into
# baseline/_ref/ (git-ignored, travels to the GPU box).  They time the reference's generic CUDA kernels for the
# neighbor search, the SFC keys + key sort and the octree build on a B200: the kernel-level numbers of BASELINE.md
# section 3b, compared with this library on the same workloads by profiles/microbench/reference_drivers.py.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../baseline/_ref
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
[ -d "$REF/domain/test/performance" ] || { echo "reference tree $REF absent: keeping prebuilt drivers"; exit 0; }
mkdir -p "$OUT"
FLAGS="-std=c++17 -O3 -DNDEBUG -DUSE_CUDA -w -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
 --extended-lambda -ccbin /usr/bin/g++ -Xcompiler -fopenmp -I$REF/domain/include -I$REF/domain/test \
 -I$HERE/sphexa_cpu"
D=$REF/domain/include/cstone
T=$REF/domain/test/performance
[ -s "$OUT/cstone_neighbors_perf" ] || $NVCC $FLAGS $T/neighbor_driver.cu -o "$OUT/cstone_neighbors_perf"
[ -s "$OUT/cstone_hilbert_perf" ] || $NVCC $FLAGS $T/hilbert.cu $D/sfc/sfc_gpu.cu -o "$OUT/cstone_hilbert_perf"
[ -s "$OUT/cstone_octree_perf" ] || $NVCC $FLAGS $T/octree.cu $D/tree/csarray_gpu.cu $D/tree/octree_gpu.cu \
    $D/traversal/collisions_gpu.cu $D/primitives/primitives_gpu.cu $D/util/reallocate.cu -o "$OUT/cstone_octree_perf"
ls -la "$OUT"/cstone_*_perf
