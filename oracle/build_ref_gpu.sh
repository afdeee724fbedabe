#!/bin/bash
# TEST INFRASTRUCTURE: baseline/_ref/libsphexa_ref_gpu.so = the reference's GPU-only block-time-step group kernels
# (sph/include/sph/ts_groups.cu, cstone/traversal/groups.cuh; cstone/primitives/primitives_gpu.cu for the scans they call) compiled for sm_100a from the sources where they lie,
# behind the C veneer oracle/ref_gpu_wrap.cu.  Loaded by tests/test_gpu_reference_kernels.py only.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../baseline/_ref
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
[ -d "$REF/sph/include/sph" ] || { echo "reference tree $REF absent: keeping prebuilt library"; exit 0; }
mkdir -p "$OUT/overlay_gpu/sph"
# the stray #endif of the fork's particles_data.hpp (line 270), removed in a temporary copy as in oracle/Makefile
sed '270d' "$REF/sph/include/sph/particles_data.hpp" > "$OUT/overlay_gpu/sph/particles_data.hpp"
$NVCC -std=c++17 -O3 -DNDEBUG -DUSE_CUDA -w -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
  --extended-lambda -ccbin /usr/bin/g++ -Xcompiler -fPIC -Xcompiler -fopenmp -shared -include map -include sstream \
  -I"$OUT/overlay_gpu" -I"$HERE/sphexa_cpu" -I"$REF/domain/include" -I"$REF/sph/include" -I"$REF/sph/include/sph" \
  -I"$REF/ryoanji/src" -I"$REF/physics/cooling/include" -I"$REF/main/src" \
  "$HERE/ref_gpu_wrap.cu" "$REF/sph/include/sph/ts_groups.cu" "$REF/domain/include/cstone/primitives/primitives_gpu.cu" -o "$OUT/libsphexa_ref_gpu.so"
rm -rf "$OUT/overlay_gpu"
ls -la "$OUT/libsphexa_ref_gpu.so"
