#!/bin/bash
# TEST INFRASTRUCTURE: builds the reference's own GPU front-end `sphexa-cuda` (main/src/sphexa/sphexa.cpp with
# USE_CUDA + cstone_gpu + sph_gpu + ryoanji + observables_gpu, the targets of main/src/sphexa/CMakeLists.txt:36-41)
# for sm_100a, from the sources where they lie under /root/reference, into baseline/_ref/sphexa-cuda (git-ignored,
# travels to the GPU box).  It is the second on-box baseline of BASELINE.md section 3b: the reference's generic CUDA
# kernels recompiled for the B200, timed by bench.py (`reference_gpu`) on the headline configuration.  Never linked
# into the product.  Same two temporary header fixes and single-rank mpi.h as oracle/build_sphexa.sh.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../baseline/_ref
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
[ -d "$REF/main/src/sphexa" ] || { echo "reference tree $REF absent: keeping prebuilt $OUT/sphexa-cuda"; exit 0; }
B=$OUT/obj_cuda   # kept (git- and gpurun-ignored): the drop-in link below reuses the objects
mkdir -p "$B/overlay/sph" "$B/gen" "$B/obj"
sed '270d' "$REF/sph/include/sph/particles_data.hpp" > "$B/overlay/sph/particles_data.hpp"
sed -e 's/using Tdu *= *decltype(d.du\[0\]);/using Tdu = std::decay_t<decltype(d.du[0])>;/' \
    "$REF/sph/include/sph/positions.hpp" > "$B/overlay/sph/positions.hpp"
printf '#pragma once\n#define GIT_BRANCH "reference"\n#define GIT_COMMIT_HASH "snapshot"\n' > "$B/gen/version.h"
INC="-I$B/overlay -I$HERE/sphexa_cpu -I$B/gen -I$REF/domain/include -I$REF/sph/include -I$REF/sph/include/sph \
 -I$REF/ryoanji/src -I$REF/physics/cooling/include -I$REF/main/src -I/usr/local/cuda/include"
# host TUs see thrust::device_vector in signatures: thrust's inline ABI namespace embeds __CUDA_ARCH_LIST__, which only
# nvcc defines (1000 for sm_100a) -- give g++ the same value or the two halves do not link
CXXFLAGS="-std=c++20 -O3 -march=x86-64-v3 -DNDEBUG -DUSE_CUDA -D__CUDA_ARCH_LIST__=1000 -fopenmp -w -include map -include sstream"
CUFLAGS="-std=c++17 -O3 -DNDEBUG -DUSE_CUDA -w -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr \
 --extended-lambda -ccbin /usr/bin/g++ -Xcompiler -fopenmp -include map -include sstream"
CU="domain/include/cstone/focus/rebalance_gpu.cu domain/include/cstone/focus/source_center_gpu.cu
 domain/include/cstone/halos/gather_halos_gpu.cu domain/include/cstone/primitives/primitives_gpu.cu
 domain/include/cstone/sfc/sfc_gpu.cu domain/include/cstone/traversal/collisions_gpu.cu
 domain/include/cstone/tree/csarray_gpu.cu domain/include/cstone/tree/octree_gpu.cu
 domain/include/cstone/util/reallocate.cu main/src/observables/conserved_gpu.cu main/src/observables/gpu_reductions.cu
 ryoanji/src/ryoanji/interface/ewald.cu ryoanji/src/ryoanji/interface/multipole_holder.cu
 ryoanji/src/ryoanji/interface/treebuilder.cu sph/include/sph/groups.cu sph/include/sph/hydro_std/eos_gpu.cu
 sph/include/sph/hydro_std/iad_gpu.cu sph/include/sph/hydro_std/momentum_energy_gpu.cu
 sph/include/sph/hydro_turb/stirring_gpu.cu
 sph/include/sph/hydro_ve/av_switches_gpu.cu sph/include/sph/hydro_ve/eos_gpu.cu
 sph/include/sph/hydro_ve/iad_divv_curlv_gpu.cu sph/include/sph/hydro_ve/momentum_energy_gpu.cu
 sph/include/sph/hydro_ve/ve_def_gradh_gpu.cu sph/include/sph/hydro_ve/xmass_gpu.cu sph/include/sph/positions_gpu.cu
 sph/include/sph/ts_groups.cu sph/include/sph/update_h_gpu.cu"
# (sph/include/sph/hydro_ve/additional_fields.cu, a fork-specific markRamp helper, does not compile in this snapshot;
#  nothing under main/src calls it)
CPP="sphexa/sphexa.cpp init/sim_init.cpp io/ifile_io_ascii.cpp io/ifile_io_hdf5.cpp observables/observables.cpp"
JOBS=${JOBS:-8}
fail=0
run() { # name, command...
  local name=$1; shift
  if ! "$@" > "$B/obj/$name.log" 2>&1; then echo "FAILED $name"; tail -5 "$B/obj/$name.log"; return 1; fi
}
export -f run
n=0
for f in $CU; do
  o=$B/obj/$(echo $f | tr '/' '_').o
  [ -s "$o" ] && continue
  ( run "$(basename $f)" $NVCC $CUFLAGS $INC -c "$REF/$f" -o "$o" ) &
  n=$((n+1)); if [ $((n % JOBS)) -eq 0 ]; then wait; fi
done
for f in $CPP; do
  o=$B/obj/$(echo $f | tr '/' '_').o
  [ -s "$o" ] && continue
  # (nvcc's front-end does not get through these C++20 translation units, so they stay with g++ like in the
  #  reference's own CMake build)
  ( run "$(basename $f)" /usr/bin/g++ $CXXFLAGS $INC -c "$REF/main/src/$f" -o "$o" ) &
  n=$((n+1)); if [ $((n % JOBS)) -eq 0 ]; then wait; fi
done
wait
ls $B/obj/*.o | wc -l
/usr/bin/g++ -O2 -I/usr/local/cuda/include -c "$HERE/sphexa_cpu/launch_guard.cpp" -o "$B/obj/zz_launch_guard.o"
$NVCC -gencode arch=compute_100a,code=sm_100a -Xcompiler -fopenmp -Xlinker --wrap=cudaLaunchKernel $B/obj/*.o \
      -o "$OUT/sphexa-cuda" -lcudart -lgomp
echo "built $OUT/sphexa-cuda"

# ---- drop-in link: the SAME reference front-end, Domain and propagators, with the five SPH-VE pair-operator objects of the
# reference (sph/include/sph/hydro_ve/*_gpu.cu) replaced by the forwarding definitions of
# sph-exa_b200/host/b200_forward.cu, which are compiled against the reference's own declarations (sph/sph_gpu.hpp) and
# call the C ABI of libsphexa_b200.so
LIB=$HERE/../sph-exa_b200
if [ -f "$LIB/libsphexa_b200.so" ] && [ -f "$LIB/host/b200_forward.cu" ]; then
  $NVCC $CUFLAGS $INC -I$HERE/../include -c "$LIB/host/b200_forward.cu" -o "$B/b200_forward.o"
  # ... and the reference's focus/rebalance_gpu.cu object (rebalanceDecisionEssentialGpu, enforceKeysGpu,
  # protectAncestorsGpu: the decisions of its FocusedOctree) replaced by host/b200_forward_focus.cu
  $NVCC $CUFLAGS $INC -I$HERE/../include -c "$LIB/host/b200_forward_focus.cu" -o "$B/b200_forward_focus.o"
  # ... and host/b200_forward_domain.cu: explicit specializations (strong symbols) of the Cornerstone GPU functions the
  # SPH front-end uses (keys, linked octree, node geometry, halo search, gather) override the reference's weak
  # instantiations; its objects stay in the link for the instantiations that are not forwarded
  $NVCC $CUFLAGS $INC -I$HERE/../include -c "$LIB/host/b200_forward_domain.cu" -o "$B/b200_forward_domain.o"
  KEEP=$(ls $B/obj/*.o | grep -v -E "hydro_ve_(xmass|ve_def_gradh|iad_divv_curlv|av_switches|momentum_energy)_gpu|focus_rebalance_gpu")
  $NVCC -gencode arch=compute_100a,code=sm_100a -Xcompiler -fopenmp -Xlinker --wrap=cudaLaunchKernel $KEEP \
        "$B/b200_forward.o" "$B/b200_forward_focus.o" "$B/b200_forward_domain.o" -o "$OUT/sphexa-cuda-b200" -L"$LIB" -lsphexa_b200 -Xlinker -rpath -Xlinker '$ORIGIN/../../sph-exa_b200' \
        -lcudart -lgomp
  echo "built $OUT/sphexa-cuda-b200"
fi
