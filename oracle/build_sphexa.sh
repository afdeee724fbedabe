#!/bin/bash
# TEST INFRASTRUCTURE: builds the reference's own CPU front-end `sphexa` (main/src/sphexa/sphexa.cpp) from the sources
# where they lie under /root/reference into baseline/_ref/sphexa (git-ignored, travels to the GPU box), following
# SURVEY.md Appendix A.  BASELINE.json configs[0] is `sphexa --init sedov -n 50 -s 10` with this binary.
# The fork snapshot does not compile as it is; two one-line fixes are applied to TEMPORARY copies of two headers that
# shadow the originals on the include path and are deleted again after the build:
#   sph/particles_data.hpp  line 270: a stray #endif
#   sph/positions.hpp       lines 127, 144: `decltype(d.du[0])` is a reference type; the value type is meant
# HDF5 is absent: built without SPH_EXA_HAVE_H5PART (run with --ascii or without output).
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/../baseline/_ref
[ -d "$REF/main/src/sphexa" ] || { echo "reference tree $REF absent: keeping prebuilt $OUT/sphexa"; exit 0; }
mkdir -p "$OUT/build/overlay/sph" "$OUT/build/gen"
sed '270d' "$REF/sph/include/sph/particles_data.hpp" > "$OUT/build/overlay/sph/particles_data.hpp"
sed -e 's/using Tdu *= *decltype(d.du\[0\]);/using Tdu = std::decay_t<decltype(d.du[0])>;/' \
    "$REF/sph/include/sph/positions.hpp" > "$OUT/build/overlay/sph/positions.hpp"
printf '#pragma once\n#define GIT_BRANCH "reference"\n#define GIT_COMMIT_HASH "snapshot"\n' > "$OUT/build/gen/version.h"
INC="-include map -include sstream -I$OUT/build/overlay -I$HERE/sphexa_cpu -I$OUT/build/gen -I$REF/domain/include \
 -I$REF/sph/include -I$REF/sph/include/sph -I$REF/ryoanji/src -I$REF/physics/cooling/include -I$REF/main/src"
FLAGS="-std=c++20 -O3 -march=x86-64-v3 -DNDEBUG -fopenmp -w"
pids=()
for f in sphexa/sphexa.cpp init/sim_init.cpp io/ifile_io_ascii.cpp io/ifile_io_hdf5.cpp observables/observables.cpp; do
  o=$OUT/build/$(basename $f .cpp).o
  /usr/bin/g++ $FLAGS $INC -c "$REF/main/src/$f" -o "$o" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
/usr/bin/g++ -fopenmp $OUT/build/*.o -o "$OUT/sphexa"
rm -rf "$OUT/build"
echo "built $OUT/sphexa"
