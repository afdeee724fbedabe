#!/bin/bash
# Profile capture of the bench workload (Sedov 200^3, f64, 1 B200), run on the GPU box through gpurun:
#   gpurun --timeout 1500 -- 'bash profiles/capture.sh r01e'
# Writes gpurun_out/launches_<tag>.csv (every launch with its device time), gpurun_out/full_<tag>.ncu-rep (ncu --set
# full of one launch of each pair kernel + the neighbor search, 4th step) and gpurun_out/courant_<tag>.ncu-rep (the
# AVswitches+Courant kernel of sphb200_sim_step_host), gpurun_out/sync_<tag>.ncu-rep and gpurun_out/sort_<tag>.ncu-rep
# (SFC keys, field gather, one radix pass).  profiles/summarize.py turns them into the committed summaries.
set -u
TAG=${1:-r01e}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-bdt --no-configs"
# The context picks the kernel-table variant of the pair operators from its own timings.  Under ncu
# every launch is replayed serialised on cold caches and those timings are not the ones of a plain run, so the captures
# pin the variant plain runs of this workload choose (lattice: the L1 table).  PARTS selects passes (default: all).
export SPHB200_TABLE=${SPHB200_TABLE:-l1}
PARTS=${2:-"plain launches full courant sync sort"}
has() { case " $PARTS " in *" $1 "*) return 0;; *) return 1;; esac; }
has plain && { $CMD --no-e2e > gpurun_out/plain_$TAG.log 2>&1 || exit 1; }
has launches && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$TAG.csv \
    $CMD --no-e2e > gpurun_out/ncu_launches_$TAG.log 2>&1
has full && ncu --set full --clock-control none --import-source on \
    -k regex:'findNeighborsKernel|xmassKernel|veDefGradhKernel|iadDivvCurlvKernel|avSwitchesKernel|momentumEnergyKernel' \
    --launch-skip 18 -c 6 -f -o gpurun_out/full_$TAG $CMD --no-e2e > gpurun_out/ncu_full_$TAG.log 2>&1
has courant && ncu --set full --clock-control none --import-source on -k regex:'avSwitchesCourantKernel' --launch-skip 1 -c 1 \
    -f -o gpurun_out/courant_$TAG $CMD > gpurun_out/ncu_courant_$TAG.log 2>&1
# the HBM-bound kernels of Domain::sync (4th step): SFC keys, the three kernels of the first radix pass of the key sort,
# the field gather
has sync && ncu --set full --clock-control none --import-source on -k regex:'sfcKeysKernel|gatherFieldsKernel' --launch-skip 6 -c 2 \
    -f -o gpurun_out/sync_$TAG $CMD --no-e2e > gpurun_out/ncu_sync_$TAG.log 2>&1
has sort && ncu --set full --clock-control none --import-source on -k regex:'radixCountKernel|radixRowScanKernel|radixScatterKernel' \
    --launch-skip 144 -c 3 -f -o gpurun_out/sort_$TAG $CMD --no-e2e > gpurun_out/ncu_sort_$TAG.log 2>&1
ls -la gpurun_out/*.ncu-rep
