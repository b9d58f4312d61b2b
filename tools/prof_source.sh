#!/bin/bash
# Source-level capture of ONE launch of the step's scoring kernel (score_both_kernel) with the source imported, for the
# per-line stall picture.  usage: tools/prof_source.sh <tag> [bench args...]   (one GPU; run after the plain command works)
TAG=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-plan --no-cpu-baseline $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'score_both_kernel' -s 4 -c 1 -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "$TAG rc=$?"
ls -la gpurun_out/${TAG}.ncu-rep
