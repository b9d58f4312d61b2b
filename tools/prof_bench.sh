#!/bin/bash
# ncu captures of the two dominant kernels of bench.py (one GPU).  usage: tools/prof_bench.sh <tag> [bench args...]
# Writes gpurun_out/<tag>.ncu-rep (full sets, source imported) after the same command has exited 0 without ncu.
TAG=$1; shift
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-plan --no-cpu-baseline $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'mig_score16_kernel|mps_score1_kernel' -s 6 -c 2 \
    -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "$TAG rc=$?"
