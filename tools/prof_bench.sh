#!/bin/bash
# ncu captures of bench.py's kernels (one GPU).  usage: tools/prof_bench.sh <tag> [bench args...]
# Writes gpurun_out/<tag>.ncu-rep: full sets (source imported) of one launch each of score_both_kernel (the step's
# scoring launch), mig_score16_kernel and mps_score1_kernel (the kinds one after the other: bench.py's second loop),
# and gpurun_out/<tag>_launches.csv: every launch of the command with its duration — both only after the same command
# has exited 0 without ncu.
TAG=$1; shift
# NCU_SRC="--import-source on" adds the source view (about 5x larger report)
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-plan --no-cpu-baseline $*"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none $NCU_SRC -k regex:'score_both_kernel|mig_score16_kernel|mps_score1_kernel' -s 4 -c 5 \
    -f -o gpurun_out/${TAG} $CMD > gpurun_out/${TAG}_ncu.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "$TAG rc=$?"
# the raw page as CSV travels back (a few hundred KB); the report itself only when asked for (KEEP_REP=1: ~25 MB, and
# gpurun returns at most 64 MiB)
ncu -i gpurun_out/${TAG}.ncu-rep --page raw --csv > gpurun_out/${TAG}_raw.csv 2>/dev/null
[ "$KEEP_REP" = 1 ] || rm -f gpurun_out/${TAG}.ncu-rep
