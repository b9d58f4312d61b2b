#!/bin/bash
# compute-sanitizer over the GPU parity tests (one GPU).  Usage: tools/run_sanitizers.sh [outdir]
# memcheck: out-of-bounds / misaligned accesses (global, shared, local); racecheck: shared-memory hazards;
# initcheck: reads of uninitialised global memory.  Each tool runs the known-answer tests plus the randomised
# and limit cases of the parity suite (small inputs: the tools slow kernels down 10-100x).
OUT=${1:-gpurun_out/sanitizer}
mkdir -p "$OUT"
SEL='tests/test_gpu_kat.py tests/test_gpu_parity.py tests/test_gpu_limits.py tests/test_zz_no_alloc.py'
KEXPR='kat or random or limits or static or empty'
for tool in memcheck racecheck initcheck; do
  extra=""
  [ "$tool" = memcheck ] && extra="--leak-check no"
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool $extra --print-limit 50 --log-file "$OUT/$tool.log" \
    python -m pytest $SEL -m gpu -q -x -k "$KEXPR" -p no:cacheprovider > "$OUT/$tool.pytest.txt" 2>&1
  echo "$tool rc=$?" >> "$OUT/summary.txt"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" "$OUT/$tool.log" >> "$OUT/summary.txt"
  tail -3 "$OUT/$tool.pytest.txt" >> "$OUT/summary.txt"
done
cat "$OUT/summary.txt"
