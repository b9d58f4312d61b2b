#!/bin/bash
# compute-sanitizer is closed on this pool ("runs under it have left GPUs needing a reset"), so the memory-safety evidence
# comes from a build of libnosplan with every array index asserted on the device (-DNOSPLAN_CHECKS), run through the
# parity suites; and the same build with the sequencer's MPS update kept out of line (-DNOSPLAN_SEQ_NOINLINE, round 1's
# form: nosplan_seq.cuh tells the story).  usage: tools/run_checked.sh [outdir]   (one GPU)
OUT=${1:-gpurun_out/checked}
mkdir -p "$OUT"
# (build them beforehand on the CPU box: `make -C walkai-nos_b200/csrc ../libnosplan_checked.so ../libnosplan_checked_noinl.so`;
# they travel with the snapshot.  Only a missing library is built here.)
for v in checked checked_noinl; do
  [ -f walkai-nos_b200/libnosplan_$v.so ] || make -C walkai-nos_b200/csrc ../libnosplan_$v.so >> "$OUT/build.log" 2>&1
done
SEL="tests/test_gpu_kat.py tests/test_gpu_parity.py tests/test_gpu_limits.py tests/test_gpu_pass.py tests/test_gpu_stream.py tests/test_gpu_stress.py tests/test_gpu_snapshot_checks.py tests/test_controller_kats.py tests/test_zz_no_alloc.py"
for v in checked checked_noinl; do
  NOSPLAN_LIB=$PWD/walkai-nos_b200/libnosplan_$v.so timeout ${CHK_TIMEOUT:-900} python -m pytest $SEL -m gpu -q -p no:cacheprovider > "$OUT/$v.pytest.txt" 2>&1
  echo "$v rc=$? $(tail -1 $OUT/$v.pytest.txt)" | tee -a "$OUT/summary.txt"
  grep -c "Assertion" "$OUT/$v.pytest.txt" | sed "s/^/$v device assertions reported: /" | tee -a "$OUT/summary.txt"
done
