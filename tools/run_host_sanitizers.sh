#!/bin/bash
# libnoshost under AddressSanitizer + UndefinedBehaviorSanitizer: the host-mirror tests (annotation wire format,
# snapshot builder, pod rows, table loader, controller known answers) run against the instrumented build.  No GPU needed.
OUT=${1:-gpurun_out/host_sanitizers}
mkdir -p "$OUT"
make -C walkai-nos_b200/csrc ../libnoshost_asan.so > "$OUT/build.log" 2>&1 || { echo "build failed"; exit 1; }
ASAN=$(/usr/bin/gcc -print-file-name=libasan.so)
UBSAN=$(/usr/bin/gcc -print-file-name=libubsan.so)
NOSHOST_LIB=$PWD/walkai-nos_b200/libnoshost_asan.so LD_PRELOAD="$ASAN $UBSAN" ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1 \
  python -m pytest tests/test_host_mirror.py tests/test_tables_loader.py tests/test_controller_kats.py tests/test_host_fuzz.py -m "not gpu" -q -p no:cacheprovider > "$OUT/pytest.txt" 2>&1
echo "host sanitizers rc=$? $(tail -1 $OUT/pytest.txt)" | tee "$OUT/summary.txt"
grep -E "ERROR: AddressSanitizer|runtime error" "$OUT/pytest.txt" | head -5 | tee -a "$OUT/summary.txt"
