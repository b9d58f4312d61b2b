"""No GPU: libnoshost built with AddressSanitizer + UndefinedBehaviorSanitizer and driven through the host-mirror,
table-loader and controller known-answer tests (tools/run_host_sanitizers.sh).  Skipped where the sanitizer
runtimes are not installed."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_library_is_clean_under_asan_and_ubsan(tmp_path):
    asan = subprocess.run(["/usr/bin/gcc", "-print-file-name=libasan.so"], capture_output=True, text=True).stdout.strip()
    if not os.path.isabs(asan) or not os.path.exists(asan):
        pytest.skip("libasan is not installed")
    r = subprocess.run(["bash", os.path.join(ROOT, "tools", "run_host_sanitizers.sh"), str(tmp_path)], cwd=ROOT, capture_output=True,
                       text=True, timeout=600)
    out = r.stdout + r.stderr
    assert "host sanitizers rc=0" in out, out[-3000:]
    assert "AddressSanitizer" not in out and "runtime error" not in out, out[-3000:]
    assert " passed" in out
