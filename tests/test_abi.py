"""No-GPU checks of the drop-in boundary: the library loads, exports every symbol include/nosplan.h
declares, and fails loudly (never silently) when no CUDA device exists."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "nosplan.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nosplan_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import walkai_nos_b200 as pkg
    pkg.build_library()
    L = C.CDLL(pkg.library_path())
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} declared in nosplan.h but not exported"
    assert sorted(pkg.nosplan.EXPORTS) == [n for n in names if n in pkg.nosplan.EXPORTS]
    assert set(pkg.nosplan.EXPORTS) <= set(names)


def test_every_entry_point_cites_the_reference():
    src = open(os.path.join(ROOT, "include", "nosplan.h")).read()
    for needle in ("pkg/gpu/mig/gpu.go", "pkg/gpu/mig/node.go", "pkg/gpu/slicing/gpu.go", "mig_controller.go",
                   "known_configs.go", "internal/partitioning/mig/state.go"):
        assert needle in src


def test_no_silent_fallback_without_a_device():
    import torch

    import walkai_nos_b200 as pkg
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(pkg.NosplanError):
        pkg.Planner(0)


def test_product_does_not_touch_the_oracle():
    """The product path must never route through oracle/ (only tests, smoke() and bench's CPU legs may)."""
    pkg_dir = os.path.join(ROOT, "walkai-nos_b200")
    pat = re.compile(r"(from\s+oracle|import\s+oracle|libnos_oracle|nos_oracle\.|#include\s+\".*oracle)")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp", "Makefile")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not pat.search(txt), f"{f} imports/links the oracle"


def _build_c_client():
    import subprocess

    import walkai_nos_b200 as pkg
    pkg.build_library()
    libdir = os.path.dirname(pkg.library_path())
    exe = os.path.join(ROOT, "tests", "abi_client.bin")
    subprocess.check_call(["gcc", "-std=c11", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "abi_client.c"), "-o", exe, "-L", libdir, "-lnosplan", "-lnoshost",
                           "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64"])
    return exe


def test_plain_c_client_links_and_fails_loudly_without_gpu():
    import subprocess
    exe = _build_c_client()
    assert subprocess.call([exe]) == 0


@pytest.mark.gpu
def test_plain_c_client_plans_on_gpu():
    import subprocess
    exe = _build_c_client()
    assert subprocess.call([exe, "gpu"]) == 0
