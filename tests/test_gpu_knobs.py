"""-m gpu: the fallback scorers must stay bit-exact too.  NOSPLAN_PLAIN_ORDER=1 (rows and nodes in input order:
mig_score_kernel's 16-bit-key path, mps_score_kernel mode 0) and NOSPLAN_KEY_BITS=32 (wide keys for every
table) are read once per process, so the known-answer and limit suites are re-run in child processes."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("knob", ["NOSPLAN_PLAIN_ORDER=1", "NOSPLAN_KEY_BITS=32", "NOSPLAN_SCORE_WAVES=1"])
def test_parity_suites_under_knob(knob):
    env = dict(os.environ)
    k, v = knob.split("=")
    env[k] = v
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-x", "-p", "no:cacheprovider",
                        os.path.join(ROOT, "tests", "test_gpu_kat.py"), os.path.join(ROOT, "tests", "test_gpu_limits.py"),
                        os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-k",
                        "kat or limits or score_mig_configs or score_mps_configs or score_random or empty_inputs"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, knob + "\n" + r.stdout[-3000:] + r.stderr[-2000:]
    import re
    m = re.search(r"(\d+) passed", r.stdout)
    assert m and int(m.group(1)) >= 40, r.stdout[-500:]  # the child really ran the suites
