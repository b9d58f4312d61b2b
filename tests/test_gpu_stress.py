"""-m gpu: 20 seeds of the randomised parity stress (tools/stress_parity.py) as collected tests: ragged random
clusters (mixed models, out-of-table profiles, invalid nodes, present-but-zero MPS keys, multi-key rows,
quantities 0..40) of 8-400 MIG + 8-300 MPS nodes, both scorers in every output mode and the planner in
every flag combination, bit-exact against the oracle."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture(scope="module")
def planner():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    yield p
    p.close()


@pytest.mark.parametrize("seed", list(range(300, 320)))
def test_stress_seed(planner, seed):
    import stress_parity
    n_mig, n_mps, rows, pods = stress_parity.run_seed(planner, seed)
    assert rows > 0 and n_mig >= 8 and n_mps >= 8
