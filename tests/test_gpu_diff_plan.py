"""-m gpu: the agent-side diff planner through the C ABI against the count-level oracle (large random
batch) and against the reference's table tests (via the device-level oracle's counts)."""
import json
import os

import numpy as np
import pytest

from oracle import diff_plan
from oracle import oracle as orc
from test_diff_plan import CASES, _spec_list, dense

pytestmark = pytest.mark.gpu


def test_diff_plan_matches_oracle():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    try:
        rng = np.random.default_rng(5)
        for G, P in ((1, 1), (7, 5), (32768, 13), (1000, 16)):
            used = rng.integers(0, 4, size=(G, P)).astype(np.int32)
            free = rng.integers(0, 4, size=(G, P)).astype(np.int32)
            spec = rng.integers(0, 8, size=(G, P)).astype(np.int32)
            mask = rng.integers(0, 1 << P, size=G).astype(np.uint32)
            mask[rng.random(G) < 0.2] = 0  # GPUs without any spec annotation
            got = p.diff_plan_mig(used, free, spec, mask)
            exp = orc.mig_diff_counts(used, free, spec, mask)
            for a, b in zip(got, exp):
                assert np.array_equal(a, b)
        profiles = ["1g.10gb", "1g.20gb", "2g.20gb", "4g.20gb"]
        for c in CASES:
            state = {int(g): [tuple(d) for d in devs] for g, devs in c["state"].items()}
            spec = _spec_list(c["spec"])
            cr, dl = diff_plan.new_mig_config_plan(state, spec)
            ec, edf, edu = diff_plan.counts(cr, dl)
            u, f, s, m = dense(state, spec, 2, profiles)
            gc, gdf, gdu = p.diff_plan_mig(u, f, s, m)
            for g in range(2):
                for i, pr in enumerate(profiles):
                    assert (gc[g, i], gdf[g, i], gdu[g, i]) == (ec.get((g, pr), 0), edf.get((g, pr), 0), edu.get((g, pr), 0))
    finally:
        p.close()
