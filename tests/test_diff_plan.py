"""Agent-side diff planner: the device-level oracle against the reference's table tests,
and the count-level oracle against the device-level one on random states.  CPU only."""
import json
import os

import numpy as np
import pytest

from oracle import diff_plan
from oracle import oracle as orc

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diff_plan_kats.json")
CASES = json.load(open(GOLDEN))["cases"]


def _spec_list(spec):
    out = []
    for k, v in spec.items():
        a = orc.parse_spec_annotation(k, v)
        assert a is not None
        out.append((a["index"], a["profile"], a["quantity"]))
    return out


@pytest.mark.parametrize("c", CASES, ids=lambda c: c["src"])
def test_reference_table(c):
    state = {int(g): [tuple(d) for d in devs] for g, devs in c["state"].items()}
    create, delete = diff_plan.new_mig_config_plan(state, _spec_list(c["spec"]))
    assert sorted(create) == sorted(tuple(x) for x in c["create"])  # ElementsMatch
    got = sorted(sorted((r[0], r[1]) for r in op) for op in delete)
    exp = sorted(sorted((g, d) for g, d in op) for op in c["delete"])
    assert got == exp


def random_case(rng, G=6, profiles=("1g.10gb", "2g.20gb", "3g.40gb", "7g.79gb")):
    state, spec = {}, []
    did = 0
    for g in range(G):
        if rng.random() < 0.8:
            devs = []
            for p in profiles:
                for _ in range(int(rng.integers(0, 4))):
                    did += 1
                    devs.append((f"{did:03d}", p, "free" if rng.random() < 0.5 else "used"))
            state[g] = devs
        if rng.random() < 0.8:
            for p in profiles:
                if rng.random() < 0.6:
                    spec.append((g, p, int(rng.integers(0, 5))))
    return state, spec


def dense(state, spec, G, profiles):
    P = len(profiles)
    used = np.zeros((G, P), np.int32)
    free = np.zeros((G, P), np.int32)
    sp = np.zeros((G, P), np.int32)
    mask = np.zeros(G, np.uint32)
    for g, devs in state.items():
        for (_, p, s) in devs:
            (free if s == "free" else used)[g, profiles.index(p)] += 1
    for (g, p, q) in spec:
        sp[g, profiles.index(p)] += q
        mask[g] |= np.uint32(1 << profiles.index(p))
    return used, free, sp, mask


@pytest.mark.parametrize("seed", range(20))
def test_count_level_matches_device_level(seed):
    rng = np.random.default_rng(seed)
    profiles = ["1g.10gb", "2g.20gb", "3g.40gb", "7g.79gb"]
    state, spec = random_case(rng, 6, profiles)
    create, delete = diff_plan.new_mig_config_plan(state, spec)
    c, df, du = diff_plan.counts(create, delete)
    used, free, sp, mask = dense(state, spec, 6, profiles)
    oc, odf, odu = orc.mig_diff_counts(used, free, sp, mask)
    for g in range(6):
        for i, p in enumerate(profiles):
            assert oc[g, i] == c.get((g, p), 0), (g, p)
            assert odf[g, i] == df.get((g, p), 0), (g, p)
            assert odu[g, i] == du.get((g, p), 0), (g, p)
