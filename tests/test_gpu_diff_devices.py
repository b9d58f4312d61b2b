"""-m gpu: the agent-side diff planner at DEVICE level through the C ABI (nosplan_diff_plan_mig_devices): the six
TestNewMigConfigPlan vectors of the reference (internal/controllers/migagent/plan/plan_test.go:29-391, transcribed in
tests/golden/diff_plan_kats.json) compared as the reference compares them — the same create operations and the same
delete operations as multisets, each delete operation the same set of device ids — plus TestMigState_Matches
(mig_state_test.go:29-95) and random states against the device-level oracle."""
import numpy as np
import pytest

from oracle import diff_plan
from test_diff_plan import CASES, _spec_list

pytestmark = pytest.mark.gpu


def _flatten(state, spec, n_gpus, profiles):
    """What the agent would hand over: per GPU its devices in ascending device-id order (plan.go:45)."""
    pid = {p: i for i, p in enumerate(profiles)}
    off, prof, status, ids = [0], [], [], []
    st_code = {"used": 0, "free": 1, "unknown": 2}
    for g in range(n_gpus):
        for (dev, p, s) in sorted(state.get(g, []), key=lambda r: r[0]):
            prof.append(pid[p])
            status.append(st_code[s])
            ids.append((g, dev, p, s))
        off.append(len(prof))
    spec_q = np.zeros((n_gpus, len(profiles)), np.int32)
    mask = np.zeros(n_gpus, np.uint32)
    for (g, p, q) in spec:
        spec_q[g, pid[p]] += q
        mask[g] |= np.uint32(1 << pid[p])
    return np.asarray(off, np.int32), np.asarray(prof or [0], np.int32), np.asarray(status or [0], np.uint8), spec_q, mask, ids


def _operations(profiles, off, ids, op, cn, ca):
    """MigConfigPlan from the arrays: DeleteOperations grouped as plan.go groups them, CreateOperations per (GPU, profile)."""
    creates = []
    for g in range(cn.shape[0]):
        for i, p in enumerate(profiles):
            if cn[g, i]:
                creates.append((g, p, int(cn[g, i])))
            if ca[g, i]:
                creates.append((g, p, int(ca[g, i])))
    groups = {}
    for i, (g, dev, p, s) in enumerate(ids):
        if op[i] == 0:
            continue
        key = (int(op[i]), g, p) if op[i] != 3 else (3, g)
        groups.setdefault(key, []).append((g, dev))
    return sorted(creates), sorted(sorted(v) for v in groups.values())


@pytest.fixture(scope="module")
def planner():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    yield p
    p.close()


@pytest.mark.parametrize("case", CASES, ids=[c["src"] for c in CASES])
def test_new_mig_config_plan_vectors_through_the_abi(planner, case):
    profiles = sorted({"1g.10gb", "1g.20gb", "2g.20gb", "4g.20gb"} | {d[1] for devs in case["state"].values() for d in devs})
    state = {int(g): [tuple(d) for d in devs] for g, devs in case["state"].items()}
    spec = _spec_list(case["spec"])
    profiles = sorted(set(profiles) | {p for (_, p, _) in spec})
    G = 1 + max([0] + list(state) + [g for (g, _, _) in spec])
    off, prof, status, spec_q, mask, ids = _flatten(state, spec, G, profiles)
    op, cn, ca, m = planner.diff_plan_mig_devices(off, prof, status, spec_q, mask)
    creates, deletes = _operations(profiles, off, ids, op, cn, ca)
    assert creates == sorted(tuple(c) for c in case["create"])
    assert deletes == sorted(sorted(tuple(d) for d in o) for o in case["delete"])


def test_mig_state_matches_vectors(planner):
    profiles = ["1g.10gb", "1g.20gb", "2g.40gb"]
    # mig_state_test.go:36-41 "Empty": no devices, no annotations
    op, cn, ca, m = planner.diff_plan_mig_devices(np.zeros(2, np.int32), [], [], np.zeros((1, 3), np.int32), np.zeros(1, np.uint32))
    assert m.all()
    # :42-83 "Matches": GPU 0 {1g.10gb x2, 2g.40gb x1}, GPU 1 {1g.20gb x2}; spec the same quantities
    state = {0: [("a", "1g.10gb", "used"), ("b", "1g.10gb", "free"), ("c", "2g.40gb", "used")], 1: [("d", "1g.20gb", "free"), ("e", "1g.20gb", "free")]}
    spec = [(0, "1g.10gb", 2), (0, "2g.40gb", 1), (1, "1g.20gb", 2)]
    off, prof, status, spec_q, mask, ids = _flatten(state, spec, 2, profiles)
    op, cn, ca, m = planner.diff_plan_mig_devices(off, prof, status, spec_q, mask)
    assert m.all() and not op.any() and not cn.any() and not ca.any()
    # and what makes the two maps differ (mig_state.go:40-58): a quantity, a missing key on either side, a zero-quantity annotation
    for bad in ([(0, "1g.10gb", 3), (0, "2g.40gb", 1), (1, "1g.20gb", 2)], [(0, "1g.10gb", 2), (1, "1g.20gb", 2)],
                spec + [(1, "1g.10gb", 1)], spec + [(1, "2g.40gb", 0)]):
        off, prof, status, spec_q, mask, ids = _flatten(state, bad, 2, profiles)
        assert not planner.diff_plan_mig_devices(off, prof, status, spec_q, mask)[3].all(), bad


def test_random_states_against_the_device_level_oracle(planner):
    rng = np.random.default_rng(12)
    profiles = ["1g.10gb", "1g.20gb", "2g.20gb", "3g.40gb", "4g.20gb", "7g.79gb"]
    for trial in range(60):
        G = int(rng.integers(1, 9))
        state, spec = {}, []
        n = 0
        for g in range(G):
            devs = []
            for _ in range(int(rng.integers(0, 8))):
                n += 1
                devs.append((f"{int(rng.integers(0, 10**6)):06d}-{n}", profiles[int(rng.integers(0, 6))], ["used", "free", "free", "unknown"][int(rng.integers(0, 4))]))
            if devs:
                state[g] = devs
            if rng.random() < 0.8:
                for p in rng.choice(6, size=int(rng.integers(0, 4)), replace=False):
                    spec.append((g, profiles[int(p)], int(rng.integers(0, 5))))
        off, prof, status, spec_q, mask, ids = _flatten(state, spec, G, profiles)
        op, cn, ca, m = planner.diff_plan_mig_devices(off, prof, status, spec_q, mask)
        creates, deletes = _operations(profiles, off, ids, op, cn, ca)
        ecr, edl = diff_plan.new_mig_config_plan(state, spec)
        assert creates == sorted(ecr), trial
        assert deletes == sorted(sorted((g, d) for (g, d, _, _) in o) for o in edl if o), trial
