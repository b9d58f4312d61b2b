"""-m gpu: BASELINE configs 3 and 4 at their NAMED sizes (256 MPS nodes x 10 000 pods; 1024 mixed nodes x
20 000 pods), bit-exact against the oracle.  The scorers run on every pod; the oracle scores one row per
distinct request vector (identical rows must give identical answers, which is asserted) and its full
per-node `updated` row is compared for the first and the last pod of every vector.  The planner runs on
all pods with the gate on (reference behaviour, mig_controller.go:121-129) and off, first-fit included,
and the final state is compared."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import walkai_nos_b200 as p
    return p


@pytest.fixture(scope="module")
def planner(pkg):
    p = pkg.Planner(0)
    yield p
    p.close()


def _dedup(key, qty):
    sig = key.astype(np.int64) * 1000003 + qty
    uniq, first_idx = np.unique(sig, return_index=True)
    return sig, uniq, first_idx


def _scorers_at_full_size(pkg, planner, o, cfg):
    pods = cfg["pods"]
    if cfg["mig"] is not None:
        off, key, qty, sel = pkg.synth.pods_of_kind(pods, 0)
        assert np.all(np.diff(off) == 1)
        sig, uniq, fi = _dedup(key, qty)
        req, mask = pkg.synth.csr_to_dense(np.arange(len(uniq) + 1, dtype=np.int32), key[fi], qty[fi], len(cfg["names"]))
        eu, ec, ef = o.score_mig(cfg["mig"], req, mask, threads=8)
        r = planner.score_mig(off, key, qty, want_updated=True, want_choice=False, want_first_choice=True)
        gs = r["first_choice"].shape[1]
        for j, s in enumerate(uniq):
            rows = np.nonzero(sig == s)[0]
            assert np.all(r["first_node"][rows] == ef[j])
            assert np.array_equal(r["updated"][rows[0]], eu[j]) and np.array_equal(r["updated"][rows[-1]], eu[j])
            if ef[j] >= 0:
                g0 = cfg["mig"]["gpu_off"][ef[j]]
                assert np.all(r["first_choice"][rows, :gs] == ec[j, g0:g0 + gs])
        assert r["n_evals"] == len(sel) * planner.mig_evals_per_demand()
    if cfg["mps"] is not None:
        off, key, qty, sel = pkg.synth.pods_of_kind(pods, 1)
        sig, uniq, fi = _dedup(key, qty)
        S = len(cfg["sizes"])
        req, mask = pkg.synth.csr_to_dense(np.arange(len(uniq) + 1, dtype=np.int32), key[fi], qty[fi], S)
        eu, efree, emask, ef = o.score_mps(cfg["mps"], req, mask, threads=8)
        r = planner.score_mps(off, key, qty, want_updated=True, want_free=False)
        for j, s in enumerate(uniq):
            rows = np.nonzero(sig == s)[0]
            assert np.all(r["first_node"][rows] == ef[j])
            assert np.array_equal(r["updated"][rows[0]], eu[j]) and np.array_equal(r["updated"][rows[-1]], eu[j])
        # the full layout table (the general kernel) on the distinct vectors
        uo = np.arange(len(uniq) + 1, dtype=np.int32)
        rf = planner.score_mps(uo, key[fi], qty[fi], want_updated=True, want_free=True)
        assert np.array_equal(rf["out_free"].astype(np.int32), efree) and np.array_equal(rf["out_free_mask"], emask)
        assert np.array_equal(rf["first_node"], ef)


def _plan_at_full_size(pkg, planner, o, cfg, gate):
    pods = cfg["pods"]
    W = max(len(cfg["names"]), len(cfg["sizes"]))
    req, mask = pkg.synth.csr_to_dense(pods["off"], pods["key"], pods["qty"], W)
    exp, ms, ps = o.plan(cfg["mig"], cfg["mps"], pods["kind"], req, mask, gate=gate, sched=True)
    got = planner.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=gate, sched=True)
    for k in ("gate_hit", "repart_node", "sched_node", "sched_gpu"):
        bad = np.nonzero(got[k] != exp[k])[0]
        assert bad.size == 0, f"{k}: first mismatch at pod {bad[:5]} got {got[k][bad[:5]]} exp {exp[k][bad[:5]]}"
    if cfg["mig"] is not None:
        used, free = planner.get_mig_state()
        assert np.array_equal(used, ms["used"]) and np.array_equal(free, ms["free"])
    if cfg["mps"] is not None:
        used, free, fm = planner.get_mps_state()
        assert np.array_equal(used, ps["used"]) and np.array_equal(free, ps["free"]) and np.array_equal(fm, ps["free_mask"])
    return exp


def _load(planner, cfg):
    planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    if cfg["mig"] is not None:
        planner.set_mig_snapshot(cfg["mig"])
    if cfg["mps"] is not None:
        planner.set_mps_snapshot(cfg["sizes"], cfg["mps"])
    return orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])


@pytest.mark.parametrize("cfgno,nodes,pods", [(3, 256, 10000), (4, 1024, 20000)])
def test_scorers_at_named_size(pkg, planner, cfgno, nodes, pods):
    cfg = pkg.synth.make_config(cfgno)
    n = (cfg["mig"]["n_nodes"] if cfg["mig"] is not None else 0) + (cfg["mps"]["n_nodes"] if cfg["mps"] is not None else 0)
    assert n == nodes and len(cfg["pods"]["kind"]) == pods
    o = _load(planner, cfg)
    _scorers_at_full_size(pkg, planner, o, cfg)


@pytest.mark.parametrize("gate", [True, False])
@pytest.mark.parametrize("cfgno", [3, 4])
def test_plan_at_named_size(pkg, planner, cfgno, gate):
    cfg = pkg.synth.make_config(cfgno)
    o = _load(planner, cfg)
    exp = _plan_at_full_size(pkg, planner, o, cfg, gate)
    if not gate:
        assert (exp["repart_node"] >= 0).any()  # the general path really repartitioned something
