"""Batch-demand planning mode — NEW BEHAVIOUR (the reference documents it, docs/en/docs/dynamic-gpu-partitioning/
configuration.md:6-11,19,38, but no longer has the code).  oracle/batch_plan.py is its definition; the CPU tests pin the
definition on small hand-checkable cases and its invariants, the -m gpu tests require nosplan_plan_batch to reproduce it
bit for bit (pods placed, GPUs chosen, nodes re-partitioned, the order kept, rounds, final state)."""
import numpy as np
import pytest

from kat_util import table
from oracle import batch_plan
from oracle import oracle as orc


def _pkg():
    import walkai_nos_b200 as pkg
    return pkg


def _a30_cluster(pkg, gpus):
    """gpus: per node a list of (used{}, free{}) on A30 GPUs; builtin table."""
    tables = table("builtin")
    names = sorted({p for geoms in tables.values() for g in geoms for p in g})
    model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
    off, used, free = [0], [], []
    for node in gpus:
        for (u, f) in node:
            used.append([u.get(n, 0) for n in names])
            free.append([f.get(n, 0) for n in names])
        off.append(len(used))
    snap = dict(n_nodes=len(gpus), gpu_off=np.asarray(off, np.int32), valid=np.ones(len(gpus), np.uint8),
                model=np.full(len(used), model_names.index("A30"), np.int32), used=np.asarray(used, np.int32).reshape(-1, len(names)),
                free=np.asarray(free, np.int32).reshape(-1, len(names)))
    snap["used_mask"] = pkg.synth._mask_of(snap["used"])
    snap["free_mask"] = pkg.synth._mask_of(snap["free"])
    return tables, names, (model_names, geom_off, geom_counts), snap


def test_definition_on_a_hand_checked_case():
    """Two empty A30 GPUs on two nodes, a free 4g.24gb on a third; pods: 4g.24gb x1, 1g.6gb x5.
    A: the 4g.24gb pod takes the free device of node 2.  B round 1: lacking {1g.6gb: 5}: node 0 -> {1g.6gb:4} (4 < 5),
    node 1 -> {1g.6gb:4} (covers); node 2's GPU holds a used 4g.24gb: no geometry with 1g.6gb keeps it.  All five pods
    are then placed: four on node 0, one on node 1.  Ascending order kept (nothing to gain from the other one)."""
    pkg = _pkg()
    tables, names, _, snap = _a30_cluster(pkg, [[({}, {})], [({}, {})], [({}, {"4g.24gb": 1})]])
    o = orc.Oracle(tables)
    k1, k4 = names.index("1g.6gb"), names.index("4g.24gb")
    r = batch_plan.batch_plan(o, "mig", snap, [k4] + [k1] * 5, [1] * 6)
    assert list(r["phase"]) == [1, 2, 2, 2, 2, 2]
    assert list(r["sched_node"]) == [2, 0, 0, 0, 0, 1]
    assert list(r["repartitioned"]) == [1, 1, 0] and r["order"] == 0 and r["rounds"] == 1
    assert r["snap"]["used"][0, k1] == 4 and r["snap"]["free"][1, k1] == 3 and r["snap"]["used"][2, k4] == 1


@pytest.mark.parametrize("kind", ["mig", "mps"])
def test_invariants_on_a_baseline_shard(kind):
    pkg = _pkg()
    cfg = pkg.synth.make_config(5, 128, 1500)
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0 if kind == "mig" else 1)
    M = len(off) - 1
    snap = cfg[kind]
    r = batch_plan.batch_plan(o, kind, snap, key[:M], qty[:M])
    fin = r["snap"]
    # used devices only grow (nothing in use is ever deleted), and by exactly what the placed pods asked for
    assert (fin["used"] >= snap["used"]).all()
    placed = r["sched_node"] >= 0
    want = np.zeros(snap["used"].shape[1], np.int64)
    np.add.at(want, key[:M][placed], qty[:M][placed])
    assert np.array_equal((fin["used"] - snap["used"]).sum(axis=0), want)
    # a pod placed in phase A sits where plain first-fit puts it; nodes not re-partitioned keep used + free
    same = r["repartitioned"] == 0
    per_gpu_same = np.repeat(same, np.diff(snap["gpu_off"]))
    assert np.array_equal((fin["used"] + fin["free"])[per_gpu_same], (snap["used"] + snap["free"])[per_gpu_same])
    assert (r["phase"][placed] > 0).all() and (r["phase"][~placed] == 0).all()
    # the mode never places fewer pods than first-fit alone, and does re-partition something on this cluster
    assert (r["phase"] == 2).sum() > 0 and r["repartitioned"].sum() > 0


# ---- GPU ---------------------------------------------------------------------------------------------------------
def _compare(pkg, p, o, kind, snap, key, qty, max_rounds=0):
    exp = batch_plan.batch_plan(o, kind, snap, key, qty, max_rounds=max_rounds or 8)
    got = p.plan_batch(0 if kind == "mig" else 1, key, qty, max_rounds)
    for k in ("phase", "sched_node", "sched_gpu", "repartitioned"):
        bad = np.nonzero(np.asarray(got[k]) != np.asarray(exp[k]))[0]
        assert bad.size == 0, (k, bad[:5], np.asarray(got[k])[bad[:5]], np.asarray(exp[k])[bad[:5]])
    assert got["order"] == exp["order"] and got["rounds"] == exp["rounds"]
    if kind == "mig":
        used, free = p.get_mig_state()
        assert np.array_equal(used, exp["snap"]["used"]) and np.array_equal(free, exp["snap"]["free"])
    else:
        used, free, fm = p.get_mps_state()
        assert np.array_equal(used, exp["snap"]["used"]) and np.array_equal(free, exp["snap"]["free"])
    return exp


@pytest.mark.gpu
def test_gpu_hand_checked_case():
    pkg = _pkg()
    tables, names, (model_names, geom_off, geom_counts), snap = _a30_cluster(pkg, [[({}, {})], [({}, {})], [({}, {"4g.24gb": 1})]])
    o = orc.Oracle(tables)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(names, geom_off, geom_counts)
        p.set_mig_snapshot(snap)
        k1, k4 = names.index("1g.6gb"), names.index("4g.24gb")
        _compare(pkg, p, o, "mig", snap, np.asarray([k4] + [k1] * 5, np.int32), np.ones(6, np.int32))
    finally:
        p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("cfgno,nodes,pods", [(2, 0, 0), (3, 64, 900), (4, 192, 2500), (5, 512, 6000)])
def test_gpu_reproduces_the_definition(cfgno, nodes, pods):
    pkg = _pkg()
    cfg = pkg.synth.make_config(cfgno, nodes, pods)
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        for kind, kd in (("mig", 0), ("mps", 1)):
            if cfg[kind] is None:
                continue
            if kd == 0:
                p.set_mig_snapshot(cfg[kind])
            else:
                p.set_mps_snapshot(cfg["sizes"], cfg[kind])
            off, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], kd)
            M = len(off) - 1
            exp = _compare(pkg, p, o, kind, cfg[kind], key[:M], qty[:M])
            assert (exp["phase"] == 2).any()
            # the snapshot is left as it was: the per-pod planner afterwards gives what it gives on a fresh handle
            _compare(pkg, p, o, kind, cfg[kind], key[:M], qty[:M], max_rounds=1)
            pods_k = dict(kind=np.full(M, kd, np.uint8), off=off, key=key, qty=qty)
            W = max(len(cfg["names"]), len(cfg["sizes"]))
            req, mask = pkg.synth.csr_to_dense(off, key, qty, W)
            e2, _, _ = o.plan(cfg["mig"] if kd == 0 else None, cfg["mps"] if kd == 1 else None, pods_k["kind"], req, mask, gate=True, sched=True)
            g2 = p.plan(pods_k["kind"], off, key, qty, gate=True, sched=True)
            assert np.array_equal(g2["sched_node"], e2["sched_node"])
    finally:
        p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [61, 62, 63])
def test_gpu_random_clusters(seed):
    import test_gpu_parity as T
    pkg = _pkg()
    cfg, _, rng = T._random_cluster(pkg, seed, n_mig=90, n_mps=70)
    p = pkg.Planner(0)
    try:
        o = T._load(p, cfg)
        for kind, kd, width in (("mig", 0, len(cfg["names"])), ("mps", 1, len(cfg["sizes"]))):
            M = 500
            key = rng.integers(0, width, size=M).astype(np.int32)
            qty = rng.choice([1, 1, 1, 2, 2, 3, 8], size=M).astype(np.int32)
            _compare(pkg, p, o, kind, cfg[kind], key, qty)
    finally:
        p.close()
