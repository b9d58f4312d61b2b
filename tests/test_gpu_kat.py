"""-m gpu: the reference's known-answer vectors driven through the C ABI of libnosplan.so on a B200,
checked against the golden expectations AND against the CPU oracle on the same inputs."""
import numpy as np
import pytest

from kat_util import (demand_csr, geometry_of, load_kats, mig_snapshot_from_nodes, mps_snapshot_from_nodes, table)
from oracle import oracle as orc

pytestmark = pytest.mark.gpu
K = load_kats()


@pytest.fixture(scope="module")
def planner():
    import walkai_nos_b200 as pkg
    p = pkg.Planner(0)
    yield p
    p.close()


def _names(tname, *dicts):
    s = set(orc.BUILTIN_MIG_PROFILES)
    for geoms in table(tname).values():
        for g in geoms:
            s |= set(g)
    for d in dicts:
        s |= set(d)
    return sorted(s)


def _setup_mig(planner, tname, names):
    from walkai_nos_b200 import synth
    t = table(tname)
    model_names, off, counts = synth.flatten_tables(t, names)
    planner.set_mig_tables(names, off, counts)
    return t, model_names


def _apply_choice(t, model, used, free, ch):
    if ch < 0:
        return dict(free)
    cand = t[model][ch]
    return {k: q - used.get(k, 0) for k, q in cand.items()}


@pytest.mark.parametrize("c", K["mig_gpu_update"], ids=lambda c: c["src"])
def test_mig_gpu_update(planner, c):
    names = _names(c["table"], c["used"], c["free"], c["req"])
    t, model_names = _setup_mig(planner, c["table"], names)
    snap = mig_snapshot_from_nodes([[(c["model"], c["used"], c["free"])]], names, model_names)
    planner.set_mig_snapshot(snap)
    off, key, qty = demand_csr([c["req"]], names.index)
    r = planner.score_mig(off, key, qty)
    assert bool(r["updated"][0, 0]) == c["updated"]
    assert (r["first_node"][0] == 0) == c["updated"]
    free = _apply_choice(t, c["model"], c["used"], c["free"], int(r["choice"][0, 0]))
    assert geometry_of(c["used"], free) == c["geometry"]
    assert int(r["first_choice"][0, 0]) == int(r["choice"][0, 0])


@pytest.mark.parametrize("c", K["mig_node_update"], ids=lambda c: c["src"])
def test_mig_node_update(planner, c):
    names = _names("builtin", c["req"])
    t, model_names = _setup_mig(planner, "builtin", names)
    gpus = [(g["model"], g["used"], g["free"]) for g in c["gpus"]]
    snap = mig_snapshot_from_nodes([gpus], names, model_names)
    planner.set_mig_snapshot(snap)
    off, key, qty = demand_csr([c["req"]], names.index)
    r = planner.score_mig(off, key, qty)
    assert bool(r["updated"][0, 0]) == c["updated"]
    out = []
    for i, (m, u, f) in enumerate(gpus):
        out.append((m, u, _apply_choice(t, m, u, f, int(r["choice"][0, i]))))
    assert orc.Oracle.node_geometry(out) == c["geometry"]
    # and the planner path (gate off, no scheduling) commits the same geometry
    kind = np.zeros(1, np.uint8)
    pr = planner.plan(kind, off, key, qty, gate=False, sched=False)
    assert (pr["repart_node"][0] == 0) == c["updated"]
    used, free = planner.get_mig_state()
    got = {}
    for g in range(len(gpus)):
        for p, n in enumerate(names):
            q = int(used[g, p] + free[g, p])
            if q > 0:
                got[n] = got.get(n, 0) + q
    assert got == {k: v for k, v in c["geometry"].items() if v > 0}


@pytest.mark.parametrize("c", K["mig_gpu_init"], ids=lambda c: c["src"])
def test_mig_gpu_init(planner, c):
    names = _names(c["table"], c["used"], c["free"])
    t, model_names = _setup_mig(planner, c["table"], names)
    planner.set_mig_snapshot(mig_snapshot_from_nodes([[(c["model"], c["used"], c["free"])]], names, model_names))
    err, free = planner.init_geometry_mig()
    assert bool(err[0]) == c["err"]
    if not c["err"]:
        f = {n: int(free[0, p]) for p, n in enumerate(names) if free[0, p] > 0}
        assert geometry_of(c["used"], f) == c["geometry"]


@pytest.mark.parametrize("c", K["mig_gpu_add_pod"], ids=lambda c: c["src"])
def test_mig_gpu_add_pod(planner, c):
    names = _names("builtin", c["req"])
    t, model_names = _setup_mig(planner, "builtin", names)
    planner.set_mig_snapshot(mig_snapshot_from_nodes([[("A30", c["used"], c["free"])]], names, model_names))
    off, key, qty = demand_csr([c["req"]], names.index)
    pr = planner.plan(np.zeros(1, np.uint8), off, key, qty, gate=True, sched=True)
    assert (pr["sched_node"][0] < 0) == c["err"]
    used, free = planner.get_mig_state()
    assert {n: int(used[0, p]) for p, n in enumerate(names) if used[0, p]} == {k: v for k, v in c["exp_used"].items() if v}
    assert {n: int(free[0, p]) for p, n in enumerate(names) if free[0, p]} == {k: v for k, v in c["exp_free"].items() if v}


@pytest.mark.parametrize("c", K["mps_gpu_update"], ids=lambda c: c["src"])
def test_mps_gpu_update(planner, c):
    sz = orc.Oracle.mps_size
    sizes = sorted({sz(k) for d in (c["used"], c["free"], c["req"]) for k in d} | {1})
    snap = mps_snapshot_from_nodes([[(c["mem"], c["used"], c["free"])]], sizes, sz)
    planner.set_mps_snapshot(sizes, snap)
    off, key, qty = demand_csr([c["req"]], lambda k: sizes.index(sz(k)))
    r = planner.score_mps(off, key, qty)
    assert bool(r["updated"][0, 0]) == c["updated"]
    free = {f"{sizes[s]}gb": int(r["out_free"][0, 0, s]) for s in range(len(sizes)) if (int(r["out_free_mask"][0, 0]) >> s) & 1}
    assert geometry_of(c["used"], free) == c["geometry"]


@pytest.mark.parametrize("c", K["mps_node_update"], ids=lambda c: c["src"])
def test_mps_node_update(planner, c):
    sz = orc.Oracle.mps_size
    gpus = orc.extract_gpus(c["labels"], c["annotations"], None, "mps")
    sizes = sorted({sz(k) for g in gpus for d in (g["used"], g["free"]) for k in d} | {sz(k) for k in c["req"]} | {1})
    snap = mps_snapshot_from_nodes([[(g["mem_gb"], g["used"], g["free"]) for g in gpus]], sizes, sz)
    planner.set_mps_snapshot(sizes, snap)
    off, key, qty = demand_csr([c["req"]], lambda k: sizes.index(sz(k)))
    r = planner.score_mps(off, key, qty)
    assert bool(r["updated"][0, 0]) == c["updated"]


@pytest.mark.parametrize("c", K["mps_node_add_pod"] + K["mig_node_add_pod"], ids=lambda c: c["src"])
def test_node_add_pod(planner, c):
    is_mps = "slicing" in c["src"]
    if is_mps:
        sz = orc.Oracle.mps_size
        gpus = orc.extract_gpus(c["labels"], c["annotations"], None, "mps")
        sizes = sorted({sz(k) for g in gpus for d in (g["used"], g["free"]) for k in d} | {sz(k) for k in c["req"]})
        planner.set_mps_snapshot(sizes, mps_snapshot_from_nodes([[(g["mem_gb"], g["used"], g["free"]) for g in gpus]], sizes, sz))
        off, key, qty = demand_csr([c["req"]], lambda k: sizes.index(sz(k)))
        pr = planner.plan(np.ones(1, np.uint8), off, key, qty)
        used, free, _ = planner.get_mps_state()
        names = [f"{s}gb" for s in sizes]
    else:
        names = _names("builtin", c["req"])
        t, model_names = _setup_mig(planner, "builtin", names)
        gpus = orc.extract_gpus(c["labels"], c["annotations"], t.keys(), "mig")
        planner.set_mig_snapshot(mig_snapshot_from_nodes([[(g["model"], g["used"], g["free"]) for g in gpus]], names, model_names))
        off, key, qty = demand_csr([c["req"]], names.index)
        pr = planner.plan(np.zeros(1, np.uint8), off, key, qty)
        used, free = planner.get_mig_state()
    assert (pr["sched_node"][0] < 0) == c["err"]
    assert pr["sched_gpu"][0] == 0  # the annotated GPU comes first in the canonical order
    us = {n: int(used[:, p].sum()) for p, n in enumerate(names) if used[:, p].sum()}
    fs = {n: int(free[:, p].sum()) for p, n in enumerate(names) if free[:, p].sum()}
    assert us == c["used_sum"] and fs == c["free_sum"]
