"""Cluster-level known answers (tests/golden/controller_kats.json): hand-derived from
internal/controllers/gpupartitioner/mig_controller.go:56-198, the package the reference has no test for.
Each case is text (node labels + status annotations, pod specs), goes through the host mirror (libnoshost:
NewNode/extractGPUs, shouldConsiderPod, GetRequestedProfiles) and then
  * the CPU oracle's replay of Reconcile (no GPU needed), and
  * nosplan_plan on the GPU through the C ABI (-m gpu),
and must give the expected gate / repartitioned node per pod and the expected geometry of every GPU."""
import json
import os

import numpy as np
import pytest

from kat_util import GOLDEN, table
from oracle import oracle as orc

KATS = json.load(open(os.path.join(GOLDEN, "controller_kats.json")))
CASES = KATS["cases"]


def _flatten(case):
    import walkai_nos_b200 as pkg
    host = pkg.host
    tables = table(KATS["table"])
    names = sorted({p for geoms in tables.values() for g in geoms for p in g})
    snap = host.Snapshot("mig", case["nodes"], tables.keys(), extra_names=names)
    arrays = snap.arrays()
    assert snap.key_names == names
    rows = host.build_pods(case["pods"], snap, None)
    return pkg, tables, names, snap, arrays, rows


def _check(case, names, arrays, rows, got, used, free):
    # per pod: considered <=> it has a planner row (shouldConsiderPod + a non-empty request)
    row_of = {int(src): j for j, src in enumerate(rows["source"])}
    for i, exp in enumerate(case["expect"]):
        if not exp["considered"] or exp.get("gate_hit") is None:
            assert i not in row_of, f"pod {i} must not reach the planner"
            continue
        j = row_of[i]
        assert bool(got["gate_hit"][j]) == exp["gate_hit"], (case["name"], i)
        assert int(got["repart_node"][j]) == exp["repart_node"], (case["name"], i, int(got["repart_node"][j]))
    # final geometry of every GPU (values; a key with value 0 is as good as absent for the plan output)
    off = arrays["gpu_off"]
    assert len(case["final"]) == arrays["n_nodes"]
    for n, gpus in enumerate(case["final"]):
        assert off[n + 1] - off[n] == len(gpus), (case["name"], n)
        for s, g in enumerate(gpus):
            for what, arr in (("used", used), ("free", free)):
                want = np.zeros(len(names), np.int32)
                for k, q in g[what].items():
                    want[names.index(k)] = q
                assert np.array_equal(arr[off[n] + s], want), (case["name"], n, s, what, arr[off[n] + s], want)


def test_there_are_enough_cases_and_each_cites_the_controller():
    assert len(CASES) >= 12
    for c in CASES:
        assert c["derivation"] and any(":" in line for line in c["derivation"])
    text = json.dumps(KATS)
    for needle in (":68-71", ":78-81", ":83-86", ":135-137", ":153-157", ":159-162", ":121-129"):
        assert needle in text, needle


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_controller_kat_oracle(case):
    pkg, tables, names, snap, arrays, rows = _flatten(case)
    o = orc.Oracle(tables)
    assert o.mig_names == names
    M = len(rows["kind"])
    req, mask = pkg.synth.csr_to_dense(rows["off"], rows["key"], rows["qty"], len(names))
    if arrays["n_nodes"] == 0 or M == 0:
        got = dict(gate_hit=np.zeros(M, np.uint8), repart_node=np.full(M, -1, np.int32))  # :78-81 / nothing to replay
        used, free = arrays["used"], arrays["free"]
    else:
        got, ms, _ = o.plan(arrays, None, rows["kind"], req, mask, gate=True, sched=False)
        used, free = ms["used"], ms["free"]
    _check(case, names, arrays, rows, got, used, free)


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_controller_kat_gpu(case):
    pkg, tables, names, snap, arrays, rows = _flatten(case)
    model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(names, geom_off, geom_counts)
        p.set_mig_snapshot(arrays)
        got = p.plan(rows["kind"], rows["off"], rows["key"], rows["qty"], gate=True, sched=False)
        if arrays["gpu_off"][-1] > 0:
            used, free = p.get_mig_state()
        else:
            used, free = arrays["used"], arrays["free"]
        _check(case, names, arrays, rows, got, used, free)
        # and with the first-fit simulation on: the gate and the repartition decisions are the same
        got2 = p.plan(rows["kind"], rows["off"], rows["key"], rows["qty"], gate=True, sched=True)
        assert np.array_equal(got2["gate_hit"], got["gate_hit"])
    finally:
        p.close()
