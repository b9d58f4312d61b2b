"""-m gpu: the whole drop-in, text to text: node labels/annotations + pod specs -> libnoshost -> libnosplan
(B200) -> planned state -> spec annotations; compared with the oracle driven from the same text."""
import numpy as np
import pytest

from kat_util import table
from oracle import oracle as orc

pytestmark = pytest.mark.gpu


def test_text_snapshot_to_spec_annotations():
    import walkai_nos_b200 as pkg
    host = pkg.host
    rng = np.random.default_rng(3)
    tables = table("yaml")
    model = "NVIDIA-H100-80GB"
    nodes = []
    for n in range(24):
        ann = {}
        for g in range(8):
            geom = tables[model][int(rng.integers(0, 12))]
            for p, q in geom.items():
                u = int(rng.integers(0, q + 1))
                if u:
                    ann[f"nos.nebuly.com/status-gpu-{g}-{p}-used"] = str(u)
                if q - u:
                    ann[f"nos.nebuly.com/status-gpu-{g}-{p}-free"] = str(q - u)
        nodes.append(dict(name=f"node-{n:04d}", labels={"nvidia.com/gpu.product": model, "nvidia.com/gpu.count": "8"},
                          annotations=ann))
    profs = ["1g.10gb", "2g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]
    pods = [dict(name=f"ns/p{i:04d}", priority=int(rng.choice([0, 100])), created=i,
                 containers=[{f"nvidia.com/mig-{profs[int(rng.integers(0, 5))]}": int(rng.integers(1, 3))}])
            for i in range(300)]
    names = sorted({p for geoms in tables.values() for g in geoms for p in g})
    snap = host.Snapshot("mig", nodes, tables.keys(), extra_names=names)
    rows = host.build_pods(pods, snap, None)
    a = snap.arrays()
    assert snap.key_names == names
    model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
    assert model_names == list(tables.keys())
    p = pkg.Planner(0)
    try:
        p.set_mig_tables(names, geom_off, geom_counts)
        p.set_mig_snapshot(a)
        for gate in (True, False):
            got = p.plan(rows["kind"], rows["off"], rows["key"], rows["qty"], gate=gate, sched=True)
            used, free = p.get_mig_state()
            o = orc.Oracle(tables)
            assert o.mig_names == names
            req, mask = pkg.synth.csr_to_dense(rows["off"], rows["key"], rows["qty"], len(names))
            exp, ms, _ = o.plan(a, None, rows["kind"], req, mask, gate=gate, sched=True)
            for k in exp:
                assert np.array_equal(got[k], exp[k]), k
            assert np.array_equal(used, ms["used"]) and np.array_equal(free, ms["free"])
            # the wire format out: spec annotations of every repartitioned node
            for n in sorted(set(int(x) for x in got["repart_node"] if x >= 0)):
                ann = snap.spec_annotations(n, used, free)
                geo = {}
                for g in range(a["gpu_off"][n], a["gpu_off"][n + 1]):
                    for k, name in enumerate(names):
                        q = int(ms["used"][g, k] + ms["free"][g, k])
                        if q > 0:
                            geo[f"nos.nebuly.com/spec-gpu-{int(a['gpu_index'][g])}-{name}"] = str(q)
                assert ann == geo
                for key, val in ann.items():
                    parsed = host.parse_spec_annotation(key, val)
                    assert parsed is not None and parsed["quantity"] == int(val)
            if not gate:
                assert (got["repart_node"] >= 0).any()
    finally:
        p.close()
