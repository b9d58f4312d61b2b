"""No GPU: the hardware counters bench.py quotes (profiles/kernel_counters.json) must have been measured on the
kernels that are in the tree.  The file carries the hash of csrc/ at capture time; a kernel edit without a new
ncu capture fails here (re-capture with the recipe in profiles/README.md, or delete the file: bench.py then
prints the roofline without a fraction instead of a stale one)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_kernel_counters_match_the_source_in_the_tree():
    import ncu_counters
    if not os.path.exists(ncu_counters.PATH):
        return
    db = json.load(open(ncu_counters.PATH))
    assert db["csrc_sha"] == ncu_counters.csrc_sha(), (
        "walkai-nos_b200/csrc changed after profiles/kernel_counters.json was captured: re-run the ncu capture")
    for e in db["entries"]:
        assert e["kernels"]
        for k in e["kernels"].values():
            assert k["inst_executed"] > 0 and k["duration_ns"] > 0


def test_lookup_refuses_counters_of_other_sources(tmp_path, monkeypatch):
    import ncu_counters
    p = tmp_path / "c.json"
    p.write_text(json.dumps({"csrc_sha": "0" * 16, "entries": [{"config": 5, "nodes": 4096, "pods": 50000, "world": 1, "kernels": {}}]}))
    monkeypatch.setattr(ncu_counters, "PATH", str(p))
    assert ncu_counters.lookup(5, 4096, 50000, 1) is None
    p.write_text(json.dumps({"csrc_sha": ncu_counters.csrc_sha(), "entries": [{"config": 5, "nodes": 4096, "pods": 50000, "world": 1, "kernels": {}}]}))
    assert ncu_counters.lookup(5, 4096, 50000, 1) is not None
    assert ncu_counters.lookup(5, 4096, 50000, 2) is None
