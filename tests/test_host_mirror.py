"""CPU tests of libnoshost.so (the host side of the drop-in) against the reference's own string-level
known-answer vectors, and against the oracle's independent restatement of NewNode/extractGPUs."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from kat_util import load_kats, table
from oracle import oracle as orc

K = load_kats()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    import walkai_nos_b200 as pkg
    pkg.build_library()
    return pkg.host


def test_exports_every_declared_symbol(host):
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "noshost.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(noshost_[a-z0-9_]+)\s*\(", src)))
    L = host.lib()
    for n in names:
        assert hasattr(L, n), n
    assert sorted(host.EXPORTS) == names


@pytest.mark.parametrize("c", K["parse_status_annotation"], ids=lambda c: c["src"])
def test_parse_status_annotation(host, c):
    a = host.parse_status_annotation(c["key"], c["value"])
    assert (a is None) == c["err"]
    if not c["err"]:
        assert a == c["expected"]
        assert host.format_status_annotation(a["index"], a["profile"], a["status"]) == c["key"]


@pytest.mark.parametrize("c", K["parse_spec_annotation"], ids=lambda c: c["src"])
def test_parse_spec_annotation(host, c):
    a = host.parse_spec_annotation(c["key"], c["value"])
    assert (a is None) == c["err"]
    if not c["err"]:
        assert a == c["expected"]
        assert host.format_spec_annotation(a["index"], a["profile"]) == c["key"]


@pytest.mark.parametrize("c", K["memory_gb"], ids=lambda c: c["src"])
def test_memory_gb(host, c):
    v = host.memory_gb(c["label"])
    assert (v is None) == c["err"]
    if not c["err"]:
        assert v == c["value"]


def test_resource_name_extraction(host):
    # pkg/gpu/mig/util.go:46-52 and pkg/gpu/slicing/util.go:34-40 (doc examples of the reference)
    assert host.mig_profile_of_resource("nvidia.com/mig-1g.10gb") == "1g.10gb"
    assert host.mig_profile_of_resource("nvidia.com/gpu") is None
    assert host.mig_profile_of_resource("nvidia.com/gpu-10gb") is None
    assert host.mps_profile_of_resource("nvidia.com/gpu-10gb") == "10gb"
    assert host.mps_profile_of_resource("nvidia.com/gpu") is None
    assert host.mps_profile_size_gb("10gb") == 10 and host.mps_profile_size_gb("nvidia.com/gpu-20gb") == 20
    assert host.mps_profile_size_gb("foo") == 0


def _as_gpus(host, snap, kind):
    a = snap.arrays()
    out = []
    for n in range(snap.n_nodes):
        if not a["valid"][n]:
            out.append(None)
            continue
        gpus = []
        for g in range(a["gpu_off"][n], a["gpu_off"][n + 1]):
            used = {snap.key_names[k]: int(a["used"][g, k]) for k in range(snap.K) if (int(a["used_mask"][g]) >> k) & 1}
            free = {snap.key_names[k]: int(a["free"][g, k]) for k in range(snap.K) if (int(a["free_mask"][g]) >> k) & 1}
            d = dict(index=int(a["gpu_index"][g]), used=used, free=free)
            if kind == "mps":
                d["mem_gb"] = int(a["mem_gb"][g])
            gpus.append(d)
        out.append(gpus)
    return out


@pytest.mark.parametrize("c", K["mig_new_node"], ids=lambda c: c["src"])
def test_mig_new_node(host, c):
    snap = host.Snapshot("mig", [dict(labels=c["labels"], annotations=c["annotations"])], table("builtin").keys())
    got = _as_gpus(host, snap, "mig")[0]
    assert (got is None) == c["err"]
    if not c["err"]:
        exp = [dict(index=g["index"], used=g["used"], free=g["free"]) for g in c["gpus"]]
        assert got == exp


@pytest.mark.parametrize("c", K["mps_new_node"], ids=lambda c: c["src"])
def test_mps_new_node(host, c):
    snap = host.Snapshot("mps", [dict(labels=c["labels"], annotations=c["annotations"])])
    got = _as_gpus(host, snap, "mps")[0]
    assert (got is None) == c["err"]
    if not c["err"]:
        assert got == [dict(index=g["index"], used=g["used"], free=g["free"], mem_gb=g["mem_gb"]) for g in c["gpus"]]


def test_duplicate_index_quirk_and_bulk_agrees_with_oracle(host):
    """node.go:89-96: missing GPUs get index len..count-1 (indices 1,1,2 in node_test.go:527-537); a bulk
    build over many nodes must agree node by node with the oracle's restatement."""
    rng = np.random.default_rng(7)
    profiles = ["1g.10gb", "2g.20gb", "3g.40gb", "7g.79gb"]
    nodes = []
    for n in range(40):
        labels = {"nvidia.com/gpu.product": "NVIDIA-A100-80GB-PCIe" if rng.random() < 0.85 else "mystery",
                  "nvidia.com/gpu.count": str(int(rng.integers(0, 6)))}
        if rng.random() < 0.1:
            del labels["nvidia.com/gpu.count"]
        ann = {"unrelated/annotation": "x", "nos.nebuly.com/spec-partitioning-plan": "123"}
        for _ in range(int(rng.integers(0, 7))):
            idx = int(rng.integers(0, 6))
            p = profiles[int(rng.integers(0, 4))]
            st = ["free", "used", "Free", "unknown"][int(rng.integers(0, 4))]
            ann[f"nos.nebuly.com/status-gpu-{idx}-{p}-{st}"] = str(int(rng.integers(0, 4)))
        ann[f"nos.nebuly.com/spec-gpu-0-{profiles[0]}"] = "2"
        nodes.append(dict(name=f"node-{n:03d}", labels=labels, annotations=ann))
    snap = host.Snapshot("mig", nodes, table("builtin").keys())
    got = _as_gpus(host, snap, "mig")
    for n, node in enumerate(nodes):
        exp = orc.extract_gpus(node["labels"], node["annotations"], table("builtin").keys(), "mig")
        if exp is None:
            assert got[n] is None
        else:
            assert got[n] == [dict(index=g["index"], used=g["used"], free=g["free"]) for g in exp]
    assert snap.key_names == sorted(snap.key_names)  # canonical vocabulary order


def test_pods_filter_order_and_requests(host):
    mig = host.Snapshot("mig", [dict(labels={"nvidia.com/gpu.product": "A30", "nvidia.com/gpu.count": "1"})],
                        table("builtin").keys(), extra_names=["1g.10gb", "2g.20gb"])
    mps = host.Snapshot("mps", [dict(labels={"nvidia.com/gpu.product": "A30", "nvidia.com/gpu.count": "1",
                                              "nvidia.com/gpu.memory": "40000"})], extra_sizes=[10, 20])
    M1, M2, S1 = "nvidia.com/mig-1g.10gb", "nvidia.com/mig-2g.20gb", "nvidia.com/gpu-10gb"
    pods = [
        dict(name="ns/a", priority=0, created=5, containers=[{M1: 1, "cpu": 2}]),
        dict(name="ns/b", priority=100, created=9, containers=[{M1: 1}, {M1: 2}], init_containers=[{M1: 5}, {M2: 1}]),
        dict(name="ns/c", priority=100, created=3, containers=[{S1: 1}]),
        dict(name="ns/d", priority=100, created=3, containers=[{M2: 1, S1: 2}]),   # both kinds -> two rows
        dict(name="ns/e", phase="Running", containers=[{M1: 1}]),                 # not pending
        dict(name="ns/f", node_name="n1", containers=[{M1: 1}]),                  # already scheduled
        dict(name="ns/g", unschedulable=False, containers=[{M1: 1}]),             # no Unschedulable condition
        dict(name="ns/h", priority=0, created=1, containers=[{"cpu": 1}]),         # no GPU request -> dropped
    ]
    r = host.build_pods(pods, mig, mps)
    k1, k2 = mig.key_names.index("1g.10gb"), mig.key_names.index("2g.20gb")
    s10 = mps.key_names.index("10gb")
    rows = []
    for i in range(len(r["source"])):
        rows.append((int(r["source"][i]), int(r["kind"][i]),
                     [(int(r["key"][e]), int(r["qty"][e])) for e in range(r["off"][i], r["off"][i + 1])]))
    assert rows == [
        (2, 1, [(s10, 1)]),                       # prio 100, created 3, "ns/c" < "ns/d"
        (3, 0, [(k2, 1)]), (3, 1, [(s10, 2)]),
        (1, 0, sorted([(k1, 5), (k2, 1)])),       # max(sum of containers = 3, init max = 5); init-only key kept
        (0, 0, [(k1, 1)]),
    ]
    with pytest.raises(ValueError):
        host.build_pods([dict(name="x", containers=[{"nvidia.com/mig-4g.40gb": 1}])], mig, mps)


@pytest.mark.parametrize("c", K["partitioning_state_equal"], ids=lambda c: c["src"])
def test_partitioning_state_equal(host, c):
    assert host.partitioning_state_equal(c["a"], c["b"]) == c["expected"]


def test_spec_annotations_from_state(host):
    nodes = [dict(name="n0", labels={"nvidia.com/gpu.product": "A30", "nvidia.com/gpu.count": "3"},
                  annotations={"nos.nebuly.com/status-gpu-1-1g.6gb-free": "2", "nos.nebuly.com/status-gpu-1-2g.12gb-used": "1"})]
    snap = host.Snapshot("mig", nodes, table("builtin").keys())
    a = snap.arrays()
    ann = snap.spec_annotations(0, a["used"], a["free"])
    # GPU order 1 (annotated), then 1, 2 (the quirk); only quantities > 0 appear (state.go:33-35)
    assert ann == {"nos.nebuly.com/spec-gpu-1-1g.6gb": "2", "nos.nebuly.com/spec-gpu-1-2g.12gb": "1"}
