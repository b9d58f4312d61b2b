"""Pins the CPU oracle against every known-answer vector the reference's own tests hold for the
hot path.  No GPU needed."""
import re

import pytest

from kat_util import geometry_of, load_kats, table
from oracle import oracle as orc

K = load_kats()
EXTRA = ["1gb"]  # never used as a MIG profile; keeps vocab building uniform


def _oracle(tname="builtin", extra=(), sizes=()):
    return orc.Oracle(table(tname), extra_mig_profiles=extra, mps_sizes=sizes)


def _names(*dicts):
    s = set()
    for d in dicts:
        s |= set(d.keys())
    return s


def _sizes(*dicts):
    return {orc.Oracle.mps_size(k) for d in dicts for k in d}


@pytest.mark.parametrize("c", K["mig_gpu_update"], ids=lambda c: c["src"])
def test_mig_gpu_update(c):
    o = _oracle(c["table"], _names(c["used"], c["free"], c["req"]))
    upd, free, _ = o.mig_gpu_update(c["model"], c["used"], c["free"], c["req"])
    assert upd == c["updated"]
    assert geometry_of(c["used"], free) == c["geometry"]


@pytest.mark.parametrize("c", K["mig_gpu_apply"], ids=lambda c: c["src"])
def test_mig_gpu_apply(c):
    o = _oracle(c["table"], _names(c["used"], c["free"], c["geometry"]))
    err, free = o.mig_gpu_apply(c["model"], c["used"], c["free"], c["geometry"])
    assert err == c["err"]
    assert free == c["exp_free"]


@pytest.mark.parametrize("c", K["mig_gpu_init"], ids=lambda c: c["src"])
def test_mig_gpu_init(c):
    o = _oracle(c["table"], _names(c["used"], c["free"]))
    err, free = o.mig_gpu_init(c["model"], c["used"], c["free"])
    assert err == c["err"]
    if not err:
        assert geometry_of(c["used"], free) == c["geometry"]


@pytest.mark.parametrize("c", K["mig_gpu_add_pod"], ids=lambda c: c["src"])
def test_mig_gpu_add_pod(c):
    o = _oracle("builtin")
    err, used, free = o.mig_gpu_add_pod(c["used"], c["free"], c["req"])
    assert err == c["err"]
    assert used == c["exp_used"]
    assert free == c["exp_free"]  # includes the present-but-zero key of gpu_test.go:162-165


@pytest.mark.parametrize("c", K["mig_node_update"], ids=lambda c: c["src"])
def test_mig_node_update(c):
    o = _oracle("builtin")
    gpus = [(g["model"], g["used"], g["free"]) for g in c["gpus"]]
    upd, out, _ = o.mig_node_update(gpus, c["req"])
    assert upd == c["updated"]
    assert o.node_geometry(out) == c["geometry"]


@pytest.mark.parametrize("c", K["mig_node_geometry"], ids=lambda c: c["src"])
def test_mig_node_geometry(c):
    gpus = [(g["model"], g["used"], g["free"]) for g in c["gpus"]]
    assert orc.Oracle.node_geometry(gpus) == c["geometry"]


@pytest.mark.parametrize("c", K["mig_node_has_free_capacity"], ids=lambda c: c["src"])
def test_mig_node_has_free_capacity(c):
    o = _oracle("builtin")
    gpus = [(g["model"], g["used"], g["free"]) for g in c["gpus"]]
    assert o.mig_node_has_free_capacity(gpus) == c["expected"]


@pytest.mark.parametrize("c", K["mig_new_node"], ids=lambda c: c["src"])
def test_mig_new_node(c):
    gpus = orc.extract_gpus(c["labels"], c["annotations"], table("builtin").keys(), "mig")
    assert (gpus is None) == c["err"]
    if not c["err"]:
        assert gpus == c["gpus"]


@pytest.mark.parametrize("c", K["mig_node_add_pod"], ids=lambda c: c["src"])
def test_mig_node_add_pod(c):
    o = _oracle("builtin")
    gpus = orc.extract_gpus(c["labels"], c["annotations"], table("builtin").keys(), "mig")
    assert [g["index"] for g in gpus] == c["gpu_indices"]  # duplicate-index quirk, node.go:89-96
    slot, out = o.mig_node_add_pod([(g["model"], g["used"], g["free"]) for g in gpus], c["req"])
    assert (slot < 0) == c["err"]
    used_sum, free_sum = {}, {}
    for _, u, f in out:
        for k, q in u.items():
            used_sum[k] = used_sum.get(k, 0) + q
        for k, q in f.items():
            free_sum[k] = free_sum.get(k, 0) + q
    assert used_sum == c["used_sum"] and free_sum == c["free_sum"]


@pytest.mark.parametrize("c", K["fewest_slices"], ids=lambda c: c["src"])
def test_fewest_slices(c):
    o = _oracle("builtin")
    assert o.fewest_slices(c["geometries"]) == c["expected_index"]


@pytest.mark.parametrize("c", K["memory_gb"], ids=lambda c: c["src"])
def test_memory_gb(c):
    v = orc.memory_gb(c["label"])
    assert (v is None) == c["err"]
    if not c["err"]:
        assert v == c["value"]


@pytest.mark.parametrize("c", K["parse_status_annotation"], ids=lambda c: c["src"])
def test_parse_status_annotation(c):
    a = orc.parse_status_annotation(c["key"], c["value"])
    assert (a is None) == c["err"]
    if not c["err"]:
        assert a == c["expected"]


@pytest.mark.parametrize("c", K["parse_spec_annotation"], ids=lambda c: c["src"])
def test_parse_spec_annotation(c):
    a = orc.parse_spec_annotation(c["key"], c["value"])
    assert (a is None) == c["err"]
    if not c["err"]:
        assert a == c["expected"]


@pytest.mark.parametrize("c", K["parse_node_annotations"], ids=lambda c: c["src"])
def test_parse_node_annotations(c):
    status, spec = orc.parse_node_annotations(c["annotations"])
    key = lambda a: sorted(a.items())
    assert sorted(status, key=key) == sorted(c["status"], key=key)  # ElementsMatch
    assert sorted(spec, key=key) == sorted(c["spec"], key=key)


@pytest.mark.parametrize("c", K["mps_new_gpu"], ids=lambda c: c["src"])
def test_mps_new_gpu(c):
    o = _oracle(sizes=_sizes(c["used"], c["free"]))
    assert o.mps_gpu_validate(c["mem"], c["used"], c["free"]) == c["err"]


@pytest.mark.parametrize("c", K["mps_gpu_update"], ids=lambda c: c["src"])
def test_mps_gpu_update(c):
    o = _oracle(sizes=_sizes(c["used"], c["free"], c["req"]))
    upd, free = o.mps_gpu_update(c["mem"], c["used"], c["free"], c["req"])
    assert upd == c["updated"]
    assert geometry_of(c["used"], free) == c["geometry"]


@pytest.mark.parametrize("c", K["mps_new_node"], ids=lambda c: c["src"])
def test_mps_new_node(c):
    gpus = orc.extract_gpus(c["labels"], c["annotations"], None, "mps")
    assert (gpus is None) == c["err"]
    if not c["err"]:
        got = [dict(index=g["index"], mem_gb=g["mem_gb"], used=g["used"], free=g["free"]) for g in gpus]
        assert got == c["gpus"]


@pytest.mark.parametrize("c", K["mps_node_geometry"], ids=lambda c: c["src"])
def test_mps_node_geometry(c):
    gpus = [(g["mem"], g["used"], g["free"]) for g in c["gpus"]]
    assert orc.Oracle.node_geometry(gpus) == c["geometry"]


@pytest.mark.parametrize("c", K["mps_node_has_free_capacity"], ids=lambda c: c["src"])
def test_mps_node_has_free_capacity(c):
    o = _oracle(sizes=_sizes(*[g["used"] for g in c["gpus"]], *[g["free"] for g in c["gpus"]]))
    gpus = [(g["mem"], g["used"], g["free"]) for g in c["gpus"]]
    assert o.mps_node_has_free_capacity(gpus) == c["expected"]


@pytest.mark.parametrize("c", K["mps_node_add_pod"], ids=lambda c: c["src"])
def test_mps_node_add_pod(c):
    gpus = orc.extract_gpus(c["labels"], c["annotations"], None, "mps")
    o = _oracle(sizes=_sizes(c["req"], *[g["used"] for g in gpus], *[g["free"] for g in gpus]))
    slot, out = o.mps_node_add_pod([(g["mem_gb"], g["used"], g["free"]) for g in gpus], c["req"])
    assert (slot < 0) == c["err"]
    used_sum, free_sum = {}, {}
    for _, u, f in out:
        for k, q in u.items():
            used_sum[k] = used_sum.get(k, 0) + q
        for k, q in f.items():
            free_sum[k] = free_sum.get(k, 0) + q
    assert used_sum == c["used_sum"] and free_sum == c["free_sum"]


@pytest.mark.parametrize("c", K["mps_node_update"], ids=lambda c: c["src"])
def test_mps_node_update(c):
    gpus = orc.extract_gpus(c["labels"], c["annotations"], None, "mps")
    assert gpus is not None
    o = _oracle(sizes=_sizes(c["req"], *[g["used"] for g in gpus], *[g["free"] for g in gpus]))
    upd, _ = o.mps_node_update([(g["mem_gb"], g["used"], g["free"]) for g in gpus], c["req"])
    assert upd == c["updated"]


MIG_NAME = re.compile(r"\d+g\.\d+gb")  # pkg/constant/constants.go:40 (unanchored MatchString)


@pytest.mark.parametrize("c", K["validate_configs"], ids=lambda c: c["src"])
def test_validate_configs(c):
    """known_configs.go:164-185: empty map, non-MIG/invalid names and quantity < 1 are rejected."""
    def validate(cfg):
        if not cfg:
            return False
        for geoms in cfg.values():
            for g in geoms:
                for name, q in g.items():
                    if not MIG_NAME.search(name) or q < 1:
                        return False
        return True
    ok = validate(c["configs"])
    assert ok != c["err"]
    if ok:
        _oracle_for = orc.Oracle(c["configs"])  # must load
    elif c["configs"] and all(q >= 0 for geoms in c["configs"].values() for g in geoms for q in g.values()):
        pass
    elif c["configs"]:
        with pytest.raises(ValueError):
            orc.Oracle(c["configs"])
    else:
        with pytest.raises(ValueError):
            orc.Oracle(c["configs"])


def test_builtin_and_yaml_tables_load_and_ids():
    """known_config_test.go:27-29 (defaults validate) + Geometry.Id format (partitioning.go:41-56)."""
    o = _oracle("builtin")
    assert o.geometry_id({"2g.20gb": 1, "1g.10gb": 5}) == "1g.10gb:5, 2g.20gb:1, "
    assert o.geometry_id({}) == ""
    y = orc.Oracle(table("yaml"))
    assert len(y.tables["NVIDIA-H100-80GB"]) == 12
    assert {"3g.20gb": 1, "2g.20gb": 2} in y.tables["NVIDIA-H100-80GB"]  # yaml typo kept (:46-47)
    assert {"2g.10gb": 1, "1g.10gb": 2, "3g.40gb": 1} in o.tables["NVIDIA-A100-80GB-PCIe"]  # builtin typo kept


def test_id_tiebreak_is_bytewise_on_rendered_string():
    """gpu.go:242 compares Geometry.Id() strings; a two-digit quantity sorts before '2' bytewise."""
    o = orc.Oracle({"M": [{"1g.5gb": 10}, {"1g.5gb": 2}]})
    assert o.geometry_id({"1g.5gb": 10}) < o.geometry_id({"1g.5gb": 2})
