"""No GPU: noshost_tables_load — the native mirror of the reference's geometry-table loader
(AllowedMigGeometries.UnmarshalJSON + AllowedMigGeometriesList.GroupByModel, pkg/gpu/mig/allowed_geometries.go:30-82)
— against the reference's own vectors (pkg/gpu/mig/allowed_geometries_test.go:27-203) and its deployed file."""
import os

import numpy as np
import pytest

from kat_util import load_tables


@pytest.fixture(scope="module")
def host():
    import walkai_nos_b200 as pkg
    return pkg.host


# allowed_geometries_test.go:35-131 (TestAllowedMigGeometriesList__UnmarshalYaml)
UNMARSHAL = [
    ("empty :35", "", [], False),
    ("field 'models' is malformed :41", '\n- models: "A30"\n  allowedGeometries:\n    - 1g.6gb: 2\n      2g.12gb: 1\n    - 4g.24gb: 1\n', None, True),
    ("field 'allowedGeometries' is malformed :52", '\n- models: [ "A30" ]\n  allowedGeometries: foo\n', None, True),
    ("missing field 'models' :60", "\n- allowedGeometries:\n    - 1g.6gb: 2\n      2g.12gb: 1\n    - 4g.24gb: 1\n", None, True),
    ("missing field 'allowedGeometries' :70", '\n- models: [ "A30" ]\n', None, True),
    ("not empty :77", '\n- models: [ "A30" ]\n  allowedGeometries:\n    - 1g.6gb: 2\n      2g.12gb: 1\n    - 4g.24gb: 1\n'
     '- models: [ "A100-SXM4-40GB", "NVIDIA-A100-40GB-PCIe" ]\n  allowedGeometries:\n    - 1g.5gb: 7\n    - 1g.5gb: 5\n      2g.10gb: 1\n',
     [{"models": ["A30"], "allowedGeometries": [{"1g.6gb": 2, "2g.12gb": 1}, {"4g.24gb": 1}]},
      {"models": ["A100-SXM4-40GB", "NVIDIA-A100-40GB-PCIe"], "allowedGeometries": [{"1g.5gb": 7}, {"1g.5gb": 5, "2g.10gb": 1}]}], False),
]


@pytest.mark.parametrize("name,text,expected,error", UNMARSHAL, ids=[u[0] for u in UNMARSHAL])
def test_unmarshal_yaml_vectors(host, name, text, expected, error):
    if error:
        with pytest.raises(ValueError):
            host.load_tables(text)
        return
    entries, grouped = host.load_tables(text)
    assert entries == expected


def test_error_messages_name_the_missing_field(host):
    with pytest.raises(ValueError, match="missing field 'models'"):  # allowed_geometries.go:39-41
        host.load_tables("- allowedGeometries:\n    - 4g.24gb: 1\n")
    with pytest.raises(ValueError, match="missing field 'allowedGeometries'"):  # :48-50
        host.load_tables('- models: [ "A30" ]\n')


def test_group_by_model_vectors(host):
    # allowed_geometries_test.go:139-143: empty list -> empty map
    entries, g = host.load_tables("")
    assert g["tables"] == {} and g["models"] == [] and list(g["geom_off"]) == [0]
    # :145-196: every model of an entry gets the entry's list
    text = ("- models: [ m1, m2 ]\n  allowedGeometries:\n    - p1: 1\n      p2: 2\n    - p1: 3\n"
            "- models: [ m3 ]\n  allowedGeometries:\n    - p3: 2\n      p2: 2\n")
    entries, g = host.load_tables(text)
    assert g["tables"] == {"m1": [{"p1": 1, "p2": 2}, {"p1": 3}], "m2": [{"p1": 1, "p2": 2}, {"p1": 3}], "m3": [{"p3": 2, "p2": 2}]}
    assert g["profiles"] == ["p1", "p2", "p3"] and list(g["geom_off"]) == [0, 2, 4, 5]
    assert np.array_equal(g["geom_counts"], [[1, 2, 0], [3, 0, 0], [1, 2, 0], [3, 0, 0], [0, 2, 2]])
    # a model named twice keeps the LAST entry's list (res[model] = ag.Geometries overwrites, :78-80)
    _, g = host.load_tables("- models: [ a ]\n  allowedGeometries:\n    - p1: 1\n- models: [ b, a ]\n  allowedGeometries:\n    - p2: 2\n")
    assert g["models"] == ["a", "b"] and g["tables"] == {"a": [{"p2": 2}], "b": [{"p2": 2}]}


def test_yaml_forms_the_reference_file_could_use(host):
    block = ("# comment\n- models:\n    - \"A30\"\n    - 'X'   # trailing comment\n  allowedGeometries:\n    - 1g.6gb: 4\n\n    - 2g.12gb: 2\n")
    entries, _ = host.load_tables(block)
    assert entries == [{"models": ["A30", "X"], "allowedGeometries": [{"1g.6gb": 4}, {"2g.12gb": 2}]}]
    with pytest.raises(ValueError):
        host.load_tables('- models: [ "A30" ]\n  allowedGeometries:\n    - 1g.6gb: "4"\n')  # a string is not an int
    with pytest.raises(ValueError):
        host.load_tables('- models: [ "A30" ]\n  allowedGeometries:\n    - 1g.6gb: four\n')
    # quantities are not judged by the loader: ValidateConfigs (known_configs.go:164-185) does that later
    entries, _ = host.load_tables('- models: [ "A30" ]\n  allowedGeometries:\n    - 1g.6gb: 0\n')
    assert entries[0]["allowedGeometries"] == [{"1g.6gb": 0}]


def test_shipped_table_is_the_transcribed_reference_table(host):
    import walkai_nos_b200 as pkg
    entries, g = host.load_tables(pkg.synth.table_yaml_text())
    gold = load_tables()["yaml"]  # tests/golden/tables.json, transcribed from the reference's YAML with its line numbers
    assert entries == [{"models": e["models"], "allowedGeometries": e["allowedGeometries"]} for e in gold]
    assert pkg.synth.yaml_tables() == g["tables"]
    ref = "/root/reference/config/gpupartitioner/manager/known_mig_geometries.yaml"
    if os.path.exists(ref):  # in the build container: the reference's own file parses to the same list
        with open(ref) as f:
            assert host.load_tables(f.read())[0] == entries
