"""Shared helpers for the known-answer tests (tests/golden/*.json)."""
import json
import os

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_kats():
    with open(os.path.join(GOLDEN, "reference_kats.json")) as f:
        return json.load(f)


def load_tables():
    with open(os.path.join(GOLDEN, "tables.json")) as f:
        return json.load(f)


def table(name):
    t = load_tables()
    if name == "builtin":
        return {k: v for k, v in t["builtin"].items()}
    if name == "gpu_test":  # pkg/gpu/mig/gpu_test.go:298-317
        d = {k: v for k, v in t["builtin"].items()}
        d["NVIDIA-A100-80GB-PCIe"] = t["gpu_test_80gb"]
        return d
    if name == "yaml":  # allowed_geometries.go:74-82 GroupByModel: last model wins
        d = {}
        for entry in t["yaml"]:
            for m in entry["models"]:
                d[m] = entry["allowedGeometries"]
        return d
    raise KeyError(name)


def geometry_of(used, free):
    """GetGeometry (mig/gpu.go:84-95; slicing/gpu.go:99-108): key union, values summed."""
    g = {}
    for k, q in list(used.items()) + list(free.items()):
        g[k] = g.get(k, 0) + q
    return g
