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


# ---- dict-level fixtures -> flattened snapshots (what the Go shim would build) -------------------------
import numpy as np  # noqa: E402


def mask_of_dict(d, idx):
    m = 0
    for k in d:
        m |= 1 << idx[k]
    return m


def mig_snapshot_from_nodes(nodes, names, model_names, valid=None):
    """nodes: list of lists of (model, used{}, free{}).  Returns the snapshot dict of synth.py."""
    pid = {n: i for i, n in enumerate(names)}
    mid = {m: i for i, m in enumerate(model_names)}
    P = len(names)
    off = [0]
    model, used, free, um, fm = [], [], [], [], []
    for gpus in nodes:
        for (m, u, f) in gpus:
            model.append(mid.get(m, -1))
            ur = np.zeros(P, np.int32)
            fr = np.zeros(P, np.int32)
            for k, q in u.items():
                ur[pid[k]] = q
            for k, q in f.items():
                fr[pid[k]] = q
            used.append(ur)
            free.append(fr)
            um.append(mask_of_dict(u, pid))
            fm.append(mask_of_dict(f, pid))
        off.append(len(model))
    tg = len(model)
    return dict(n_nodes=len(nodes), gpu_off=np.asarray(off, np.int32),
                valid=np.asarray(valid if valid is not None else [1] * len(nodes), np.uint8).reshape(-1)[:len(nodes)]
                if len(nodes) else np.zeros(0, np.uint8),
                model=np.asarray(model, np.int32).reshape(tg), used=np.stack(used).astype(np.int32) if tg else np.zeros((0, P), np.int32),
                free=np.stack(free).astype(np.int32) if tg else np.zeros((0, P), np.int32),
                used_mask=np.asarray(um, np.uint32).reshape(tg), free_mask=np.asarray(fm, np.uint32).reshape(tg))


def mps_snapshot_from_nodes(nodes, sizes, size_of, valid=None):
    """nodes: list of lists of (mem_gb, used{}, free{}) keyed by "<k>gb" names."""
    sid = {s: i for i, s in enumerate(sizes)}
    S = len(sizes)
    off = [0]
    mem, used, free, um, fm = [], [], [], [], []
    for gpus in nodes:
        for (m, u, f) in gpus:
            mem.append(m)
            ur = np.zeros(S, np.int32)
            fr = np.zeros(S, np.int32)
            umk = fmk = 0
            for k, q in u.items():
                ur[sid[size_of(k)]] = q
                umk |= 1 << sid[size_of(k)]
            for k, q in f.items():
                fr[sid[size_of(k)]] = q
                fmk |= 1 << sid[size_of(k)]
            used.append(ur)
            free.append(fr)
            um.append(umk)
            fm.append(fmk)
        off.append(len(mem))
    tg = len(mem)
    return dict(n_nodes=len(nodes), gpu_off=np.asarray(off, np.int32),
                valid=np.asarray(valid if valid is not None else [1] * len(nodes), np.uint8).reshape(-1)[:len(nodes)]
                if len(nodes) else np.zeros(0, np.uint8),
                mem_gb=np.asarray(mem, np.int32).reshape(tg), used=np.stack(used).astype(np.int32) if tg else np.zeros((0, S), np.int32),
                free=np.stack(free).astype(np.int32) if tg else np.zeros((0, S), np.int32),
                used_mask=np.asarray(um, np.uint32).reshape(tg), free_mask=np.asarray(fm, np.uint32).reshape(tg))


def demand_csr(rows, idx_of):
    """rows: list of {name: qty} -> CSR with keys ascending."""
    off, key, qty = [0], [], []
    for r in rows:
        items = sorted((idx_of(k), q) for k, q in r.items())
        for k, q in items:
            key.append(k)
            qty.append(q)
        off.append(len(key))
    return (np.asarray(off, np.int32), np.asarray(key if key else [0], np.int32), np.asarray(qty if qty else [0], np.int32))
