"""Synthetic cluster snapshots for the five BASELINE.json configs.

Deterministic: splitmix64 seeded 0x6e6f730000 + config number, consumed in the order documented in
each function.  Used by the parity tests and by bench.py (the same arrays go to the CUDA library and
to the CPU oracle).  Geometry tables are the reference's deployed YAML table
(config/gpupartitioner/manager/known_mig_geometries.yaml:1-59), kept as data/mig_geometries.json, rendered into the
reference's wire format and read through the native loader of libnoshost.
"""
from __future__ import annotations

import os
from typing import Dict, List, Tuple

import numpy as np

MASK64 = (1 << 64) - 1

# pkg/gpu/mig/profile.go:29-47 — the 13 named profiles, in canonical (bytewise ascending) order
MIG_PROFILES = sorted(["1g.6gb", "2g.12gb", "4g.24gb", "1g.5gb", "2g.10gb", "3g.20gb", "4g.20gb", "7g.40gb",
                       "1g.10gb", "2g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"])

_TABLE_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "mig_geometries.json")
_tables_cache = None


def table_yaml_text() -> str:
    """The deployed geometry table (data/mig_geometries.json: the entries of the reference's
    config/gpupartitioner/manager/known_mig_geometries.yaml:1-59) in the wire format the reference reads: a block sequence
    of {models: [flow list], allowedGeometries: [block mappings]}."""
    import json
    with open(_TABLE_FILE) as f:
        entries = json.load(f)["entries"]
    lines = []
    for e in entries:
        lines.append("- models: [ " + ", ".join('"%s"' % m for m in e["models"]) + " ]")
        lines.append("  allowedGeometries:")
        for geom in e["allowedGeometries"]:
            for j, (name, qty) in enumerate(geom.items()):
                lines.append(("    - " if j == 0 else "      ") + f"{name}: {qty}")
    return "\n".join(lines) + "\n"


def yaml_tables() -> Dict[str, List[Dict[str, int]]]:
    """model -> allowed geometries of the deployed table, read through the native loader (noshost_tables_load:
    AllowedMigGeometries.UnmarshalJSON + GroupByModel, pkg/gpu/mig/allowed_geometries.go:30-82)."""
    global _tables_cache
    if _tables_cache is None:
        from . import host
        _tables_cache = host.load_tables(table_yaml_text())[1]["tables"]
    return {m: [dict(g) for g in geoms] for m, geoms in _tables_cache.items()}


def flatten_tables(tables: Dict[str, List[Dict[str, int]]], names: List[str]):
    pid = {n: i for i, n in enumerate(names)}
    off = [0]
    rows = []
    for m in tables:
        for g in tables[m]:
            r = np.zeros(len(names), np.int32)
            for k, v in g.items():
                r[pid[k]] = v
            rows.append(r)
        off.append(len(rows))
    return list(tables.keys()), np.asarray(off, np.int32), np.stack(rows).astype(np.int32)


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & MASK64

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & MASK64
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK64
        return z ^ (z >> 31)

    def uniform(self) -> float:
        return (self.next() >> 11) / float(1 << 53)

    def below(self, n: int) -> int:
        return self.next() % n

    def choice_w(self, weights) -> int:
        u = self.uniform()
        acc = 0.0
        for i, w in enumerate(weights):
            acc += w
            if u < acc:
                return i
        return len(weights) - 1


def seed_for(config: int) -> int:
    return 0x6E6F730000 + config


def gen_mig_nodes(rng: SplitMix64, n_nodes: int, gpus_per_node: int, model: str, tables, names):
    """Per GPU, in order: one draw picks an allowed geometry uniformly; then, for each profile of
    that geometry in ascending name order and each of its devices, one draw marks it used (u < 0.5)."""
    P = len(names)
    pid = {n: i for i, n in enumerate(names)}
    model_names = list(tables.keys())
    mid = model_names.index(model)
    geoms = tables[model]
    tg = n_nodes * gpus_per_node
    used = np.zeros((tg, P), np.int32)
    free = np.zeros((tg, P), np.int32)
    for g in range(tg):
        geom = geoms[rng.below(len(geoms))]
        for name in sorted(geom):
            for _ in range(geom[name]):
                if rng.uniform() < 0.5:
                    used[g, pid[name]] += 1
                else:
                    free[g, pid[name]] += 1
    return dict(n_nodes=n_nodes, gpu_off=np.arange(0, tg + 1, gpus_per_node, dtype=np.int32),
                valid=np.ones(n_nodes, np.uint8), model=np.full(tg, mid, np.int32), used=used, free=free,
                used_mask=_mask_of(used), free_mask=_mask_of(free))


def gen_mps_nodes(rng: SplitMix64, n_nodes: int, gpus_per_node: int, mem_gb: int, sizes: List[int]):
    """Per GPU: one draw = fill fraction; then repeatedly one draw picks a size uniformly from
    `sizes` and — if it still fits under fill*mem — one draw marks it used (u < 0.5) else free;
    the first slice that does not fit ends the GPU."""
    S = len(sizes)
    tg = n_nodes * gpus_per_node
    used = np.zeros((tg, S), np.int32)
    free = np.zeros((tg, S), np.int32)
    for g in range(tg):
        target = rng.uniform() * mem_gb
        tot = 0
        while True:
            si = rng.below(S)
            if tot + sizes[si] > target:
                break
            if rng.uniform() < 0.5:
                used[g, si] += 1
            else:
                free[g, si] += 1
            tot += sizes[si]
    return dict(n_nodes=n_nodes, gpu_off=np.arange(0, tg + 1, gpus_per_node, dtype=np.int32),
                valid=np.ones(n_nodes, np.uint8), mem_gb=np.full(tg, mem_gb, np.int32), used=used, free=free,
                used_mask=_mask_of(used), free_mask=_mask_of(free))


def _mask_of(a: np.ndarray) -> np.ndarray:
    m = np.zeros(a.shape[0], np.uint32)
    for k in range(a.shape[1]):
        m |= (a[:, k] != 0).astype(np.uint32) << np.uint32(k)
    return m


def gen_pods(rng: SplitMix64, n_pods: int, mig_choices, mig_weights, mps_choices, mig_frac: float, qty2_prob: float):
    """Per pod, in order: [kind draw if both kinds] profile draw, quantity draw (if qty2_prob > 0),
    priority draw (uniform over {0,100,1000}).  Creation time = pod index (strictly increasing).
    Returns pods sorted by the planner order: priority desc, creation asc, name asc."""
    pods = []
    for i in range(n_pods):
        if mig_choices and mps_choices:
            kind = 0 if rng.uniform() < mig_frac else 1
        else:
            kind = 0 if mig_choices else 1
        if kind == 0:
            key = mig_choices[rng.choice_w(mig_weights)]
        else:
            key = mps_choices[rng.below(len(mps_choices))]
        qty = 1
        if qty2_prob > 0 and kind == 0:
            qty = 2 if rng.uniform() < qty2_prob else 1
        prio = (0, 100, 1000)[rng.below(3)]
        pods.append((kind, key, qty, prio, i, f"ns/pod-{i:06d}"))
    pods.sort(key=lambda p: (-p[3], p[4], p[5]))
    kind = np.asarray([p[0] for p in pods], np.uint8)
    key = np.asarray([p[1] for p in pods], np.int32)
    qty = np.asarray([p[2] for p in pods], np.int32)
    off = np.arange(0, n_pods + 1, dtype=np.int32)  # one key per pod
    return dict(kind=kind, off=off, key=key, qty=qty, priority=np.asarray([p[3] for p in pods], np.int32),
                created=np.asarray([p[4] for p in pods], np.int64))


MPS_SIZES = [5, 10, 20, 40]
MPS_MEM_GB = 82  # label 81920 MB -> ceil(81920/1000) (pkg/gpu/util.go:59-73)


def make_config(cfg: int, scale_nodes: int = 0, scale_pods: int = 0):
    """Returns dict(tables, names, model_names, geom_off, geom_counts, mig, mps, sizes, pods)."""
    names = MIG_PROFILES
    tables = yaml_tables()
    model_names, geom_off, geom_counts = flatten_tables(tables, names)
    pid = {n: i for i, n in enumerate(names)}
    rng = SplitMix64(seed_for(cfg))
    h100 = ["1g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]
    out = dict(cfg=cfg, tables=tables, names=names, model_names=model_names, geom_off=geom_off, geom_counts=geom_counts,
               mig=None, mps=None, sizes=MPS_SIZES)
    if cfg == 1:
        n, m = scale_nodes or 1, scale_pods or 20
        out["mig"] = gen_mig_nodes(rng, n, 8, "NVIDIA-A100-80GB-PCIe", tables, names)
        out["pods"] = gen_pods(rng, m, [pid["1g.10gb"], pid["2g.20gb"], pid["3g.40gb"]], [0.5, 0.3, 0.2], None, 1.0, 0.0)
    elif cfg == 2:
        n, m = scale_nodes or 64, scale_pods or 2000
        out["mig"] = gen_mig_nodes(rng, n, 8, "NVIDIA-H100-80GB", tables, names)
        out["pods"] = gen_pods(rng, m, [pid[x] for x in h100], [1 / 6.0] * 6, None, 1.0, 0.1)
    elif cfg == 3:
        n, m = scale_nodes or 256, scale_pods or 10000
        out["mps"] = gen_mps_nodes(rng, n, 8, MPS_MEM_GB, MPS_SIZES)
        out["pods"] = gen_pods(rng, m, None, None, list(range(len(MPS_SIZES))), 0.0, 0.0)
    elif cfg in (4, 5):
        n, m = (1024, 20000) if cfg == 4 else (4096, 50000)
        n, m = scale_nodes or n, scale_pods or m
        # even node index = MIG (H100 table), odd = MPS: generated as two node lists, MIG first
        out["mig"] = gen_mig_nodes(rng, (n + 1) // 2, 8, "NVIDIA-H100-80GB", tables, names)
        out["mps"] = gen_mps_nodes(rng, n // 2, 8, MPS_MEM_GB, MPS_SIZES)
        out["pods"] = gen_pods(rng, m, [pid[x] for x in h100], [1 / 6.0] * 6, list(range(len(MPS_SIZES))), 0.5, 0.1)
    else:
        raise ValueError(f"unknown config {cfg}")
    return out


def pods_of_kind(pods, kind: int):
    """CSR of the pods of one kind, order kept."""
    sel = np.nonzero(pods["kind"] == kind)[0]
    off = [0]
    key, qty = [], []
    for i in sel:
        for e in range(pods["off"][i], pods["off"][i + 1]):
            key.append(int(pods["key"][e]))
            qty.append(int(pods["qty"][e]))
        off.append(len(key))
    return (np.asarray(off, np.int32), np.asarray(key if key else [0], np.int32), np.asarray(qty if qty else [0], np.int32),
            sel)


def csr_to_dense(off, key, qty, width: int):
    D = len(off) - 1
    req = np.zeros((D, max(width, 1)), np.int32)
    mask = np.zeros(D, np.uint32)
    for d in range(D):
        for e in range(off[d], off[d + 1]):
            req[d, key[e]] = qty[e]
            mask[d] |= np.uint32(1 << int(key[e]))
    return req, mask
