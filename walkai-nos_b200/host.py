"""ctypes binding of libnoshost.so (include/noshost.h): the host side of the drop-in — annotation wire
format <-> SoA snapshot, pod filtering / ordering / request extraction, plan -> spec annotations."""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.environ.get("NOSHOST_LIB") or os.path.join(_HERE, "libnoshost.so")  # NOSHOST_LIB: the ASan/UBSan build
_lib = None

LABEL_PRODUCT = "nvidia.com/gpu.product"  # pkg/constant/constants.go:67-77
LABEL_COUNT = "nvidia.com/gpu.count"
LABEL_MEMORY = "nvidia.com/gpu.memory"
STATUS = {0: "used", 1: "free", 2: "unknown"}

EXPORTS = [
    "noshost_parse_status_annotation", "noshost_parse_spec_annotation", "noshost_format_status_annotation",
    "noshost_format_spec_annotation", "noshost_memory_gb", "noshost_mig_profile_of_resource",
    "noshost_mps_profile_of_resource", "noshost_mps_profile_size_gb", "noshost_snapshot_build", "noshost_snapshot_free",
    "noshost_snapshot_n_nodes", "noshost_snapshot_total_gpus", "noshost_snapshot_n_keys", "noshost_snapshot_key_name",
    "noshost_snapshot_key_size", "noshost_snapshot_gpu_off", "noshost_snapshot_valid", "noshost_snapshot_model",
    "noshost_snapshot_mem_gb", "noshost_snapshot_gpu_index", "noshost_snapshot_used", "noshost_snapshot_free_",
    "noshost_snapshot_used_mask", "noshost_snapshot_free_mask", "noshost_pods_build", "noshost_pods_free",
    "noshost_pods_count", "noshost_pods_source", "noshost_pods_kind", "noshost_pods_off", "noshost_pods_key",
    "noshost_pods_qty", "noshost_pods_error", "noshost_spec_annotations", "noshost_node_partitioning_equal",
    "noshost_tables_load", "noshost_tables_free", "noshost_tables_n_entries", "noshost_tables_entry_n_models",
    "noshost_tables_entry_model", "noshost_tables_entry_n_geometries", "noshost_tables_entry_geometry_size",
    "noshost_tables_entry_geometry_profile", "noshost_tables_entry_geometry_quantity", "noshost_tables_n_models",
    "noshost_tables_model", "noshost_tables_n_profiles", "noshost_tables_profile", "noshost_tables_geom_off",
    "noshost_tables_geom_counts",
]


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_LIB)
        L.noshost_snapshot_build.restype = C.c_void_p
        L.noshost_pods_build.restype = C.c_void_p
        L.noshost_snapshot_key_name.restype = C.c_char_p
        L.noshost_pods_error.restype = C.c_char_p
        for n in ("gpu_off", "model", "mem_gb", "gpu_index", "used", "free_"):
            getattr(L, f"noshost_snapshot_{n}").restype = C.POINTER(C.c_int32)
        L.noshost_snapshot_valid.restype = C.POINTER(C.c_uint8)
        L.noshost_snapshot_used_mask.restype = C.POINTER(C.c_uint32)
        L.noshost_snapshot_free_mask.restype = C.POINTER(C.c_uint32)
        for n in ("source", "off", "key", "qty"):
            getattr(L, f"noshost_pods_{n}").restype = C.POINTER(C.c_int32)
        L.noshost_pods_kind.restype = C.POINTER(C.c_uint8)
        L.noshost_tables_load.restype = C.c_void_p
        L.noshost_tables_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.noshost_tables_free.argtypes = [C.c_void_p]
        for n, rt, extra in (("n_entries", C.c_int, 0), ("entry_n_models", C.c_int, 1), ("entry_model", C.c_char_p, 2),
                             ("entry_n_geometries", C.c_int, 1), ("entry_geometry_size", C.c_int, 2),
                             ("entry_geometry_profile", C.c_char_p, 3), ("entry_geometry_quantity", C.c_int64, 3),
                             ("n_models", C.c_int, 0), ("model", C.c_char_p, 1), ("n_profiles", C.c_int, 0), ("profile", C.c_char_p, 1),
                             ("geom_off", C.POINTER(C.c_int32), 0), ("geom_counts", C.POINTER(C.c_int32), 0)):
            f = getattr(L, f"noshost_tables_{n}")
            f.restype = rt
            f.argtypes = [C.c_void_p] + [C.c_int] * extra
        for name in EXPORTS:
            f = getattr(L, name)
            if name.startswith("noshost_tables_"):
                continue
            if name.startswith(("noshost_snapshot_", "noshost_pods_")) and not name.endswith("_build"):
                f.argtypes = [C.c_void_p] + ([C.c_int] if name.endswith(("key_name", "key_size")) else [])
        _lib = L
    return _lib


def _strs(xs: Sequence[Optional[str]]):
    return (C.c_char_p * max(len(xs), 1))(*[None if x is None else x.encode() for x in xs])


def parse_status_annotation(key: str, value: str):
    idx, st, q = C.c_int32(), C.c_int32(), C.c_int32()
    buf = C.create_string_buffer(256)
    if lib().noshost_parse_status_annotation(key.encode(), value.encode(), C.byref(idx), buf, 256, C.byref(st), C.byref(q)):
        return None
    return dict(index=idx.value, profile=buf.value.decode(), status=STATUS[st.value], quantity=q.value)


def parse_spec_annotation(key: str, value: str):
    idx, q = C.c_int32(), C.c_int32()
    buf = C.create_string_buffer(256)
    if lib().noshost_parse_spec_annotation(key.encode(), value.encode(), C.byref(idx), buf, 256, C.byref(q)):
        return None
    return dict(index=idx.value, profile=buf.value.decode(), quantity=q.value)


def format_status_annotation(index: int, profile: str, status: str) -> str:
    buf = C.create_string_buffer(512)
    lib().noshost_format_status_annotation(index, profile.encode(), {"used": 0, "free": 1, "unknown": 2}[status], buf, 512)
    return buf.value.decode()


def format_spec_annotation(index: int, profile: str) -> str:
    buf = C.create_string_buffer(512)
    lib().noshost_format_spec_annotation(index, profile.encode(), buf, 512)
    return buf.value.decode()


def memory_gb(label: str):
    out = C.c_int32()
    return None if lib().noshost_memory_gb(label.encode(), C.byref(out)) else out.value


def mig_profile_of_resource(r: str):
    buf = C.create_string_buffer(256)
    return None if lib().noshost_mig_profile_of_resource(r.encode(), buf, 256) else buf.value.decode()


def mps_profile_of_resource(r: str):
    buf = C.create_string_buffer(256)
    return None if lib().noshost_mps_profile_of_resource(r.encode(), buf, 256) else buf.value.decode()


def mps_profile_size_gb(p: str) -> int:
    return int(lib().noshost_mps_profile_size_gb(p.encode()))


class Snapshot:
    """Flattened []mig.Node / []slicing.Node built from node labels + annotations."""

    def __init__(self, kind: str, nodes: List[dict], model_names: Iterable[str] = (), extra_names: Iterable[str] = (),
                 extra_sizes: Iterable[int] = ()):
        L = lib()
        self.kind = kind
        self.node_names = [n.get("name", f"node-{i}") for i, n in enumerate(nodes)]
        product = [n.get("labels", {}).get(LABEL_PRODUCT) for n in nodes]
        count = [n.get("labels", {}).get(LABEL_COUNT) for n in nodes]
        memory = [n.get("labels", {}).get(LABEL_MEMORY) for n in nodes]
        off, keys, vals = [0], [], []
        for n in nodes:
            for k, v in n.get("annotations", {}).items():
                keys.append(k)
                vals.append(v)
            off.append(len(keys))
        off = np.asarray(off, np.int32)
        models = list(model_names)
        en = list(extra_names)
        es = np.asarray(list(extra_sizes) or [0], np.int32)
        self.h = C.c_void_p(L.noshost_snapshot_build(
            0 if kind == "mig" else 1, len(nodes), _strs(product), _strs(count), _strs(memory),
            off.ctypes.data_as(C.POINTER(C.c_int32)), _strs(keys), _strs(vals), len(models), _strs(models), len(en), _strs(en),
            len(list(extra_sizes)), es.ctypes.data_as(C.POINTER(C.c_int32))))
        if not self.h.value:
            raise ValueError("noshost_snapshot_build refused its arguments")
        self.n_nodes = L.noshost_snapshot_n_nodes(self.h)
        self.tg = L.noshost_snapshot_total_gpus(self.h)
        self.K = L.noshost_snapshot_n_keys(self.h)
        self.key_names = [L.noshost_snapshot_key_name(self.h, k).decode() for k in range(self.K)]
        self.key_sizes = [L.noshost_snapshot_key_size(self.h, k) for k in range(self.K)]

    def _arr(self, name, n, dtype):
        p = getattr(lib(), f"noshost_snapshot_{name}")(self.h)
        return np.ctypeslib.as_array(p, shape=(max(n, 1),))[:n].astype(dtype).copy() if n else np.zeros(0, dtype)

    def arrays(self) -> Dict[str, np.ndarray]:
        d = dict(n_nodes=self.n_nodes, gpu_off=self._arr("gpu_off", self.n_nodes + 1, np.int32),
                 valid=self._arr("valid", self.n_nodes, np.uint8), gpu_index=self._arr("gpu_index", self.tg, np.int32),
                 used=self._arr("used", self.tg * self.K, np.int32).reshape(self.tg, self.K),
                 free=self._arr("free_", self.tg * self.K, np.int32).reshape(self.tg, self.K),
                 used_mask=self._arr("used_mask", self.tg, np.uint32), free_mask=self._arr("free_mask", self.tg, np.uint32))
        if self.kind == "mig":
            d["model"] = self._arr("model", self.tg, np.int32)
        else:
            d["mem_gb"] = self._arr("mem_gb", self.tg, np.int32)
        return d

    def spec_annotations(self, node: int, used: np.ndarray, free: np.ndarray) -> Dict[str, str]:
        used = np.ascontiguousarray(used, np.int32)
        free = np.ascontiguousarray(free, np.int32)
        n = lib().noshost_spec_annotations(self.h, node, used.ctypes.data_as(C.POINTER(C.c_int32)),
                                           free.ctypes.data_as(C.POINTER(C.c_int32)), None, 0)
        buf = C.create_string_buffer(n + 1)
        lib().noshost_spec_annotations(self.h, node, used.ctypes.data_as(C.POINTER(C.c_int32)),
                                       free.ctypes.data_as(C.POINTER(C.c_int32)), buf, n + 1)
        return dict(line.split("=", 1) for line in buf.value.decode().splitlines())

    def close(self):
        if getattr(self, "h", None):
            lib().noshost_snapshot_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def build_pods(pods: List[dict], mig: Optional[Snapshot], mps: Optional[Snapshot]):
    """pods: dicts with phase ("Pending"...), node_name, unschedulable (bool), priority, created, name,
    containers / init_containers = lists of {resource: qty}.  Returns planner-ordered CSR rows."""
    L = lib()
    M = len(pods)
    pending = np.asarray([p.get("phase", "Pending") == "Pending" for p in pods] or [0], np.uint8)
    scheduled = np.asarray([bool(p.get("node_name", "")) for p in pods] or [0], np.uint8)
    unsched = np.asarray([bool(p.get("unschedulable", True)) for p in pods] or [0], np.uint8)
    prio = np.asarray([p.get("priority", 0) for p in pods] or [0], np.int32)
    created = np.asarray([p.get("created", 0) for p in pods] or [0], np.int64)
    names = [p.get("name", f"default/pod-{i}") for i, p in enumerate(pods)]
    cont_off, cont_init, res_off, res_name, res_qty = [0], [], [0], [], []
    for p in pods:
        for init, lst in ((0, p.get("containers", [])), (1, p.get("init_containers", []))):
            for c in lst:
                cont_init.append(init)
                for r, q in c.items():
                    res_name.append(r)
                    res_qty.append(q)
                res_off.append(len(res_name))
        cont_off.append(len(cont_init))
    i32 = lambda a: np.asarray(a or [0], np.int32)
    cont_off_a, res_off_a = np.asarray(cont_off, np.int32), np.asarray(res_off, np.int32)
    cont_init_a = np.asarray(cont_init or [0], np.uint8)
    res_qty_a = np.asarray(res_qty or [0], np.int64)
    P = lambda a, t: a.ctypes.data_as(C.POINTER(t))
    h = C.c_void_p(L.noshost_pods_build(M, P(pending, C.c_uint8), P(scheduled, C.c_uint8), P(unsched, C.c_uint8),
                                        P(prio, C.c_int32), P(created, C.c_int64), _strs(names), P(cont_off_a, C.c_int32),
                                        P(cont_init_a, C.c_uint8), P(res_off_a, C.c_int32), _strs(res_name),
                                        P(res_qty_a, C.c_int64), mig.h if mig else None, mps.h if mps else None))
    try:
        err = L.noshost_pods_error(h).decode()
        if err:
            raise ValueError(err)
        n = L.noshost_pods_count(h)
        def get(name, cnt, dt):
            ptr = getattr(L, f"noshost_pods_{name}")(h)
            if cnt == 0 or not ptr:  # no planner rows at all (every pod filtered out): empty std::vector, null data()
                return np.zeros(max(cnt, 0), dt)
            return np.ctypeslib.as_array(ptr, shape=(cnt,)).astype(dt).copy()
        off = get("off", n + 1, np.int32)
        if off.size == 0:
            off = np.zeros(1, np.int32)
        nnz = int(off[-1]) if n else 0
        key, qty = get("key", nnz, np.int32), get("qty", nnz, np.int32)
        if nnz == 0:
            key, qty = np.zeros(1, np.int32), np.zeros(1, np.int32)
        return dict(source=get("source", n, np.int32), kind=get("kind", n, np.uint8), off=off, key=key, qty=qty)
    finally:
        L.noshost_pods_free(h)


def node_partitioning_equal(a: List[dict], b: List[dict], resources: Sequence[str]) -> bool:
    """a, b: lists of {"index": int, "resources": {name: qty}} (state.NodePartitioning.GPUs)."""
    K = max(len(resources), 1)
    col = {r: i for i, r in enumerate(resources)}

    def flat(gs):
        idx = np.asarray([g["index"] for g in gs] or [0], np.int32)
        res = np.zeros((max(len(gs), 1), K), np.int32)
        for i, g in enumerate(gs):
            for r, q in g["resources"].items():
                res[i, col[r]] = q
        return idx, res
    ia, ra = flat(a)
    ib, rb = flat(b)
    P = lambda x: x.ctypes.data_as(C.POINTER(C.c_int32))
    return bool(lib().noshost_node_partitioning_equal(K, len(a), P(ia), P(ra), len(b), P(ib), P(rb)))


def partitioning_state_equal(a: Dict[str, List[dict]], b: Dict[str, List[dict]]) -> bool:
    """state.PartitioningState.Equal (internal/partitioning/state/partitioning.go:46-56)."""
    if len(a) != len(b):
        return False
    resources = sorted({r for st in (a, b) for gs in st.values() for g in gs for r in g["resources"]})
    return all(node_partitioning_equal(gs, b.get(node, []), resources) for node, gs in a.items())


def load_tables(text: str):
    """known_mig_geometries.yaml -> (entries, grouped): entries = the AllowedMigGeometriesList as parsed
    ([{"models": [...], "allowedGeometries": [{profile: qty}, ...]}, ...]); grouped = dict(models, profiles, geom_off,
    geom_counts, tables) = GroupByModel flattened for nosplan_set_mig_tables (tables: {model: [geometry dicts]}).
    Raises ValueError where the reference's yaml.Unmarshal returns an error (allowed_geometries.go:30-59)."""
    L = lib()
    err = C.create_string_buffer(256)
    h = L.noshost_tables_load(text.encode(), err, 256)
    if not h:
        raise ValueError(err.value.decode())
    h = C.c_void_p(h)
    try:
        entries = []
        for e in range(L.noshost_tables_n_entries(h)):
            models = [L.noshost_tables_entry_model(h, e, i).decode() for i in range(L.noshost_tables_entry_n_models(h, e))]
            geoms = []
            for g in range(L.noshost_tables_entry_n_geometries(h, e)):
                geoms.append({L.noshost_tables_entry_geometry_profile(h, e, g, k).decode(): int(L.noshost_tables_entry_geometry_quantity(h, e, g, k))
                              for k in range(L.noshost_tables_entry_geometry_size(h, e, g))})
            entries.append({"models": models, "allowedGeometries": geoms})
        models = [L.noshost_tables_model(h, m).decode() for m in range(L.noshost_tables_n_models(h))]
        profiles = [L.noshost_tables_profile(h, p).decode() for p in range(L.noshost_tables_n_profiles(h))]
        P = len(profiles)
        off = np.ctypeslib.as_array(L.noshost_tables_geom_off(h), shape=(len(models) + 1,)).astype(np.int32).copy()
        total = int(off[-1])
        counts = (np.ctypeslib.as_array(L.noshost_tables_geom_counts(h), shape=(total * P,)).astype(np.int32).reshape(total, P).copy()
                  if total * P else np.zeros((total, P), np.int32))
        tables = {m: [{profiles[p]: int(counts[g, p]) for p in range(P) if counts[g, p] != 0} for g in range(off[i], off[i + 1])]
                  for i, m in enumerate(models)}
        return entries, dict(models=models, profiles=profiles, geom_off=off, geom_counts=counts, tables=tables)
    finally:
        L.noshost_tables_free(h)
