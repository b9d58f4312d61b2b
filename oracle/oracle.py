"""ctypes front-end of the CPU oracle (oracle/nos_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package never imports this module.

Dict-based helpers mirror the reference's Go maps: a Python dict {profile: qty} is a Go
map[ProfileName]int, including zero-valued keys (a key present with value 0 is not the same as an
absent key, see SURVEY §7 "Map semantics").
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libnos_oracle.so")

i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
i8p = C.POINTER(C.c_int8)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "nos_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libnos_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


class MigSnap(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("gpu_off", i32p), ("valid", u8p), ("model", i32p), ("used", i32p),
                ("used_mask", u32p), ("free", i32p), ("free_mask", u32p)]


class MpsSnap(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("gpu_off", i32p), ("valid", u8p), ("mem_gb", i32p), ("used", i32p),
                ("used_mask", u32p), ("free", i32p), ("free_mask", u32p)]


def _p(a: Optional[np.ndarray], t):
    return None if a is None else a.ctypes.data_as(t)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_new.restype = C.c_void_p
        _lib.orc_error.restype = C.c_char_p
        for name in dir(_lib):
            pass
    return _lib


# Builtin MIG profile constants, pkg/gpu/mig/profile.go:29-47
BUILTIN_MIG_PROFILES = ["1g.6gb", "2g.12gb", "4g.24gb", "1g.5gb", "2g.10gb", "3g.20gb", "4g.20gb", "7g.40gb",
                        "1g.10gb", "2g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]


class Oracle:
    """One oracle context: MIG vocabulary + model tables, MPS size vocabulary."""

    def __init__(self, tables: Dict[str, List[Dict[str, int]]], extra_mig_profiles: Iterable[str] = (),
                 mps_sizes: Iterable[int] = (), builtin_profiles: bool = True):
        L = lib()
        self.L = L
        self.h = C.c_void_p(L.orc_new())
        names = (set(BUILTIN_MIG_PROFILES) if builtin_profiles else set()) | set(extra_mig_profiles)
        for geoms in tables.values():
            for g in geoms:
                names |= set(g.keys())
        self.mig_names = sorted(names)  # canonical order: ascending bytewise (ASCII)
        self.P = len(self.mig_names)
        self.pid = {n: i for i, n in enumerate(self.mig_names)}
        arr = (C.c_char_p * self.P)(*[n.encode() for n in self.mig_names])
        if L.orc_set_mig_vocab(self.h, self.P, arr) != 0:
            raise ValueError(L.orc_error(self.h).decode())
        self.set_tables(tables)
        self.set_mps_sizes(mps_sizes)

    def __del__(self):
        try:
            self.L.orc_free(self.h)
        except Exception:
            pass

    # ---- tables ------------------------------------------------------------------------------
    def set_tables(self, tables: Dict[str, List[Dict[str, int]]]):
        self.model_names = list(tables.keys())
        self.mid = {m: i for i, m in enumerate(self.model_names)}
        self.tables = tables
        off = [0]
        rows = []
        for m in self.model_names:
            for g in tables[m]:
                row = np.zeros(self.P, np.int32)
                for k, v in g.items():
                    row[self.pid[k]] = v
                rows.append(row)
            off.append(len(rows))
        self.geom_off = np.asarray(off, np.int32)
        self.geom_counts = np.stack(rows).astype(np.int32) if rows else np.zeros((0, self.P), np.int32)
        rc = self.L.orc_set_mig_models(self.h, len(self.model_names), _p(self.geom_off, i32p),
                                       _p(self.geom_counts, i32p))
        if rc != 0:
            raise ValueError(self.L.orc_error(self.h).decode())

    def set_mps_sizes(self, sizes: Iterable[int]):
        self.mps_sizes = sorted(set(int(s) for s in sizes))
        self.S = len(self.mps_sizes)
        self.sid = {s: i for i, s in enumerate(self.mps_sizes)}
        a = np.asarray(self.mps_sizes, np.int32)
        if self.L.orc_set_mps_sizes(self.h, self.S, _p(a, i32p)) != 0:
            raise ValueError(self.L.orc_error(self.h).decode())

    # ---- dict <-> dense ------------------------------------------------------------------------
    def mig_dense(self, d: Dict[str, int]) -> Tuple[np.ndarray, int]:
        v = np.zeros(self.P, np.int32)
        m = 0
        for k, q in d.items():
            v[self.pid[k]] = q
            m |= 1 << self.pid[k]
        return v, m

    def mig_dict(self, v: np.ndarray, m: int) -> Dict[str, int]:
        return {self.mig_names[p]: int(v[p]) for p in range(self.P) if (m >> p) & 1}

    @staticmethod
    def mps_size(name: str) -> int:
        return int(lib().orc_mps_profile_size_gb(name.encode()))

    def mps_dense(self, d: Dict[str, int]) -> Tuple[np.ndarray, int]:
        v = np.zeros(max(self.S, 1), np.int32)
        m = 0
        for k, q in d.items():
            i = self.sid[self.mps_size(k)]
            v[i] = q
            m |= 1 << i
        return v, m

    def mps_dict(self, v: np.ndarray, m: int) -> Dict[str, int]:
        return {f"{self.mps_sizes[p]}gb": int(v[p]) for p in range(self.S) if (m >> p) & 1}

    # ---- misc ----------------------------------------------------------------------------------
    def geometry_id(self, g: Dict[str, int]) -> str:
        v, m = self.mig_dense(g)
        buf = C.create_string_buffer(1024)
        self.L.orc_geometry_id(self.h, _p(v, i32p), C.c_uint32(m), buf, 1024)
        return buf.value.decode()

    def fewest_slices(self, geoms: Sequence[Dict[str, int]]) -> int:
        masks = np.asarray([self.mig_dense(g)[1] for g in geoms], np.uint32)
        return int(self.L.orc_fewest_slices(len(geoms), _p(masks, u32p)))

    # ---- MIG GPU --------------------------------------------------------------------------------
    def mig_gpu_update(self, model: str, used: Dict[str, int], free: Dict[str, int], req: Dict[str, int]):
        uv, um = self.mig_dense(used)
        fv, fm = self.mig_dense(free)
        rv, rm = self.mig_dense(req)
        fmc = C.c_uint32(fm)
        ch = C.c_int32(-1)
        upd = self.L.orc_mig_gpu_update(self.h, self.mid[model], _p(uv, i32p), C.c_uint32(um), _p(fv, i32p),
                                        C.byref(fmc), _p(rv, i32p), C.c_uint32(rm), C.byref(ch))
        return bool(upd), self.mig_dict(fv, fmc.value), ch.value

    def mig_gpu_apply(self, model: str, used, free, geom):
        uv, um = self.mig_dense(used)
        fv, fm = self.mig_dense(free)
        gv, gm = self.mig_dense(geom)
        fmc = C.c_uint32(fm)
        err = self.L.orc_mig_gpu_apply_geometry(self.h, self.mid[model], _p(uv, i32p), C.c_uint32(um), _p(fv, i32p),
                                                C.byref(fmc), _p(gv, i32p), C.c_uint32(gm))
        return bool(err), self.mig_dict(fv, fmc.value)

    def mig_gpu_init(self, model: str, used, free):
        uv, um = self.mig_dense(used)
        fv, fm = self.mig_dense(free)
        fmc = C.c_uint32(fm)
        err = self.L.orc_mig_gpu_init_geometry(self.h, self.mid[model], _p(uv, i32p), C.c_uint32(um), _p(fv, i32p),
                                               C.byref(fmc))
        return bool(err), self.mig_dict(fv, fmc.value)

    def mig_gpu_add_pod(self, used, free, req):
        uv, um = self.mig_dense(used)
        fv, fm = self.mig_dense(free)
        rv, rm = self.mig_dense(req)
        umc, fmc = C.c_uint32(um), C.c_uint32(fm)
        err = self.L.orc_mig_gpu_add_pod(self.h, _p(uv, i32p), C.byref(umc), _p(fv, i32p), C.byref(fmc),
                                         _p(rv, i32p), C.c_uint32(rm))
        return bool(err), self.mig_dict(uv, umc.value), self.mig_dict(fv, fmc.value)

    # ---- MIG node: gpus = [(model, used, free), ...] ------------------------------------------------
    def _mig_node_arrays(self, gpus):
        G = len(gpus)
        models = np.zeros(max(G, 1), np.int32)
        used = np.zeros((max(G, 1), self.P), np.int32)
        free = np.zeros((max(G, 1), self.P), np.int32)
        um = np.zeros(max(G, 1), np.uint32)
        fm = np.zeros(max(G, 1), np.uint32)
        for i, (model, u, f) in enumerate(gpus):
            models[i] = self.mid[model]
            used[i], um[i] = self.mig_dense(u)
            free[i], fm[i] = self.mig_dense(f)
        return G, models, used, um, free, fm

    def mig_node_update(self, gpus, req):
        G, models, used, um, free, fm = self._mig_node_arrays(gpus)
        rv, rm = self.mig_dense(req)
        ch = np.full(max(G, 1), -1, np.int32)
        upd = self.L.orc_mig_node_update(self.h, G, _p(models, i32p), _p(used, i32p), _p(um, u32p), _p(free, i32p),
                                         _p(fm, u32p), _p(rv, i32p), C.c_uint32(rm), _p(ch, i32p))
        out = [(gpus[i][0], dict(gpus[i][1]), self.mig_dict(free[i], int(fm[i]))) for i in range(G)]
        return bool(upd), out, ch[:G].tolist()

    def mig_node_add_pod(self, gpus, req):
        G, models, used, um, free, fm = self._mig_node_arrays(gpus)
        rv, rm = self.mig_dense(req)
        slot = self.L.orc_mig_node_add_pod(self.h, G, _p(used, i32p), _p(um, u32p), _p(free, i32p), _p(fm, u32p),
                                           _p(rv, i32p), C.c_uint32(rm))
        out = [(gpus[i][0], self.mig_dict(used[i], int(um[i])), self.mig_dict(free[i], int(fm[i]))) for i in range(G)]
        return int(slot), out

    def mig_node_has_free_capacity(self, gpus) -> bool:
        G, models, used, um, free, fm = self._mig_node_arrays(gpus)
        return bool(self.L.orc_mig_node_has_free_capacity(self.h, G, _p(models, i32p), _p(used, i32p), _p(um, u32p),
                                                          _p(free, i32p), _p(fm, u32p)))

    @staticmethod
    def node_geometry(gpus) -> Dict[str, int]:
        """mig/node.go:113-120 and slicing/node.go:176-185: sum of used+free over GPUs (key union)."""
        res: Dict[str, int] = {}
        for _, u, f in gpus:
            for k, q in list(u.items()) + list(f.items()):
                res[k] = res.get(k, 0) + q
        return res

    # ---- MPS: gpus = [(mem_gb, used, free), ...] --------------------------------------------------
    def mps_gpu_validate(self, mem, used, free) -> bool:
        uv, um = self.mps_dense(used)
        fv, fm = self.mps_dense(free)
        return bool(self.L.orc_mps_gpu_validate(self.h, mem, _p(uv, i32p), C.c_uint32(um), _p(fv, i32p), C.c_uint32(fm)))

    def mps_gpu_update(self, mem, used, free, req):
        uv, um = self.mps_dense(used)
        fv, fm = self.mps_dense(free)
        rv, rm = self.mps_dense(req)
        fmc = C.c_uint32(fm)
        upd = self.L.orc_mps_gpu_update(self.h, mem, _p(uv, i32p), C.c_uint32(um), _p(fv, i32p), C.byref(fmc),
                                        _p(rv, i32p), C.c_uint32(rm))
        return bool(upd), self.mps_dict(fv, fmc.value)

    def _mps_node_arrays(self, gpus):
        G = len(gpus)
        S = max(self.S, 1)
        mem = np.zeros(max(G, 1), np.int32)
        used = np.zeros((max(G, 1), S), np.int32)
        free = np.zeros((max(G, 1), S), np.int32)
        um = np.zeros(max(G, 1), np.uint32)
        fm = np.zeros(max(G, 1), np.uint32)
        for i, (m, u, f) in enumerate(gpus):
            mem[i] = m
            used[i], um[i] = self.mps_dense(u)
            free[i], fm[i] = self.mps_dense(f)
        return G, mem, used, um, free, fm

    def mps_node_update(self, gpus, req):
        G, mem, used, um, free, fm = self._mps_node_arrays(gpus)
        rv, rm = self.mps_dense(req)
        gu = np.zeros(max(G, 1), np.uint8)
        upd = self.L.orc_mps_node_update(self.h, G, _p(mem, i32p), _p(used, i32p), _p(um, u32p), _p(free, i32p),
                                         _p(fm, u32p), _p(rv, i32p), C.c_uint32(rm), _p(gu, u8p))
        out = [(gpus[i][0], dict(gpus[i][1]), self.mps_dict(free[i], int(fm[i]))) for i in range(G)]
        return bool(upd), out

    def mps_node_add_pod(self, gpus, req):
        G, mem, used, um, free, fm = self._mps_node_arrays(gpus)
        rv, rm = self.mps_dense(req)
        slot = self.L.orc_mps_node_add_pod(self.h, G, _p(mem, i32p), _p(used, i32p), _p(um, u32p), _p(free, i32p),
                                           _p(fm, u32p), _p(rv, i32p), C.c_uint32(rm))
        out = [(gpus[i][0], self.mps_dict(used[i], int(um[i])), self.mps_dict(free[i], int(fm[i]))) for i in range(G)]
        return int(slot), out

    def mps_node_has_free_capacity(self, gpus) -> bool:
        G, mem, used, um, free, fm = self._mps_node_arrays(gpus)
        return bool(self.L.orc_mps_node_has_free_capacity(self.h, G, _p(mem, i32p), _p(used, i32p), _p(um, u32p),
                                                          _p(free, i32p), _p(fm, u32p)))

    # ---- flattened (array) level: snapshots are dicts of numpy arrays, see tests/synth.py ------------
    def _mig_snap(self, s) -> MigSnap:
        return MigSnap(int(s["n_nodes"]), _p(s["gpu_off"], i32p), _p(s["valid"], u8p), _p(s["model"], i32p),
                       _p(s["used"], i32p), _p(s["used_mask"], u32p), _p(s["free"], i32p), _p(s["free_mask"], u32p))

    def _mps_snap(self, s) -> MpsSnap:
        return MpsSnap(int(s["n_nodes"]), _p(s["gpu_off"], i32p), _p(s["valid"], u8p), _p(s["mem_gb"], i32p),
                       _p(s["used"], i32p), _p(s["used_mask"], u32p), _p(s["free"], i32p), _p(s["free_mask"], u32p))

    def score_mig(self, snap, req: np.ndarray, req_mask: np.ndarray, threads: int = 1, want_choice: bool = True):
        D = req.shape[0]
        N = int(snap["n_nodes"])
        tg = int(snap["gpu_off"][N])
        updated = np.zeros((D, N), np.uint8)
        choice = np.full((D, tg), -1, np.int8) if want_choice else None
        first = np.full(D, -1, np.int32)
        req = np.ascontiguousarray(req, np.int32)
        req_mask = np.ascontiguousarray(req_mask, np.uint32)
        st = self._mig_snap(snap)
        self.L.orc_score_mig(self.h, C.byref(st), D, _p(req, i32p), _p(req_mask, u32p), _p(updated, u8p),
                             _p(choice, i8p), _p(first, i32p), threads)
        return updated, choice, first

    def score_mps(self, snap, req: np.ndarray, req_mask: np.ndarray, threads: int = 1, want_free: bool = True):
        D = req.shape[0]
        N = int(snap["n_nodes"])
        tg = int(snap["gpu_off"][N])
        updated = np.zeros((D, N), np.uint8)
        out_free = np.zeros((D, tg, self.S), np.int32) if want_free else None
        out_fm = np.zeros((D, tg), np.uint32) if want_free else None
        first = np.full(D, -1, np.int32)
        req = np.ascontiguousarray(req, np.int32)
        req_mask = np.ascontiguousarray(req_mask, np.uint32)
        st = self._mps_snap(snap)
        self.L.orc_score_mps(self.h, C.byref(st), D, _p(req, i32p), _p(req_mask, u32p), _p(updated, u8p),
                             _p(out_free, i32p), _p(out_fm, u32p), _p(first, i32p), threads)
        return updated, out_free, out_fm, first

    def plan(self, mig_snap, mps_snap, pod_kind: np.ndarray, pod_req: np.ndarray, pod_req_mask: np.ndarray,
             gate: bool = True, sched: bool = True):
        """Mutates copies of the snapshots; returns (per-pod outputs, final mig snap, final mps snap)."""
        M = len(pod_kind)
        ms = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in mig_snap.items()} if mig_snap else None
        ps = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in mps_snap.items()} if mps_snap else None
        gate_hit = np.zeros(M, np.uint8)
        repart = np.full(M, -1, np.int32)
        snode = np.full(M, -1, np.int32)
        sgpu = np.full(M, -1, np.int32)
        pod_kind = np.ascontiguousarray(pod_kind, np.uint8)
        pod_req = np.ascontiguousarray(pod_req, np.int32)
        pod_req_mask = np.ascontiguousarray(pod_req_mask, np.uint32)
        mst = self._mig_snap(ms) if ms else None
        pst = self._mps_snap(ps) if ps else None
        flags = (1 if gate else 0) | (2 if sched else 0)
        self.L.orc_plan(self.h, C.byref(mst) if mst else None, C.byref(pst) if pst else None, M, _p(pod_kind, u8p),
                        _p(pod_req, i32p), _p(pod_req_mask, u32p), pod_req.shape[1] if pod_req.ndim == 2 else 0, flags,
                        _p(gate_hit, u8p), _p(repart, i32p), _p(snode, i32p), _p(sgpu, i32p))
        return dict(gate_hit=gate_hit, repart_node=repart, sched_node=snode, sched_gpu=sgpu), ms, ps

    def build_partitioning(self, snap, P: int) -> np.ndarray:
        tg = snap["used"].shape[0]
        out = np.zeros((tg, P), np.int32)
        self.L.orc_build_partitioning(tg, P, _p(snap["used"], i32p), _p(snap["used_mask"], u32p), _p(snap["free"], i32p),
                                      _p(snap["free_mask"], u32p), _p(out, i32p))
        return out


# ---- wire format (string level) --------------------------------------------------------------------
_STATUS = {0: "used", 1: "free", 2: "unknown"}


def parse_status_annotation(key: str, value: str):
    idx, st, q = C.c_int32(), C.c_int32(), C.c_int32()
    buf = C.create_string_buffer(256)
    if lib().orc_parse_status_annotation(key.encode(), value.encode(), C.byref(idx), buf, 256, C.byref(st), C.byref(q)):
        return None
    return dict(index=idx.value, profile=buf.value.decode(), status=_STATUS[st.value], quantity=q.value)


def parse_spec_annotation(key: str, value: str):
    idx, q = C.c_int32(), C.c_int32()
    buf = C.create_string_buffer(256)
    if lib().orc_parse_spec_annotation(key.encode(), value.encode(), C.byref(idx), buf, 256, C.byref(q)):
        return None
    return dict(index=idx.value, profile=buf.value.decode(), quantity=q.value)


def parse_node_annotations(annotations: Dict[str, str]):
    """pkg/gpu/annotation.go:87-101: spec is tried first, then status; others ignored."""
    status, spec = [], []
    for k, v in annotations.items():
        a = parse_spec_annotation(k, v)
        if a is not None:
            spec.append(a)
            continue
        a = parse_status_annotation(k, v)
        if a is not None:
            status.append(a)
    return status, spec


def memory_gb(label: str):
    out = C.c_int32()
    if lib().orc_memory_gb(label.encode(), C.byref(out)):
        return None
    return out.value


LABEL_PRODUCT = "nvidia.com/gpu.product"  # pkg/constant/constants.go:67-77
LABEL_COUNT = "nvidia.com/gpu.count"
LABEL_MEMORY = "nvidia.com/gpu.memory"


def _go_atoi(s: str):
    import re
    if not re.fullmatch(r"[+-]?[0-9]+", s):
        return None
    return int(s)


def extract_gpus(labels: Dict[str, str], annotations: Dict[str, str], known_models: Optional[Iterable[str]],
                 kind: str):
    """mig/node.go:40-100 NewNode+extractGPUs and slicing/node.go:32-105, under the canonical GPU order
    of SURVEY §D (annotated GPUs ascending by index, then "missing" GPUs with index len..count-1 — the
    duplicate-index quirk is kept).  Returns None where the reference returns an error, else a list
    of dict(index, used, free[, mem_gb])."""
    if LABEL_PRODUCT not in labels:
        return None
    model = labels[LABEL_PRODUCT]
    if LABEL_COUNT not in labels:
        return None
    count = _go_atoi(labels[LABEL_COUNT])
    if count is None:
        return None
    mem = None
    if kind == "mps":
        if LABEL_MEMORY not in labels:
            return None
        mem = memory_gb(labels[LABEL_MEMORY])
        if mem is None:
            return None
    status, _ = parse_node_annotations(annotations)
    by_idx: Dict[int, list] = {}
    for a in status:
        by_idx.setdefault(a["index"], []).append(a)
    gpus = []
    for idx in sorted(by_idx):
        used, free = {}, {}
        for a in by_idx[idx]:
            if a["status"] == "used":
                used[a["profile"]] = a["quantity"]
            if a["status"] == "free":
                free[a["profile"]] = a["quantity"]
        if kind == "mig":
            if known_models is not None and model not in known_models:
                return None  # NewGPU error, mig/gpu.go:46-49
            gpus.append(dict(index=idx, model=model, used=used, free=free))
        else:
            # slicing.NewGPU validates (slicing/gpu.go:67-97)
            tot = 0
            for d in (used, free):
                for k, q in d.items():
                    sz = Oracle.mps_size(k)
                    if sz < 1:
                        return None
                    tot += sz * q
            if tot > mem:
                return None
            gpus.append(dict(index=idx, model=model, mem_gb=mem, used=used, free=free))
    for i in range(len(gpus), count):
        if kind == "mig":
            if known_models is not None and model not in known_models:
                return None
            gpus.append(dict(index=i, model=model, used={}, free={}))
        else:
            gpus.append(dict(index=i, model=model, mem_gb=mem, used={}, free={}))
    return gpus


def mig_diff_counts(used: np.ndarray, free: np.ndarray, spec: np.ndarray, spec_mask: np.ndarray):
    """Count-level plan.NewMigConfigPlan (orc_mig_diff_counts): returns (create, del_free, del_used) [G][P]."""
    used = np.ascontiguousarray(used, np.int32)
    free = np.ascontiguousarray(free, np.int32)
    spec = np.ascontiguousarray(spec, np.int32)
    spec_mask = np.ascontiguousarray(spec_mask, np.uint32)
    G, P = used.shape
    create = np.zeros((G, P), np.int32)
    df = np.zeros((G, P), np.int32)
    du = np.zeros((G, P), np.int32)
    lib().orc_mig_diff_counts(G, P, _p(used, i32p), _p(free, i32p), _p(spec, i32p), _p(spec_mask, u32p), _p(create, i32p),
                              _p(df, i32p), _p(du, i32p))
    return create, df, du
