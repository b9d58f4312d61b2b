"""ctypes binding of libnosplan.so (include/nosplan.h).  Thin marshalling only: numpy arrays in,
numpy arrays out.  There is no fallback: if the library or a CUDA device is missing, it raises."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Dict, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.environ.get("NOSPLAN_LIB") or os.path.join(_HERE, "libnosplan.so")  # NOSPLAN_LIB: A/B builds of the same library

i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)
u8p = C.POINTER(C.c_uint8)
i8p = C.POINTER(C.c_int8)
i64p = C.POINTER(C.c_int64)

PLAN_GATE = 1
PLAN_SCHED = 2
PASS_SINGLE_KEY = 1
PASS_SERIAL = 2
PASS_COUNT = 4

EXPORTS = [
    "nosplan_create", "nosplan_destroy", "nosplan_last_error", "nosplan_version", "nosplan_set_mig_tables",
    "nosplan_set_mig_snapshot", "nosplan_set_mps_snapshot", "nosplan_score_mig", "nosplan_score_mps", "nosplan_plan",
    "nosplan_get_mig_state", "nosplan_get_mps_state", "nosplan_init_geometry_mig", "nosplan_score_mig_dev",
    "nosplan_score_mps_dev", "nosplan_mig_evals_per_demand", "nosplan_mps_evals_per_demand", "nosplan_score_mig_static", "nosplan_diff_plan_mig",
    "nosplan_set_node_base", "nosplan_exchange_create", "nosplan_exchange_open", "nosplan_exchange_min_dev",
    "nosplan_exchange_status", "nosplan_exchange_destroy", "nosplan_score_snapshot", "nosplan_presence",
    "nosplan_diff_plan_mig_devices", "nosplan_plan_batch", "nosplan_score_pass_dev", "nosplan_pass_status", "nosplan_kernel_timing", "nosplan_kernel_time_ms", "nosplan_executed_evals",
]


IPC_HANDLE_BYTES = 64


class _MigSnap(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("gpu_off", C.c_void_p), ("valid", C.c_void_p), ("model", C.c_void_p),
                ("used", C.c_void_p), ("free_", C.c_void_p)]


class _MpsSnap(C.Structure):
    _fields_ = [("n_sizes", C.c_int32), ("sizes_gb", C.c_void_p), ("n_nodes", C.c_int32), ("gpu_off", C.c_void_p),
                ("valid", C.c_void_p), ("mem_gb", C.c_void_p), ("used", C.c_void_p), ("free_", C.c_void_p),
                ("free_mask", C.c_void_p)]


class _Demands(C.Structure):
    _fields_ = [("n_rows", C.c_int32), ("off", C.c_void_p), ("key", C.c_void_p), ("qty", C.c_void_p)]


def _addr(a):
    return None if a is None else a.ctypes.data


class NosplanError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"nosplan error {code}: {msg}")
        self.code = code


def library_path() -> str:
    return _LIB


def build_library(force: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a … (csrc/Makefile).  Cross-compiles without a GPU."""
    src_dir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(src_dir, f) for f in os.listdir(src_dir)] + [os.path.join(_HERE, "..", "include", "nosplan.h")]
    stale = not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", src_dir, "all"] + (["-B"] if force else []), stdout=subprocess.DEVNULL,
                              stderr=subprocess.DEVNULL)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            raise FileNotFoundError(f"{_LIB} is missing: build it with __graft_entry__.build() (no CPU fallback exists)")
        L = C.CDLL(_LIB)
        L.nosplan_last_error.restype = C.c_char_p
        L.nosplan_version.restype = C.c_char_p
        L.nosplan_mig_evals_per_demand.restype = C.c_int64
        L.nosplan_mps_evals_per_demand.restype = C.c_int64
        L.nosplan_mig_evals_per_demand.argtypes = [C.c_void_p]
        L.nosplan_mps_evals_per_demand.argtypes = [C.c_void_p]
        L.nosplan_last_error.argtypes = [C.c_void_p]
        L.nosplan_destroy.argtypes = [C.c_void_p]
        _lib = L
    return _lib


_CZ = C.c_char * 0


def _p(a, t=None):
    """Pointer argument for a numpy array.  A zero-length ctypes view of the buffer (0.9 us) instead of
    numpy's .ctypes machinery (4 us): the wrapper is inside bench.py's end-to-end timing."""
    if a is None:
        return None
    if a.size and a.flags.writeable:
        return _CZ.from_buffer(a)
    return C.c_void_p(a.ctypes.data)


def dense_to_csr(req: np.ndarray, mask: Optional[np.ndarray] = None):
    """[D][K] dense request rows (+ optional key-presence masks) -> CSR (off, key, qty); keys ascending."""
    req = np.asarray(req, np.int32)
    D, K = req.shape if req.ndim == 2 else (0, 0)
    off = [0]
    key: List[int] = []
    qty: List[int] = []
    for d in range(D):
        for k in range(K):
            present = ((int(mask[d]) >> k) & 1) if mask is not None else (req[d, k] != 0)
            if present:
                key.append(k)
                qty.append(int(req[d, k]))
        off.append(len(key))
    return (np.asarray(off, np.int32), np.asarray(key if key else [0], np.int32)[:max(len(key), 1)],
            np.asarray(qty if qty else [0], np.int32)[:max(len(qty), 1)])


class Planner:
    """One libnosplan handle bound to a CUDA device."""

    def __init__(self, device: int = 0):
        self.L = lib()
        h = C.c_void_p()
        rc = self.L.nosplan_create(int(device), C.byref(h))
        if rc != 0:
            raise NosplanError(rc, "nosplan_create failed (no CUDA device? there is no CPU fallback)")
        self.h = h
        self.P = 0
        self.S = 0
        self.mig = None
        self.mps = None

    def close(self):
        if getattr(self, "h", None):
            self.L.nosplan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int):
        if rc != 0:
            raise NosplanError(rc, self.L.nosplan_last_error(self.h).decode())

    # ---- tables / snapshots ----------------------------------------------------------------------
    def set_mig_tables(self, profile_names: Sequence[str], geom_off: np.ndarray, geom_counts: np.ndarray):
        names = (C.c_char_p * len(profile_names))(*[n.encode() for n in profile_names])
        geom_off = np.ascontiguousarray(geom_off, np.int32)
        geom_counts = np.ascontiguousarray(geom_counts, np.int32)
        self._ck(self.L.nosplan_set_mig_tables(self.h, len(profile_names), names, len(geom_off) - 1, _p(geom_off, i32p),
                                               _p(geom_counts, i32p)))
        self.P = len(profile_names)

    def set_mig_snapshot(self, snap: Dict[str, np.ndarray]):
        # ascontiguousarray does not copy when dtype and layout already match (pinned buffers stay pinned)
        a = lambda k, t: np.ascontiguousarray(snap[k], t)
        s = dict(n_nodes=snap["n_nodes"], gpu_off=a("gpu_off", np.int32), valid=a("valid", np.uint8),
                 model=a("model", np.int32), used=a("used", np.int32), free=a("free", np.int32))
        self.mig = None  # a rejected snapshot leaves the handle without one
        self._ck(self.L.nosplan_set_mig_snapshot(self.h, int(s["n_nodes"]), _p(s["gpu_off"], i32p), _p(s["valid"], u8p),
                                                 _p(s["model"], i32p), _p(s["used"], i32p), _p(s["free"], i32p)))
        go = s["gpu_off"]
        if getattr(self, "_gs_of", None) is not go:  # the same offsets array again: keep its widest node
            self._gs_of, self._gs = go, (int(np.max(np.diff(go))) if int(s["n_nodes"]) else 0)
        self.mig = dict(n_nodes=int(s["n_nodes"]), tg=int(go[int(s["n_nodes"])]), gs=max(self._gs, 1))

    def set_mps_snapshot(self, sizes_gb: Sequence[int], snap: Dict[str, np.ndarray], with_mask: bool = True):
        a = lambda k, t: np.ascontiguousarray(snap[k], t)
        s = dict(n_nodes=snap["n_nodes"], gpu_off=a("gpu_off", np.int32), valid=a("valid", np.uint8),
                 mem_gb=a("mem_gb", np.int32), used=a("used", np.int32), free=a("free", np.int32))
        sizes = np.ascontiguousarray(sizes_gb, np.int32)
        fm = a("free_mask", np.uint32) if (with_mask and "free_mask" in snap) else None
        self.mps = None
        self._ck(self.L.nosplan_set_mps_snapshot(self.h, len(sizes), _p(sizes, i32p), int(s["n_nodes"]),
                                                 _p(s["gpu_off"], i32p), _p(s["valid"], u8p), _p(s["mem_gb"], i32p),
                                                 _p(s["used"], i32p), _p(s["free"], i32p), _p(fm, u32p)))
        self.S = len(sizes)
        self.mps = dict(n_nodes=int(s["n_nodes"]), tg=int(s["gpu_off"][int(s["n_nodes"])]))

    # ---- scorers -----------------------------------------------------------------------------------
    def score_mig(self, off, key, qty, want_updated=True, want_choice=True, want_first_choice=True):
        D = len(off) - 1
        if self.mig is None:
            raise NosplanError(-1, "no MIG snapshot")
        N, tg, gs = self.mig["n_nodes"], self.mig["tg"], self.mig["gs"]
        first = np.empty(D, np.int32)  # every output is written in full by the library
        fch = np.empty((D, gs), np.int8) if want_first_choice else None
        upd = np.empty((D, N), np.uint8) if want_updated else None
        ch = np.empty((D, tg), np.int8) if want_choice else None
        ne = C.c_int64(0)
        off, key, qty = (np.ascontiguousarray(x, np.int32) for x in (off, key, qty))
        self._ck(self.L.nosplan_score_mig(self.h, D, _p(off, i32p), _p(key, i32p), _p(qty, i32p), _p(first, i32p),
                                          _p(fch, i8p), _p(upd, u8p), _p(ch, i8p), C.byref(ne)))
        return dict(first_node=first, first_choice=fch, updated=upd, choice=ch, n_evals=ne.value)

    def score_mps(self, off, key, qty, want_updated=True, want_free=True):
        D = len(off) - 1
        if self.mps is None:
            raise NosplanError(-1, "no MPS snapshot")
        N, tg = self.mps["n_nodes"], self.mps["tg"]
        first = np.empty(D, np.int32)
        upd = np.empty((D, N), np.uint8) if want_updated else None
        of = np.empty((D, tg, self.S), np.uint8) if want_free else None
        om = np.empty((D, tg), np.uint32) if want_free else None
        ne = C.c_int64(0)
        off, key, qty = (np.ascontiguousarray(x, np.int32) for x in (off, key, qty))
        self._ck(self.L.nosplan_score_mps(self.h, D, _p(off, i32p), _p(key, i32p), _p(qty, i32p), _p(first, i32p),
                                          _p(upd, u8p), _p(of, u8p), _p(om, u32p), C.byref(ne)))
        return dict(first_node=first, updated=upd, out_free=of, out_free_mask=om, n_evals=ne.value)

    _SNAP_DT = dict(gpu_off=np.int32, valid=np.uint8, model=np.int32, used=np.int32, free=np.int32, mem_gb=np.int32,
                    free_mask=np.uint32)

    @classmethod
    def _as_abi(cls, d, keys):
        """The snapshot's arrays in the dtype and layout the ABI takes; arrays that already match are passed as they
        are (pinned buffers stay pinned), the others are converted."""
        out = []
        for k in keys:
            v = d.get(k)
            if v is not None:
                t = cls._SNAP_DT[k]
                if v.dtype != t or not v.flags.c_contiguous:
                    v = np.ascontiguousarray(v, t)
            out.append(v)
        return out

    @staticmethod
    def _rows_abi(rows):
        return tuple(x if (x.dtype == np.int32 and x.flags.c_contiguous) else np.ascontiguousarray(x, np.int32) for x in rows)

    def score_snapshot(self, mig=None, mig_rows=None, sizes_gb=None, mps=None, mps_rows=None, want_first_choice=True):
        """One call for a whole scoring pass (nosplan_score_snapshot): mig / mps are snapshot dicts as for
        set_*_snapshot, *_rows = (off, key, qty).  Nothing is installed on the handle.  With an exchange open on the
        handle the winners are indices over the whole sharded node list (the call is collective)."""
        ms = ps = md = pd = None
        out = {}
        fm = ff = fp = None
        keep = []
        if mig is not None:
            N = int(mig["n_nodes"])
            arrs = self._as_abi(mig, ("gpu_off", "valid", "model", "used", "free"))
            off, key, qty = rows = self._rows_abi(mig_rows)
            keep += [arrs, rows]
            ms = _MigSnap(N, *[_addr(a) for a in arrs])
            md = _Demands(len(off) - 1, off.ctypes.data, key.ctypes.data, qty.ctypes.data)
            go = arrs[0]
            if getattr(self, "_gs_of", None) is not go:  # the same offsets array again: keep its widest node
                self._gs_of, self._gs = go, (int(np.max(np.diff(go))) if N else 0)
            fm = np.empty(len(off) - 1, np.int32)
            ff = np.empty((len(off) - 1, max(self._gs, 1)), np.int8) if want_first_choice else None
            out["mig_first_node"] = fm
            out["mig_first_choice"] = ff
        if mps is not None:
            sizes = np.ascontiguousarray(sizes_gb, np.int32)
            arrs = self._as_abi(mps, ("gpu_off", "valid", "mem_gb", "used", "free", "free_mask"))
            off, key, qty = rows = self._rows_abi(mps_rows)
            keep += [sizes, arrs, rows]
            ps = _MpsSnap(len(sizes), sizes.ctypes.data, int(mps["n_nodes"]), *[_addr(a) for a in arrs])
            pd = _Demands(len(off) - 1, off.ctypes.data, key.ctypes.data, qty.ctypes.data)
            fp = np.empty(len(off) - 1, np.int32)
            out["mps_first_node"] = fp
        ne = C.c_int64(0)
        ref = lambda x: None if x is None else C.byref(x)
        self._ck(self.L.nosplan_score_snapshot(self.h, ref(ms), ref(md), _p(fm), _p(ff), ref(ps), ref(pd), _p(fp), C.byref(ne)))
        out["n_evals"] = ne.value
        return out

    # ---- planner -------------------------------------------------------------------------------------
    def plan(self, kind, off, key, qty, gate=True, sched=True):
        M = len(kind)
        kind = np.ascontiguousarray(kind, np.uint8)
        off, key, qty = (np.ascontiguousarray(x, np.int32) for x in (off, key, qty))
        gate_hit = np.zeros(M, np.uint8)
        rep = np.full(M, -2, np.int32)
        node = np.full(M, -2, np.int32)
        gpu = np.full(M, -2, np.int32)
        flags = (PLAN_GATE if gate else 0) | (PLAN_SCHED if sched else 0)
        self._ck(self.L.nosplan_plan(self.h, M, _p(kind, u8p), _p(off, i32p), _p(key, i32p), _p(qty, i32p), flags,
                                     _p(gate_hit, u8p), _p(rep, i32p), _p(node, i32p), _p(gpu, i32p)))
        return dict(gate_hit=gate_hit, repart_node=rep, sched_node=node, sched_gpu=gpu)

    def plan_batch(self, kind: int, pod_key, pod_qty, max_rounds: int = 0):
        """Batch-demand planning mode (nosplan_plan_batch; new behaviour, defined in oracle/batch_plan.py)."""
        pod_key = np.ascontiguousarray(pod_key, np.int32)
        pod_qty = np.ascontiguousarray(pod_qty, np.int32)
        M = len(pod_key)
        N = (self.mps if kind else self.mig)["n_nodes"]
        node, gpu = np.full(max(M, 1), -2, np.int32), np.full(max(M, 1), -2, np.int32)
        phase = np.zeros(max(M, 1), np.uint8)
        rep = np.zeros(max(N, 1), np.uint8)
        order, rounds = C.c_int32(0), C.c_int32(0)
        self._ck(self.L.nosplan_plan_batch(self.h, int(kind), M, _p(pod_key), _p(pod_qty), int(max_rounds), _p(node), _p(gpu),
                                           _p(phase), _p(rep), C.byref(order), C.byref(rounds)))
        return dict(sched_node=node[:M], sched_gpu=gpu[:M], phase=phase[:M], repartitioned=rep[:N], order=order.value,
                    rounds=rounds.value)

    def get_mig_state(self):
        tg = self.mig["tg"]
        used = np.zeros((tg, self.P), np.int32)
        free = np.zeros((tg, self.P), np.int32)
        self._ck(self.L.nosplan_get_mig_state(self.h, _p(used, i32p), _p(free, i32p)))
        return used, free

    def get_mps_state(self):
        tg = self.mps["tg"]
        used = np.zeros((tg, self.S), np.int32)
        free = np.zeros((tg, self.S), np.int32)
        fm = np.zeros(tg, np.uint32)
        self._ck(self.L.nosplan_get_mps_state(self.h, _p(used, i32p), _p(free, i32p), _p(fm, u32p)))
        return used, free, fm

    def init_geometry_mig(self):
        tg = self.mig["tg"]
        err = np.zeros(tg, np.uint8)
        free = np.zeros((tg, self.P), np.int32)
        self._ck(self.L.nosplan_init_geometry_mig(self.h, _p(err, u8p), _p(free, i32p)))
        return err, free

    def diff_plan_mig(self, used, free, spec, spec_mask):
        used, free, spec = (np.ascontiguousarray(x, np.int32) for x in (used, free, spec))
        spec_mask = np.ascontiguousarray(spec_mask, np.uint32)
        G, P = used.shape
        create, df, du = (np.zeros((G, P), np.int32) for _ in range(3))
        self._ck(self.L.nosplan_diff_plan_mig(self.h, G, P, _p(used, i32p), _p(free, i32p), _p(spec, i32p),
                                              _p(spec_mask, u32p), _p(create, i32p), _p(df, i32p), _p(du, i32p)))
        return create, df, du

    def diff_plan_mig_devices(self, dev_off, dev_profile, dev_status, spec, spec_mask):
        """plan.NewMigConfigPlan at device level (nosplan_diff_plan_mig_devices): returns dev_op [n_devices],
        create_new / create_again [G][P], matches [G]."""
        dev_off = np.ascontiguousarray(dev_off, np.int32)
        G = len(dev_off) - 1
        nd = int(dev_off[-1]) if G else 0
        dev_profile = np.ascontiguousarray(dev_profile if nd else [0], np.int32)
        dev_status = np.ascontiguousarray(dev_status if nd else [0], np.uint8)
        spec = np.ascontiguousarray(spec, np.int32)
        spec_mask = np.ascontiguousarray(spec_mask, np.uint32)
        P = spec.shape[1]
        op = np.zeros(max(nd, 1), np.uint8)
        cn, ca = np.zeros((G, P), np.int32), np.zeros((G, P), np.int32)
        m = np.zeros(max(G, 1), np.uint8)
        self._ck(self.L.nosplan_diff_plan_mig_devices(self.h, G, P, _p(dev_off), _p(dev_profile), _p(dev_status), _p(spec), _p(spec_mask),
                                                      _p(op), _p(cn), _p(ca), _p(m)))
        return op[:nd], cn, ca, m[:G]

    # ---- device-resident (pointers are ints: torch tensor .data_ptr()) -------------------------------
    def score_mig_dev(self, D, off_ptr, key_ptr, qty_ptr, first_ptr, updated_ptr=0, choice_ptr=0, stream=0):
        self._ck(self.L.nosplan_score_mig_dev(self.h, int(D), C.c_void_p(off_ptr), C.c_void_p(key_ptr), C.c_void_p(qty_ptr),
                                              C.c_void_p(first_ptr), C.c_void_p(updated_ptr or None),
                                              C.c_void_p(choice_ptr or None), C.c_void_p(stream or None)))

    def score_mps_dev(self, D, off_ptr, key_ptr, qty_ptr, first_ptr, updated_ptr=0, free_ptr=0, mask_ptr=0, stream=0):
        self._ck(self.L.nosplan_score_mps_dev(self.h, int(D), C.c_void_p(off_ptr), C.c_void_p(key_ptr), C.c_void_p(qty_ptr),
                                              C.c_void_p(first_ptr), C.c_void_p(updated_ptr or None),
                                              C.c_void_p(free_ptr or None), C.c_void_p(mask_ptr or None),
                                              C.c_void_p(stream or None)))

    def score_pass_dev(self, Dm, m_off, m_key, m_qty, Dp, p_off, p_key, p_qty, first_ptr, flags=0, stream=0):
        """Both kinds in one call (nosplan_score_pass_dev); pointers are device addresses (ints), 0 = none."""
        v = lambda x: C.c_void_p(x or None)
        self._ck(self.L.nosplan_score_pass_dev(self.h, int(Dm), v(m_off), v(m_key), v(m_qty), int(Dp), v(p_off), v(p_key), v(p_qty),
                                               v(first_ptr), int(flags), v(stream)))

    def pass_status(self):
        self._ck(self.L.nosplan_pass_status(self.h))

    def kernel_timing(self, enable: bool):
        self._ck(self.L.nosplan_kernel_timing(self.h, 1 if enable else 0))

    def kernel_time_ms(self, kind: int):
        """(summed ms, launches) of the kind's dominant scorer kernel since the last read (synchronises)."""
        ms, n = C.c_double(0), C.c_int(0)
        self._ck(self.L.nosplan_kernel_time_ms(self.h, int(kind), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def executed_evals(self):
        """(MIG, MPS) evaluations counted by passes run with PASS_COUNT since the last read (synchronises)."""
        a, b = C.c_int64(0), C.c_int64(0)
        self._ck(self.L.nosplan_executed_evals(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def score_mig_static(self, n_nodes, gpus_per_node, valid_ptr, model_ptr, used_ptr, free_ptr, D, off_ptr, key_ptr, qty_ptr,
                         first_ptr, stream=0):
        """Allocation-free scorer: every pointer is a device address owned by the caller."""
        self._ck(self.L.nosplan_score_mig_static(self.h, int(n_nodes), int(gpus_per_node), C.c_void_p(valid_ptr or None),
                                                 C.c_void_p(model_ptr or None), C.c_void_p(used_ptr or None),
                                                 C.c_void_p(free_ptr or None), int(D), C.c_void_p(off_ptr or None),
                                                 C.c_void_p(key_ptr or None), C.c_void_p(qty_ptr or None),
                                                 C.c_void_p(first_ptr or None), C.c_void_p(stream or None)))

    # ---- sharded scoring (one process per GPU) -----------------------------------------------------
    def set_node_base(self, mig_first_node: int, mps_first_node: int):
        self._ck(self.L.nosplan_set_node_base(self.h, int(mig_first_node), int(mps_first_node)))

    def presence(self, kind: int) -> np.ndarray:
        """Sum of used + free per key over the valid nodes of the handle's working state (the gate's ingredient)."""
        out = np.zeros(self.S if kind else self.P, np.int64)
        self._ck(self.L.nosplan_presence(self.h, int(kind), _p(out)))
        return out

    def exchange_create(self, rank: int, world: int, max_keys: int) -> bytes:
        """Allocates this rank's exchange buffer; returns its 64-byte CUDA IPC handle."""
        out = (C.c_ubyte * IPC_HANDLE_BYTES)()
        self._ck(self.L.nosplan_exchange_create(self.h, int(rank), int(world), int(max_keys), out))
        return bytes(out)

    def exchange_open(self, handles: Sequence[bytes]):
        blob = b"".join(handles)
        buf = (C.c_ubyte * max(len(blob), 1)).from_buffer_copy(blob or b"\0")
        self._ck(self.L.nosplan_exchange_open(self.h, buf))

    def exchange_min_dev(self, n_keys, in_ptr, out_ptr, stream=0):
        self._ck(self.L.nosplan_exchange_min_dev(self.h, int(n_keys), C.c_void_p(in_ptr or None), C.c_void_p(out_ptr or None),
                                                 C.c_void_p(stream or None)))

    def exchange_status(self):
        self._ck(self.L.nosplan_exchange_status(self.h))

    def exchange_destroy(self):
        self._ck(self.L.nosplan_exchange_destroy(self.h))

    def mig_evals_per_demand(self) -> int:
        return int(self.L.nosplan_mig_evals_per_demand(self.h))

    def mps_evals_per_demand(self) -> int:
        return int(self.L.nosplan_mps_evals_per_demand(self.h))
