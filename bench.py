#!/usr/bin/env python
"""bench.py — planner hot-path benchmark (contract: see the task statement / DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--impl b200|reference]

A "step" is one pass of the batched scorer over one cluster snapshot: every pending pod's request
vector is scored against every node (MIG pods x MIG nodes, MPS pods x MPS nodes).  The metric is
candidate evaluations per second (one evaluation = one iteration of pkg/gpu/mig/gpu.go:168-188 for
a (demand, node, GPU, allowed geometry) tuple; for MPS one greedy run per (demand, node, GPU)).
Planner latency (Plan(snapshot, pods) through the C ABI, p50) is reported in the same line.

N > 1: one process per GPU (torchrun), nodes sharded contiguously across ranks, pods replicated; the per-row
"first node that would update" keys are exchanged by the library itself (the scorers' epilogues atomicMin their
winners into every rank's buffer over NVLink, one finishing kernel reads the minimum: "first node wins"), and
every run checks the sharded answer against an unsharded GPU pass and the CPU oracle before it is timed.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_EVAL_MIG = 68.0 / 12.0  # (64 B state read + 4 B result) / C, C = 12
NOMINAL_PEAK_FALLBACK = 6650.0    # B200_PROFILING.md fallback (GB/s)


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return NOMINAL_PEAK_FALLBACK, "fallback"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU, sampled every 50 ms while the timed regions run: NVML in-process
    (a fork of nvidia-smi every 100 ms measurably slowed the host-bound end-to-end loop), nvidia-smi if NVML is missing."""
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []  # (sm_mhz, sm_max_mhz, [reason flags])
        self.stop_flag = False
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index(index))
        except Exception:
            self.nvml = None

    @staticmethod
    def _physical_index(index):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x.strip() for x in vis.split(",") if x.strip()]
            if index < len(ids) and ids[index].isdigit():
                return int(ids[index])
        return index

    def _sample_nvml(self):
        n = self.nvml
        sm = n.nvmlDeviceGetClockInfo(self.dev, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(self.dev, n.NVML_CLOCK_SM)
        r = n.nvmlDeviceGetCurrentClocksEventReasons(self.dev) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
            else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev)
        bits = [getattr(n, "nvmlClocksEventReasonHwSlowdown", 0x8), getattr(n, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                getattr(n, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), getattr(n, "nvmlClocksEventReasonSwPowerCap", 0x4)]
        self.rows.append((int(sm), int(mx), [bool(r & b) for b in bits]))

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            f = [x.strip() for x in out.split(",")]
            if f[0].isdigit() and f[1].isdigit():
                self.rows.append((int(f[0]), int(f[1]), [x.lower().startswith("active") for x in f[2:6]]))

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            time.sleep(0.02 if self.nvml else 0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for i, n in enumerate(self.NAMES) if any(r[2][i] for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(r[1] for r in self.rows), "reasons": reasons,
                "samples": len(self.rows), "source": "nvml" if self.nvml else "nvidia-smi"}


def run_reference(args, cfg, rank):
    """--impl reference: the reference's CPU implementation of the path.  The reference is Go and
    cannot be built in this image (no Go toolchain), so this leg times the C++ oracle
    port (oracle/) with all host threads on a bounded sample of the same workload."""
    if rank != 0:
        return
    import walkai_nos_b200 as pkg
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    mo, mk, mq, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
    po, pk, pq, _ = pkg.synth.pods_of_kind(cfg["pods"], 1)
    # bounded sample: the first `sample` pods of each kind against the full cluster
    sample = args.ref_sample
    Dm, Dp = min(sample, len(mo) - 1), min(sample, len(po) - 1)
    mreq, mmask = pkg.synth.csr_to_dense(mo[:Dm + 1], mk, mq, len(cfg["names"]))
    preq, pmask = pkg.synth.csr_to_dense(po[:Dp + 1], pk, pq, len(cfg["sizes"]))
    ev_m = Dm * sum(len(cfg["tables"][cfg["model_names"][m]]) for m in cfg["mig"]["model"]) if cfg["mig"] else 0
    ev_p = Dp * int(cfg["mps"]["gpu_off"][-1]) if cfg["mps"] else 0

    def step():
        if cfg["mig"] and Dm:
            o.score_mig(cfg["mig"], mreq, mmask, threads=cores, want_choice=False)
        if cfg["mps"] and Dp:
            o.score_mps(cfg["mps"], preq, pmask, threads=cores, want_free=False)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = (ev_m + ev_p) / dt
    line = {
        "impl": "reference", "metric": "candidate_evaluations_per_second", "value": val, "unit": "evals/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak" if args.weak else "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
        "config": workload_config(cfg, args),
        "cpu_baseline": {"value": val, "unit": "evals/s", "cores": cores, "kind": "port",
                         "sample": f"first {Dm} MIG + {Dp} MPS pods x all nodes per step (Go reference unbuildable here: no Go toolchain)"},
        "e2e": {"value": val, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def oracle_agrees_on_distinct_rows(pkg, cfg, kind, rows, first_node):
    """first_node of every row == the oracle's answer for the row's request vector (one oracle row per distinct
    vector; bench rows are single-key)."""
    from oracle import oracle as orc
    off, key, qty = rows
    D = len(off) - 1
    if not np.all(np.diff(off) == 1):
        raise SystemExit("bench.py: the parity check expects single-key rows")
    sig = key[:D].astype(np.int64) * 1000003 + qty[:D]
    uniq, fi = np.unique(sig, return_index=True)
    o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
    width = len(cfg["sizes"]) if kind else len(cfg["names"])
    req, mask = pkg.synth.csr_to_dense(np.arange(len(uniq) + 1, dtype=np.int32), key[fi], qty[fi], width)
    if kind == 0:
        _, _, ef = o.score_mig(cfg["mig"], req, mask, threads=os.cpu_count() or 1, want_choice=False)
    else:
        _, _, _, ef = o.score_mps(cfg["mps"], req, mask, threads=os.cpu_count() or 1, want_free=False)
    return bool(np.array_equal(ef[np.searchsorted(uniq, sig)], np.asarray(first_node)[:D]))


def workload_config(cfg, args):
    nm = int(cfg["mig"]["n_nodes"]) if cfg["mig"] else 0
    np_ = int(cfg["mps"]["n_nodes"]) if cfg["mps"] else 0
    return {"workload": f"BASELINE configs[{cfg['cfg'] - 1}]: {nm + np_} nodes x 8 GPUs ({nm} MIG H100 table + {np_} MPS 82GB), "
                        f"{len(cfg['pods']['kind'])} pending pods, all allowed geometries",
            "nodes": nm + np_, "pods": int(len(cfg["pods"]["kind"])), "l2": "flushed between timed steps (256 MiB write)",
            "sharding": f"nodes/{args.gpus} contiguous, pods replicated"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-sample", type=int, default=4096)
    ap.add_argument("--nodes", type=int, default=0)
    ap.add_argument("--pods", type=int, default=0)
    ap.add_argument("--weak", action="store_true", help="weak scaling: the config's node count PER GPU (nodes x N in all), pods unchanged; the line says scaling=weak")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-plan", action="store_true", help="skip the planner-latency section (profiling runs)")
    ap.add_argument("--loopback-exchange", action="store_true", help="one GPU: open a one-rank exchange (measures what the exchange path costs without peers)")
    ap.add_argument("--shard-of", type=int, default=1, help="one GPU, rank 0's shard of a W-rank run (ncu captures of the shard sizes)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    import walkai_nos_b200 as pkg
    if args.weak and not args.nodes:
        args.nodes = {1: 1, 2: 64, 3: 256, 4: 1024, 5: 4096}[args.config] * max(args.gpus, 1)
    cfg = pkg.synth.make_config(args.config, args.nodes, args.pods)

    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the planner has no CPU path (use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries the JSON line only
        dist.init_process_group("nccl", device_id=dev)

    planner = pkg.Planner(local_rank)
    planner.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
    mig, mig_lo = pkg.dist.shard_snapshot(cfg["mig"], rank, world if world > 1 else args.shard_of)
    mps, mps_lo = pkg.dist.shard_snapshot(cfg["mps"], rank, world if world > 1 else args.shard_of)
    if mig is not None:
        planner.set_mig_snapshot(mig)
    if mps is not None:
        planner.set_mps_snapshot(cfg["sizes"], mps)
    mo, mk, mq, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
    po, pk, pq, _ = pkg.synth.pods_of_kind(cfg["pods"], 1)
    Dm, Dp = (len(mo) - 1 if mig is not None else 0), (len(po) - 1 if mps is not None else 0)
    ev_m_local = Dm * planner.mig_evals_per_demand()
    ev_p_local = Dp * planner.mps_evals_per_demand()

    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    d_mo, d_mk, d_mq = t(mo), t(mk), t(mq)
    d_po, d_pk, d_pq = t(po), t(pk), t(pq)
    Dall = Dm + Dp
    # one step = ONE library call (nosplan_score_pass_dev): both kinds grouped by one launch, scored side by side;
    # N > 1: the scorers' epilogues fold their winners straight into every rank's exchange buffer over NVLink and a
    # finishing kernel hands back the minimum over ranks ("first node wins" over the whole list)
    first_glob = torch.full((max(Dall, 1),), -1, dtype=torch.int32, device=dev)
    first_loc = torch.full((max(Dall, 1),), -1, dtype=torch.int32, device=dev)
    exchange = "none"
    if world > 1:
        exchange = "peer atomicMin over NVLink, pushed from the scorers' epilogue (libnosplan)"
        try:
            pkg.dist.open_exchange(planner, max(Dall, 1))
        except pkg.NosplanError as e:  # no peer mapping on this box: say so, use the library collective
            exchange = f"nccl all-reduce (peer mapping failed: {e})"
        flags_t = torch.tensor([0 if exchange.startswith("peer") else 1], device=dev)
        dist.all_reduce(flags_t, op=dist.ReduceOp.MAX)
        if int(flags_t.item()) and exchange.startswith("peer"):
            exchange = "nccl all-reduce (a peer could not map)"
            planner.exchange_destroy()
        planner.set_node_base(mig_lo, mps_lo)  # the scorers report global node indices
    if world == 1 and args.loopback_exchange:
        planner.exchange_open([planner.exchange_create(0, 1, max(Dall, 1))])
        exchange = "one-rank loopback"
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    PASS = pkg.nosplan.PASS_SINGLE_KEY  # every bench row has one key (checked on the device, nosplan_pass_status)
    ptrs = (Dm, d_mo.data_ptr(), d_mk.data_ptr(), d_mq.data_ptr(), Dp, d_po.data_ptr(), d_pk.data_ptr(), d_pq.data_ptr())

    def step_device(ev=None, flags=PASS):
        """inputs resident in HBM; ev = (e0, e1) brackets the step on the launching stream"""
        if ev:
            ev[0].record()
        if world > 1 and not exchange.startswith("peer"):
            planner.score_pass_dev(*ptrs, first_loc.data_ptr(), flags, stream)
            first_glob.copy_(pkg.dist.merge_first_node_keys(first_loc, torch.zeros_like(first_loc)))
        else:
            planner.score_pass_dev(*ptrs, first_glob.data_ptr(), flags, stream)
        if ev:
            ev[1].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush.zero_()
        step_device()
    barrier()
    planner.pass_status()
    if world > 1 and exchange.startswith("peer"):
        # cross-check of the fused exchange against an NCCL all-reduce(MIN) of the separately scored shards
        if Dm:
            planner.score_mig_dev(Dm, d_mo.data_ptr(), d_mk.data_ptr(), d_mq.data_ptr(), first_loc.data_ptr(), 0, 0, stream)
        if Dp:
            planner.score_mps_dev(Dp, d_po.data_ptr(), d_pk.data_ptr(), d_pq.data_ptr(), first_loc[Dm:].data_ptr(), 0, 0, 0, stream)
        ref = pkg.dist.merge_first_node_keys(first_loc, torch.zeros_like(first_loc))
        if not torch.equal(ref, first_glob):
            raise SystemExit("bench.py: the peer-memory exchange disagrees with the NCCL all-reduce")
    sharded_parity = None
    full = None
    if world > 1:
        # The sharded answer must be the unsharded one: rank 0 scores the WHOLE snapshot on its GPU (one handle, no
        # exchange) and checks that against the oracle on the distinct request vectors; every rank then compares
        # what the sharded pass + exchange left in first_glob with rank 0's unsharded winners.
        full = torch.full_like(first_glob, -1)
        oracle_ok = 1
        if rank == 0:
            fp = pkg.Planner(local_rank)
            fp.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
            parts = []
            for kind, snap, rows, D in ((0, cfg["mig"], (mo, mk, mq), Dm), (1, cfg["mps"], (po, pk, pq), Dp)):
                if snap is None or not D:
                    continue
                if kind == 0:
                    fp.set_mig_snapshot(snap)
                    got = fp.score_mig(*rows, want_updated=False, want_choice=False, want_first_choice=False)["first_node"]
                else:
                    fp.set_mps_snapshot(cfg["sizes"], snap)
                    got = fp.score_mps(*rows, want_updated=False, want_free=False)["first_node"]
                parts.append(got)
                oracle_ok &= int(oracle_agrees_on_distinct_rows(pkg, cfg, kind, rows, got))
            fp.close()
            if parts:
                full = torch.from_numpy(np.concatenate(parts)).to(dev)
        dist.broadcast(full, src=0)
        same = torch.tensor([int(torch.equal(full, first_glob)), oracle_ok], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        sharded_parity = {"vs_unsharded_gpu": "ok" if int(same[0]) else "MISMATCH",
                          "vs_oracle_dedup": "ok" if int(same[1]) else "MISMATCH", "rows": int(Dall)}
        if not int(same[0]) or not int(same[1]):
            raise SystemExit(f"bench.py: sharded scoring disagrees with the unsharded pass / the oracle: {sharded_parity}")
    sampler = ClockSampler(local_rank)
    sampler.start()
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(args.steps)]
    barrier()
    for i in range(args.steps):
        flush.zero_()
        step_device(evs[i])
    barrier()
    total_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
    planner.pass_status()
    # the step again with CUDA events around the scoring launch (both kinds share it), then with the two kinds one
    # after the other and events around each scorer kernel: a kernel alone on the GPU is the roofline's denominator
    planner.kernel_timing(True)
    for i in range(min(args.steps, 20)):
        flush.zero_()
        step_device(None)
    barrier()
    (mig_ms, mig_n), (mps_ms, mps_n) = planner.kernel_time_ms(0), planner.kernel_time_ms(1)
    for i in range(min(args.steps, 20)):
        flush.zero_()
        step_device(None, PASS | pkg.nosplan.PASS_SERIAL)
    barrier()
    (mig_alone_ms, mig_alone_n), (mps_alone_ms, mps_alone_n) = planner.kernel_time_ms(0), planner.kernel_time_ms(1)
    planner.kernel_timing(False)
    # executed work: one pass of the counting instantiations, outside every timed region
    step_device(None, PASS | pkg.nosplan.PASS_COUNT)
    barrier()
    ex_m, ex_p = planner.executed_evals()
    tt = torch.tensor([total_ms, (mig_ms + mps_ms) / max(mig_n, mps_n, 1), 0.0, mig_alone_ms / max(mig_alone_n, 1),
                       mps_alone_ms / max(mps_alone_n, 1)], dtype=torch.float64, device=dev)
    evt = torch.tensor([float(ev_m_local), float(ev_p_local), float(ex_m), float(ex_p)], dtype=torch.float64, device=dev)
    per_rank = None
    if world > 1:
        allt = [torch.zeros_like(tt) for _ in range(world)]
        dist.all_gather(allt, tt)
        per_rank = {"step_ms": [float(x[0]) / args.steps for x in allt], "scoring_launch_ms": [float(x[1]) for x in allt],
                    "mig_alone_ms": [float(x[3]) for x in allt], "mps_alone_ms": [float(x[4]) for x in allt]}
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(evt, op=dist.ReduceOp.SUM)
    total_ms, both_k_ms, _, mig_k_alone_ms, mps_k_alone_ms = (float(x) for x in tt.tolist())
    ev_m, ev_p, ex_m, ex_p = (float(x) for x in evt.tolist())
    ms_per_step = total_ms / args.steps
    value = (ev_m + ev_p) / (ms_per_step * 1e-3)

    # ---- end to end through the C ABI with host buffers (snapshot + pods in, winners out) -----------------
    # the step's inputs live in pinned host memory (numpy views of pinned torch tensors)
    pin = lambda x: torch.from_numpy(np.ascontiguousarray(x)).pin_memory().numpy()
    if mig is not None:
        mig = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in mig.items()}
    if mps is not None:
        mps = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in mps.items()}
    mo, mk, mq, po, pk, pq = (pin(x) for x in (mo, mk, mq, po, pk, pq))

    # dtypes the ABI takes (score_snapshot passes the arrays as they are)
    _dt = dict(gpu_off=np.int32, valid=np.uint8, model=np.int32, used=np.int32, free=np.int32, mem_gb=np.int32,
               free_mask=np.uint32, used_mask=np.uint32)
    for snap in (mig, mps):
        if snap is not None:
            for k, v in list(snap.items()):
                if isinstance(v, np.ndarray) and v.dtype != _dt.get(k, v.dtype):
                    snap[k] = pin(v.astype(_dt[k]))
    mig_rows = (mo, mk, mq) if mig is not None else None
    mps_rows = (po, pk, pq) if mps is not None else None

    peer = exchange.startswith("peer") or world == 1

    def step_e2e():
        # ONE C-ABI call per pass: both snapshots and all rows in from host memory, both kinds scored, the shards'
        # winners exchanged on the GPUs (N > 1), global winners + the winners' per-GPU choices back in host memory
        got = planner.score_snapshot(mig, mig_rows, cfg["sizes"], mps, mps_rows)
        keys = np.concatenate([got[k] for k in ("mig_first_node", "mps_first_node") if k in got]) if Dall else np.zeros(0, np.int32)
        if not peer and Dall:  # no peer mapping on this box: the shards' winners go through NCCL instead
            kin = torch.from_numpy(keys).to(dev)
            keys = pkg.dist.merge_first_node_keys(kin, torch.zeros_like(kin)).cpu().numpy()
        return keys
    if not peer:
        planner.set_node_base(mig_lo, mps_lo)
    h2d = d2h = 0
    if mig is not None:
        h2d += mig["used"].nbytes + mig["free"].nbytes + mig["model"].nbytes + mig["gpu_off"].nbytes + mig["valid"].nbytes
        h2d += mo.nbytes + mk.nbytes + mq.nbytes
        d2h += Dm * 4 + Dm * planner.mig["gs"]
    if mps is not None:
        h2d += mps["used"].nbytes + mps["free"].nbytes + mps["mem_gb"].nbytes + mps["gpu_off"].nbytes + mps["valid"].nbytes + mps["free_mask"].nbytes
        h2d += po.nbytes + pk.nbytes + pq.nbytes
        d2h += Dp * 4
    if not peer:
        h2d += Dall * 4
        d2h += Dall * 4
    for _ in range(args.warmup):
        last = step_e2e()
    barrier()
    if world > 1 and sharded_parity is not None:
        # the end-to-end path (host buffers in, exchanged winners back on the host) against the same unsharded winners
        e2e_same = torch.tensor([int(np.array_equal(last, full.cpu().numpy()[:Dall]))], device=dev)
        dist.all_reduce(e2e_same, op=dist.ReduceOp.MIN)
        sharded_parity["e2e_vs_unsharded_gpu"] = "ok" if int(e2e_same[0]) else "MISMATCH"
        if not int(e2e_same[0]):
            raise SystemExit("bench.py: the end-to-end sharded pass disagrees with the unsharded winners")
    elif world == 1 and Dall:
        if not np.array_equal(last, first_glob.cpu().numpy()[:Dall]):
            raise SystemExit("bench.py: the end-to-end pass disagrees with the device-resident pass")
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / args.steps
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_val = (ev_m + ev_p) / float(e2e_t.item())

    # ---- planner latency: Plan(snapshot, pods) through the C ABI, p50 over steps ---------------------------
    # The plan is one dependency chain over the pod order: it is not sharded.  With
    # N > 1 rank 0 plans the whole snapshot on its GPU (the state is a few MB) while the others wait.
    pods = cfg["pods"]
    lat, lat_off, lat_absent = [], [], []
    absent = None
    if rank == 0 and not args.no_plan:
        pl = planner
        if world > 1:
            pl = pkg.Planner(local_rank)
            pl.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        if world > 1 or args.shard_of > 1:
            if cfg["mig"] is not None:
                pl.set_mig_snapshot(cfg["mig"])
            if cfg["mps"] is not None:
                pl.set_mps_snapshot(cfg["sizes"], cfg["mps"])
        plan_args = (pods["kind"], pods["off"], pods["key"], pods["qty"])
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pl.plan(*plan_args, gate=True, sched=True)
            if i >= args.warmup:
                lat.append((time.perf_counter() - t0) * 1e3)
        for i in range(1 + min(args.steps, 3)):
            t0 = time.perf_counter()
            pl.plan(*plan_args, gate=False, sched=True)
            if i >= 1:
                lat_off.append((time.perf_counter() - t0) * 1e3)
        # reference behaviour (gate on) where the gate does NOT stop everything: one requested MIG profile is
        # taken out of the snapshot, so its pods fall through to tryRepartition until a commit brings it back
        if cfg["mig"] is not None:
            mk_all = pods["key"][pods["off"][:-1]][pods["kind"] == 0]
            if mk_all.size:
                col = int(np.bincount(mk_all).argmax())
                snap = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
                snap["used"][:, col] = 0
                snap["free"][:, col] = 0
                pl.set_mig_snapshot(snap)
                absent = cfg["names"][col]
                for i in range(1 + min(args.steps, 5)):
                    t0 = time.perf_counter()
                    r_abs = pl.plan(*plan_args, gate=True, sched=True)
                    if i >= 1:
                        lat_absent.append((time.perf_counter() - t0) * 1e3)
                absent = {"profile": absent, "repartitions": int((r_abs["repart_node"] >= 0).sum()),
                          "gate_hits": int(r_abs["gate_hit"].sum())}
        if world > 1:
            pl.close()
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import ncu_counters
        peak, peak_src = hbm_peak()
        clocks = sampler.summary()
        props = torch.cuda.get_device_properties(local_rank)
        sm_count = props.multi_processor_count
        sm_hz = (clocks["sm_mhz"] or clocks["sm_max_mhz"] or 1965) * 1e6
        issue_peak = sm_count * 4 * sm_hz  # warp-instructions per second: 4 schedulers per SM, one issue per clock each
        nodes_tot = (int(cfg["mig"]["n_nodes"]) if cfg["mig"] else 0) + (int(cfg["mps"]["n_nodes"]) if cfg["mps"] else 0)
        shard_world = world if world > 1 else max(args.shard_of, 1)
        ctr = ncu_counters.lookup(args.config, nodes_tot, int(len(pods["kind"])), shard_world)  # rank 0's shard
        src = f"profiles/kernel_counters.json@csrc {ncu_counters.csrc_sha()}"

        def roof(kernel, alone_ms):
            """issue-slot roofline of one kernel: ncu's executed warp-instructions of one launch / the launch's duration
            measured here with CUDA events (the kernel alone on the GPU), against SMs x 4 schedulers x SM clock"""
            r = {"bound": "issue", "kernel": kernel, "unit": "warp-inst/s", "peak": issue_peak, "peak_source":
                 f"{sm_count} SMs x 4 schedulers x {sm_hz / 1e6:.0f} MHz (median SM clock sampled during the timed region)",
                 "kernel_ms": alone_ms, "achieved": None, "frac": None, "inst_per_launch": None,
                 "traffic": None, "source": None}
            k = (ctr or {}).get("kernels", {}).get(kernel)
            if k and alone_ms > 0:
                r["inst_per_launch"] = k["inst_executed"]
                r["achieved"] = k["inst_executed"] / (alone_ms * 1e-3)
                r["frac"] = r["achieved"] / issue_peak
                r["traffic"] = k.get("dram_bytes_read", 0) + k.get("dram_bytes_write", 0)
                r["source"] = src + f" ({ctr['capture']})"
                r["hbm"] = {"dram_bytes_per_launch": r["traffic"], "achieved_gbs": r["traffic"] / (alone_ms * 1e-3) / 1e9,
                            "peak_gbs": peak, "peak_source": peak_src,
                            "frac_of_measured_peak": r["traffic"] / (alone_ms * 1e-3) / 1e9 / peak}
            else:
                r["source"] = "no ncu capture of this workload on the kernels in the tree (tools/ncu_counters.py)"
            return r
        roofline = roof("mig_score16_kernel", mig_k_alone_ms)
        roofline["mps"] = roof("mps_score1_kernel", mps_k_alone_ms)
        roofline["scoring_launch_ms_in_step"] = both_k_ms  # score_both_kernel: both scorers' CTAs in one launch
        if Dm and Dp:
            roofline["step_launch"] = roof("score_both_kernel", both_k_ms)  # the step's one scoring launch, timed inside the step
        # SURVEY 8(d)'s byte model (64 B state + 4 B result per (row, node, GPU), no credit for on-chip reuse) as a work
        # rate: the tiles are read once per CTA and re-used from shared memory, so this is NOT HBM traffic
        if Dm and mig_k_alone_ms > 0:
            roofline["algorithmic_gbs_no_reuse"] = (ev_m / world) * BYTES_PER_EVAL_MIG / (mig_k_alone_ms * 1e-3) / 1e9
            roofline["algorithmic_bytes_per_eval"] = BYTES_PER_EVAL_MIG
        n_launch = (1 if Dall else 0) + (1 if (Dm and Dp) else (1 if Dall else 0)) + (1 if (world > 1 and exchange.startswith("peer")) else 0)
        line = {
            "metric": "candidate_evaluations_per_second", "value": value, "unit": "evals/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak" if args.weak else "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
            "config": workload_config(cfg, args),
            "e2e": {"value": e2e_val, "unit": "evals/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": float(e2e_t.item()) * 1e3},
            "gpu_launches": int(args.steps * n_launch),
            "launches_per_step": {"rows_group2_kernel": 1 if Dall else 0,
                                  "score_both_kernel (mig_score16 + mps_score1 bodies)": 1 if (Dm and Dp) else 0,
                                  "mig_score16_kernel": 1 if (Dm and not Dp) else 0, "mps_score1_kernel": 1 if (Dp and not Dm) else 0,
                                  "xchg_finish_kernel": 1 if (world > 1 and exchange.startswith("peer")) else 0},
            "exchange": exchange,
            "sharded_parity": sharded_parity,
            "per_rank": per_rank,
            "planner_latency_ms_p50": float(np.median(lat)) if lat else None,
            "planner_latency_ms_p95": float(np.percentile(lat, 95)) if lat else None,
            "planner_latency_gate_off_ms_p50": float(np.median(lat_off)) if lat_off else None,
            "planner_latency_absent_profile_ms_p50": float(np.median(lat_absent)) if lat_absent else None,
            "planner_absent_profile_case": absent,
            "evals_per_step": {"mig": ev_m, "mps": ev_p},
            "evals_counted": "nominal: every (row, node, GPU, allowed geometry) tuple the reference loop gpu.go:168-188 would visit "
                             "(MPS: every (row, node, GPU) run of slicing/gpu.go:162-220), identical rows included",
            # what the kernels really execute: a node leaves its GPU loop once the request is covered (node.go:164-169 makes
            # every later candidate provide 0), a GPU that already holds enough skips its candidates, and the
            # demand-independent half of an evaluation (CanApplyGeometry, distance, total, id rank) runs once per
            # snapshot (inside e2e, outside value)
            "evals_executed": {"mig": ex_m, "mps": ex_p,
                               "definition": "MIG: (row, node, GPU, candidate) tuples whose provided-count was computed by the DPX core; "
                                             "MPS: (row, node, GPU) runs of the closed form; counted by an instrumented "
                                             "instantiation of the same kernels in one extra pass"},
            "value_executed": (ex_m + ex_p) / (ms_per_step * 1e-3),
            "roofline": roofline,
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            from oracle import oracle as orc
            cores = os.cpu_count() or 1
            o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
            sample = args.ref_sample
            Sm, Sp = min(sample, Dm), min(sample, Dp)
            t0 = time.perf_counter()
            evs_cpu = 0
            if Sm:
                req, mask = pkg.synth.csr_to_dense(mo[:Sm + 1], mk, mq, len(cfg["names"]))
                o.score_mig(cfg["mig"], req, mask, threads=cores, want_choice=False)
                evs_cpu += Sm * planner.mig_evals_per_demand()
            if Sp:
                req, mask = pkg.synth.csr_to_dense(po[:Sp + 1], pk, pq, len(cfg["sizes"]))
                o.score_mps(cfg["mps"], req, mask, threads=cores, want_free=False)
                evs_cpu += Sp * planner.mps_evals_per_demand()
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": evs_cpu / dt, "unit": "evals/s", "cores": cores, "kind": "port",
                                    "sample": f"first {Sm} MIG + {Sp} MPS pods x all nodes, one pass ({dt:.2f} s); "
                                              "C++ port of the Go functions (Go toolchain absent)"}
        print(json.dumps(line), flush=True)
    if world > 1 and exchange.startswith("peer"):
        pkg.dist.close_exchange(planner)  # status, barrier, then free: no peer may still be writing flags
    planner.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
