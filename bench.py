#!/usr/bin/env python
"""bench.py — planner hot-path benchmark (contract: see the task statement / DESIGN.md "Measurement").

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--impl b200|reference]

A "step" is one pass of the batched scorer over one cluster snapshot: every pending pod's request
vector is scored against every node (MIG pods x MIG nodes, MPS pods x MPS nodes).  The metric is
candidate evaluations per second (one evaluation = one iteration of pkg/gpu/mig/gpu.go:168-188 for
a (demand, node, GPU, allowed geometry) tuple; for MPS one greedy run per (demand, node, GPU)).
Planner latency (Plan(snapshot, pods) through the C ABI, p50) is reported in the same line.

N > 1: one process per GPU (torchrun), nodes sharded contiguously across ranks, pods replicated, one
NCCL all-reduce(min) of the per-demand first-updating-node keys per step ("first node wins").
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
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.rows = []
        self.stop_flag = False

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons}


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
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-plan", action="store_true", help="skip the planner-latency section (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    import walkai_nos_b200 as pkg
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
    mig, mig_lo = pkg.dist.shard_snapshot(cfg["mig"], rank, world)
    mps, mps_lo = pkg.dist.shard_snapshot(cfg["mps"], rank, world)
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
    first_all = torch.full((max(Dm, 1) + max(Dp, 1),), -1, dtype=torch.int32, device=dev)
    first_m, first_p = first_all[:max(Dm, 1)], first_all[max(Dm, 1):]
    # N > 1: the per-row "first node" keys of all shards are merged by the library's own exchange
    # (atomicMin into every peer's buffer over NVLink + step flags); NCCL is only the cross-check
    first_glob = torch.empty_like(first_all)
    exchange = "none"
    if world > 1:
        exchange = "peer atomicMin over NVLink (libnosplan)"
        try:
            pkg.dist.open_exchange(planner, first_all.numel())
        except pkg.NosplanError as e:  # no peer mapping on this box: say so, use the library collective
            exchange = f"nccl all-reduce (peer mapping failed: {e})"
        flags_t = torch.tensor([0 if exchange.startswith("peer") else 1], device=dev)
        dist.all_reduce(flags_t, op=dist.ReduceOp.MAX)
        if int(flags_t.item()) and exchange.startswith("peer"):
            exchange = "nccl all-reduce (a peer could not map)"
        if exchange.startswith("peer"):
            planner.set_node_base(mig_lo, mps_lo)  # the scorers report global node indices
    lo_vec = torch.cat([torch.full((max(Dm, 1),), mig_lo, dtype=torch.int32, device=dev),
                        torch.full((max(Dp, 1),), mps_lo, dtype=torch.int32, device=dev)])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    BIG = 0x7FFFFFFF

    def step_device(ev=None):
        """inputs resident in HBM; returns nothing. ev = (e0, e1, e2) to time mig / mps parts."""
        if ev:
            ev[0].record()
        if Dm:
            planner.score_mig_dev(Dm, d_mo.data_ptr(), d_mk.data_ptr(), d_mq.data_ptr(), first_m.data_ptr(), 0, 0, stream)
        if ev:
            ev[1].record()
        if Dp:
            planner.score_mps_dev(Dp, d_po.data_ptr(), d_pk.data_ptr(), d_pq.data_ptr(), first_p.data_ptr(), 0, 0, 0, stream)
        if world > 1:
            # first node wins across shards: global index = local + shard offset, min over ranks
            if exchange.startswith("peer"):
                planner.exchange_min_dev(first_all.numel(), first_all.data_ptr(), first_glob.data_ptr(), stream)
            else:
                first_glob.copy_(pkg.dist.merge_first_node_keys(first_all, lo_vec))
        if ev:
            ev[2].record()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush.zero_()
        step_device()
    barrier()
    if world > 1 and exchange.startswith("peer"):
        # cross-check of the library's exchange against the NCCL all-reduce on the same keys
        planner.exchange_status()
        ref = pkg.dist.merge_first_node_keys(first_all, torch.zeros_like(first_all))
        if not torch.equal(ref, first_glob):
            raise SystemExit("bench.py: the peer-memory exchange disagrees with the NCCL all-reduce")
    sharded_parity = None
    if world > 1:
        # The sharded answer must be the unsharded one: rank 0 scores the WHOLE snapshot on its GPU (one handle, no
        # exchange) and checks that against the oracle on the distinct request vectors; every rank then compares
        # what the sharded pass + exchange left in first_glob with rank 0's unsharded winners.
        full = torch.full_like(first_all, -1)
        oracle_ok = 1
        if rank == 0:
            fp = pkg.Planner(local_rank)
            fp.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
            parts = []
            for kind, snap, rows, D in ((0, cfg["mig"], (mo, mk, mq), Dm), (1, cfg["mps"], (po, pk, pq), Dp)):
                if snap is None or not D:
                    parts.append(np.full(1, -1, np.int32))
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
            full = torch.from_numpy(np.concatenate(parts)).to(dev)
        dist.broadcast(full, src=0)
        same = torch.tensor([int(torch.equal(full, first_glob)), oracle_ok], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        sharded_parity = {"vs_unsharded_gpu": "ok" if int(same[0]) else "MISMATCH",
                          "vs_oracle_dedup": "ok" if int(same[1]) else "MISMATCH", "rows": int(first_all.numel())}
        if not int(same[0]) or not int(same[1]):
            raise SystemExit(f"bench.py: sharded scoring disagrees with the unsharded pass / the oracle: {sharded_parity}")
    sampler = ClockSampler(local_rank)
    sampler.start()
    total_ms = mig_ms = mps_ms = 0.0
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    barrier()
    for i in range(args.steps):
        flush.zero_()
        step_device(evs[i])
    barrier()
    for e in evs:
        total_ms += e[0].elapsed_time(e[2])
        mig_ms += e[0].elapsed_time(e[1])
        mps_ms += e[1].elapsed_time(e[2])
    tt = torch.tensor([total_ms, mig_ms, mps_ms], dtype=torch.float64, device=dev)
    evt = torch.tensor([float(ev_m_local), float(ev_p_local)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_reduce(evt, op=dist.ReduceOp.SUM)
    total_ms, mig_ms, mps_ms = (float(x) for x in tt.tolist())
    ev_m, ev_p = (float(x) for x in evt.tolist())
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

    def step_e2e():
        # one C-ABI call per pass: both snapshots in, both kinds scored (side by side inside the library), winners out
        got = planner.score_snapshot(mig, mig_rows, cfg["sizes"], mps, mps_rows)
        out = [got]
        keys = []
        if mig is not None and Dm:
            keys.append((got["mig_first_node"], mig_lo))
        if mps is not None and Dp:
            keys.append((got["mps_first_node"], mps_lo))
        if world > 1 and keys:
            # the shards' winners (host arrays, local indices) -> global keys -> exchanged on the GPUs -> host
            k = np.concatenate([np.where(f >= 0, f + lo, -1).astype(np.int32) for f, lo in keys])
            kin = torch.from_numpy(k).to(dev)
            kout = first_glob[:kin.numel()]
            if exchange.startswith("peer"):
                planner.exchange_min_dev(kin.numel(), kin.data_ptr(), kout.data_ptr(), stream)
            else:
                kout.copy_(pkg.dist.merge_first_node_keys(kin, torch.zeros_like(kin)))
            out.append(kout.cpu())
        return out
    h2d = d2h = 0
    if mig is not None:
        h2d += mig["used"].nbytes + mig["free"].nbytes + mig["model"].nbytes + mig["gpu_off"].nbytes + mig["valid"].nbytes
        h2d += mo.nbytes + mk.nbytes + mq.nbytes
        d2h += Dm * 4 + Dm * planner.mig["gs"]
    if mps is not None:
        h2d += mps["used"].nbytes + mps["free"].nbytes + mps["mem_gb"].nbytes + mps["gpu_off"].nbytes + mps["valid"].nbytes + mps["free_mask"].nbytes
        h2d += po.nbytes + pk.nbytes + pq.nbytes
        d2h += Dp * 4
    if world > 1:
        h2d += (Dm + Dp) * 4
        d2h += (Dm + Dp) * 4
    for _ in range(args.warmup):
        last = step_e2e()
    barrier()
    if world > 1 and sharded_parity is not None:
        # the end-to-end path (host buffers in, exchanged winners back on the host) against the same unsharded winners
        want = torch.cat([full[:max(Dm, 1)][:Dm], full[max(Dm, 1):][:Dp]]).cpu()
        e2e_same = torch.tensor([int(torch.equal(last[1], want))], device=dev)
        dist.all_reduce(e2e_same, op=dist.ReduceOp.MIN)
        sharded_parity["e2e_vs_unsharded_gpu"] = "ok" if int(e2e_same[0]) else "MISMATCH"
        if not int(e2e_same[0]):
            raise SystemExit("bench.py: the end-to-end sharded pass disagrees with the unsharded winners")
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
    lat, lat_off = [], []
    if rank == 0 and not args.no_plan:
        pl = planner
        if world > 1:
            pl = pkg.Planner(local_rank)
            pl.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
            if cfg["mig"] is not None:
                pl.set_mig_snapshot(cfg["mig"])
            if cfg["mps"] is not None:
                pl.set_mps_snapshot(cfg["sizes"], cfg["mps"])
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            pl.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=True, sched=True)
            if i >= args.warmup:
                lat.append((time.perf_counter() - t0) * 1e3)
        for i in range(1 + min(args.steps, 3)):
            t0 = time.perf_counter()
            pl.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=False, sched=True)
            if i >= 1:
                lat_off.append((time.perf_counter() - t0) * 1e3)
        if world > 1:
            pl.close()
    barrier()
    sampler.stop_flag = True
    sampler.join(timeout=2)

    if rank == 0:
        peak, peak_src = hbm_peak()
        mig_kernel_s = (mig_ms / args.steps) * 1e-3
        ach = (ev_m / world) * BYTES_PER_EVAL_MIG / mig_kernel_s / 1e9 if Dm and mig_kernel_s > 0 else 0.0
        line = {
            "metric": "candidate_evaluations_per_second", "value": value, "unit": "evals/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "u8/int32", "data": "synthetic",
            "config": workload_config(cfg, args),
            "e2e": {"value": e2e_val, "unit": "evals/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": float(e2e_t.item()) * 1e3},
            "gpu_launches": int(args.steps * ((3 if Dm else 0) + (3 if Dp else 0) + (2 if exchange.startswith("peer") else 0))),
            "exchange": exchange,
            "sharded_parity": sharded_parity,
            "planner_latency_ms_p50": float(np.median(lat)) if lat else None,
            "planner_latency_ms_p95": float(np.percentile(lat, 95)) if lat else None,
            "planner_latency_gate_off_ms_p50": float(np.median(lat_off)) if lat_off else None,
            "evals_per_step": {"mig": ev_m, "mps": ev_p},
            "evals_definition": "MIG: one per (row, node, GPU, allowed geometry) = iterations of the loop body gpu.go:168-188, "
                                "identical rows included; its demand-independent half (CanApplyGeometry, distance, total, id rank) "
                                "is evaluated once per snapshot (inside e2e, outside value), the demand-dependent half per tuple in "
                                "the timed kernels.  MPS: one per (row, node, GPU) run of slicing/gpu.go:162-220",
            # achieved/frac follow SURVEY §8(d): algorithmic bytes (no credit for on-chip reuse) / time of the MIG
            # scorer's launches (row grouping, mig_score16_kernel, the multi-key pass, the finishing pass; CUDA
            # events).  traffic = ncu dram__bytes_read+write per mig_score16_kernel launch of the same command
            # (profiles/r1k_bench_cfg5.md): the tile state is re-used from shared memory, HBM is idle and the
            # binding roof is instruction issue (ncu: issue-active 73 %, ALU pipe 72 %, LSU 40 %): frac is above 1.
            "roofline": {"bound": "hbm", "kernel": "mig_score16_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "peak_source": peak_src, "traffic": 1406208 if args.config == 5 and world == 1 else None,
                         "algorithmic_bytes_per_eval": BYTES_PER_EVAL_MIG, "kernel_ms": mig_ms / args.steps,
                         "mps_kernel_ms": mps_ms / args.steps, "binding_roof": "instruction issue (integer pipes)",
                         "alu_pipe_pct_of_peak_ncu": 71.9, "issue_active_pct_ncu": 72.9},
            "clocks": sampler.summary(),
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
    planner.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
