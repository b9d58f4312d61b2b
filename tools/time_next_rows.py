"""Measurement of the two "next" rows of SURVEY 8(f) at cfg5 scale.
f2  nosplan_diff_plan_mig (agent-side plan.NewMigConfigPlan at count level): C-ABI wall time and kernel-only
    device time for 32 768 GPUs x 13 profiles, next to the count-level oracle on one host core.
f1  libnoshost (annotation wire format -> SoA snapshot, pod specs -> request rows, state -> spec annotations):
    host code by design; wall time for 4096 nodes x 8 GPUs of status annotations.
usage: python tools/time_next_rows.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import walkai_nos_b200 as pkg
from oracle import oracle as orc


def med(f, n=20, warm=3):
    for _ in range(warm):
        f()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        f()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts))


def main():
    rng = np.random.default_rng(11)
    # ---- f2 ------------------------------------------------------------------------------------------------
    G, P = 32768, 13
    used = rng.integers(0, 4, size=(G, P)).astype(np.int32)
    free = rng.integers(0, 4, size=(G, P)).astype(np.int32)
    spec = rng.integers(0, 8, size=(G, P)).astype(np.int32)
    mask = rng.integers(0, 1 << P, size=G).astype(np.uint32)
    p = pkg.Planner(0)
    got = p.diff_plan_mig(used, free, spec, mask)
    exp = orc.mig_diff_counts(used, free, spec, mask)
    assert all(np.array_equal(a, b) for a, b in zip(got, exp))
    t_abi = med(lambda: p.diff_plan_mig(used, free, spec, mask))
    t_cpu = med(lambda: orc.mig_diff_counts(used, free, spec, mask), n=5, warm=1)
    print(f"f2 diff plan, {G} GPUs x {P} profiles: C ABI (5.1 MB in, 5.1 MB out) {t_abi * 1e3:.3f} ms = "
          f"{G / t_abi:.3g} GPU rows/s; oracle on one core {t_cpu * 1e3:.2f} ms = {G / t_cpu:.3g} rows/s")
    p.close()
    # ---- f1 ------------------------------------------------------------------------------------------------
    from kat_util import table
    host = pkg.host
    tables = table("yaml")
    model = "NVIDIA-H100-80GB"
    nodes = []
    n_ann = 0
    for n in range(4096):
        ann = {}
        for g in range(8):
            geom = tables[model][int(rng.integers(0, 12))]
            for pr, q in geom.items():
                u = int(rng.integers(0, q + 1))
                if u:
                    ann[f"nos.nebuly.com/status-gpu-{g}-{pr}-used"] = str(u)
                if q - u:
                    ann[f"nos.nebuly.com/status-gpu-{g}-{pr}-free"] = str(q - u)
        n_ann += len(ann)
        nodes.append(dict(name=f"node-{n:04d}", labels={"nvidia.com/gpu.product": model, "nvidia.com/gpu.count": "8"},
                          annotations=ann))
    names = sorted({pr for geoms in tables.values() for g in geoms for pr in g})
    t0 = time.perf_counter()
    snap = host.Snapshot("mig", nodes, tables.keys(), extra_names=names)
    a = snap.arrays()
    t_snap = time.perf_counter() - t0
    profs = ["1g.10gb", "2g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]
    pods = [dict(name=f"ns/p{i:05d}", priority=int(rng.choice([0, 100])), created=i,
                 containers=[{f"nvidia.com/mig-{profs[int(rng.integers(0, 5))]}": int(rng.integers(1, 3))}])
            for i in range(50000)]
    t0 = time.perf_counter()
    rows = host.build_pods(pods, snap, None)
    t_pods = time.perf_counter() - t0
    print(f"f1 wire format -> SoA: {len(nodes)} nodes, {n_ann} status annotations in {t_snap * 1e3:.1f} ms "
          f"({n_ann / t_snap:.3g} annotations/s, Python marshalling included); {len(pods)} pod specs -> "
          f"{len(rows['off']) - 1} request rows in {t_pods * 1e3:.1f} ms")
    assert int(a["gpu_off"][-1]) == 4096 * 8
    more_rows(rng)


def more_rows(rng):
    """Round-2 rows: the device-level diff planner, the table loader, the batch-demand mode and the allocation-free scorer."""
    import torch
    from oracle import batch_plan
    p = pkg.Planner(0)
    # f2 at device level: 32 768 GPUs, ~6 devices each
    G, P = 32768, 13
    n_dev = rng.integers(0, 8, size=G)
    off = np.concatenate([[0], np.cumsum(n_dev)]).astype(np.int32)
    prof = rng.integers(0, P, size=int(off[-1])).astype(np.int32)
    status = rng.integers(0, 3, size=int(off[-1])).astype(np.uint8)
    spec = rng.integers(0, 4, size=(G, P)).astype(np.int32)
    mask = rng.integers(0, 1 << P, size=G).astype(np.uint32)
    t = med(lambda: p.diff_plan_mig_devices(off, prof, status, spec, mask))
    print(f"f2 device level, {G} GPUs / {int(off[-1])} devices: C ABI {t * 1e3:.3f} ms = {G / t:.3g} GPU rows/s")
    # a16 loader: the deployed table
    text = pkg.synth.table_yaml_text()
    t = med(lambda: pkg.host.load_tables(text), n=50)
    print(f"a16 loader: known_mig_geometries.yaml ({len(text)} bytes, 3 entries, 8 models) -> flat tables in {t * 1e6:.0f} us (Python marshalling included)")
    # f4 batch-demand mode: parity + CPU time of the definition on a cfg5 shard, GPU time at the full cfg5 size
    def batch_once(cfg, kind, kd, key, qty):
        (p.set_mig_snapshot if kd == 0 else lambda s: p.set_mps_snapshot(cfg["sizes"], s))(cfg[kind])
        t0 = time.perf_counter()
        r = p.plan_batch(kd, key, qty)
        return r, time.perf_counter() - t0
    for nodes, pods, check in ((256, 3000, True), (0, 0, False)):
        cfg = pkg.synth.make_config(5, nodes, pods)
        p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
        o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
        for kind, kd in (("mig", 0), ("mps", 1)):
            off_, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], kd)
            M = len(off_) - 1
            batch_once(cfg, kind, kd, key[:M], qty[:M])
            r, t = min((batch_once(cfg, kind, kd, key[:M], qty[:M]) for _ in range(5)), key=lambda x: x[1])
            extra = ""
            if check:
                t0 = time.perf_counter()
                e = batch_plan.batch_plan(o, kind, cfg[kind], key[:M], qty[:M])
                extra = f"; CPU definition {(time.perf_counter() - t0) * 1e3:.0f} ms, same result"
                assert np.array_equal(r["sched_node"], e["sched_node"]) and np.array_equal(r["repartitioned"], e["repartitioned"])
            print(f"f4 batch mode {kind}: {M} pods x {cfg[kind]['n_nodes']} nodes: {t * 1e3:.2f} ms through the C ABI ({r['rounds']} rounds, "
                  f"{int((r['phase'] == 1).sum())} placed at once + {int((r['phase'] == 2).sum())} after re-partitioning "
                  f"{int(r['repartitioned'].sum())} nodes, {int((r['phase'] == 0).sum())} pending){extra}")
    p.set_mig_snapshot(cfg["mig"])
    # the allocation-free MIG scorer next to the ordinary device-resident one (cfg5: 2048 MIG nodes, 25 000 rows)
    mig = cfg["mig"]
    off_, key, qty, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
    D = len(off_) - 1
    dev = torch.device("cuda", 0)
    tt = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    bufs = [tt(mig[k]) for k in ("valid", "model", "used", "free")] + [tt(off_), tt(key), tt(qty)]
    first = torch.empty(D, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def dev_time(f, n=20):
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n
    ts = dev_time(lambda: p.score_mig_static(int(mig["n_nodes"]), 8, *[b.data_ptr() for b in bufs[:4]], D, *[b.data_ptr() for b in bufs[4:]],
                                             first.data_ptr(), st))
    a = first.cpu().numpy().copy()
    td = dev_time(lambda: p.score_mig_dev(D, bufs[4].data_ptr(), bufs[5].data_ptr(), bufs[6].data_ptr(), first.data_ptr(), 0, 0, st))
    assert np.array_equal(a, first.cpu().numpy())
    print(f"nosplan_score_mig_static (snapshot packed inside the call, general scorer): {ts:.3f} ms per call; nosplan_score_mig_dev "
          f"(snapshot resident, ordered scorer): {td:.3f} ms; same winners")
    p.close()


if __name__ == "__main__":
    main()
