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


if __name__ == "__main__":
    main()
