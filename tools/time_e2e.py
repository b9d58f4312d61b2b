"""Where the end-to-end scoring step spends its time: each C-ABI call of bench.py's step_e2e timed alone
(host wall clock; every call ends with its own stream synchronisation), inputs in pinned host memory."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import walkai_nos_b200 as pkg

cfg = pkg.synth.make_config(int(sys.argv[1]) if len(sys.argv) > 1 else 5)
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
p = pkg.Planner(0)
p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
pin = lambda x: torch.from_numpy(np.ascontiguousarray(x)).pin_memory().numpy()
mig = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in cfg["mig"].items()}
mps = {k: (pin(v) if isinstance(v, np.ndarray) else v) for k, v in cfg["mps"].items()}
mo, mk, mq, _ = pkg.synth.pods_of_kind(cfg["pods"], 0)
po, pk, pq, _ = pkg.synth.pods_of_kind(cfg["pods"], 1)
mo, mk, mq, po, pk, pq = (pin(x) for x in (mo, mk, mq, po, pk, pq))
calls = [
    ("set_mig_snapshot", lambda: p.set_mig_snapshot(mig)),
    ("set_mps_snapshot", lambda: p.set_mps_snapshot(cfg["sizes"], mps)),
    ("score_mig", lambda: p.score_mig(mo, mk, mq, want_updated=False, want_choice=False, want_first_choice=True)),
    ("score_mps", lambda: p.score_mps(po, pk, pq, want_updated=False, want_free=False)),
]
acc = {n: [] for n, _ in calls}
tot = []
for r in range(reps + 3):
    t00 = time.perf_counter()
    for n, f in calls:
        t0 = time.perf_counter()
        f()
        if r >= 3:
            acc[n].append((time.perf_counter() - t0) * 1e3)
    if r >= 3:
        tot.append((time.perf_counter() - t00) * 1e3)
for n, _ in calls:
    print(f"{n:18s} p50 {np.median(acc[n]):.3f} ms  min {min(acc[n]):.3f}")
print(f"{'step':18s} p50 {np.median(tot):.3f} ms  min {min(tot):.3f}")

# the one-call form (nosplan_score_snapshot), with NOSPLAN_TRACE=1 its own phase breakdown goes to stderr
c32 = lambda *xs: tuple(np.ascontiguousarray(x, np.int32) for x in xs)
one = []
for r in range(reps + 3):
    t0 = time.perf_counter()
    p.score_snapshot(mig, (mo, mk, mq), cfg["sizes"], mps, (po, pk, pq))
    if r >= 3:
        one.append((time.perf_counter() - t0) * 1e3)
print(f"{'score_snapshot':18s} p50 {np.median(one):.3f} ms  min {min(one):.3f}")
