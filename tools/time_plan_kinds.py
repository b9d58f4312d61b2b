import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import walkai_nos_b200 as pkg
cfg = pkg.synth.make_config(5)
p = pkg.Planner(0)
p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
p.set_mig_snapshot(cfg["mig"]); p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
pods = cfg["pods"]
for kind in (0, 1):
    off, key, qty, sel = pkg.synth.pods_of_kind(pods, kind)
    kd = np.full(len(sel), kind, np.uint8)
    for gate in (True, False):
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); r = p.plan(kd, off, key, qty, gate=gate, sched=True); ts.append((time.perf_counter()-t0)*1e3)
        print(f"kind={kind} gate={gate}: {['%.2f'%t for t in ts]} repart={int((r['repart_node']>=0).sum())} sched={int((r['sched_node']>=0).sum())}")
