"""Times nosplan_plan (host wall clock through the C ABI) on a BASELINE config."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import walkai_nos_b200 as pkg
cfgno = int(sys.argv[1]) if len(sys.argv) > 1 else 5
cfg = pkg.synth.make_config(cfgno)
p = pkg.Planner(0)
p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
if cfg["mig"] is not None: p.set_mig_snapshot(cfg["mig"])
if cfg["mps"] is not None: p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
pods = cfg["pods"]
for gate, sched, reps in ((True, True, 8), (True, False, 4), (False, True, 2)):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        r = p.plan(pods["kind"], pods["off"], pods["key"], pods["qty"], gate=gate, sched=sched)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"cfg{cfgno} gate={gate} sched={sched}: ms {['%.2f' % t for t in ts]} scheduled={int((r['sched_node']>=0).sum())} repart={int((r['repart_node']>=0).sum())}")
