import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import walkai_nos_b200 as pkg
cfg = pkg.synth.make_config(5)
p = pkg.Planner(0)
p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
p.set_mig_snapshot(cfg["mig"]); p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
pods = cfg["pods"]
for _ in range(4):
    t0 = time.perf_counter(); p.plan(pods["kind"], pods["off"], pods["key"], pods["qty"]); print("ms", (time.perf_counter()-t0)*1e3)
