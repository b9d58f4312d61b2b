import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import walkai_nos_b200 as pkg
from oracle import oracle as orc
cfg = pkg.synth.make_config(3, 0, 3000)
p = pkg.Planner(0)
p.set_mig_tables(cfg["names"], cfg["geom_off"], cfg["geom_counts"])
p.set_mps_snapshot(cfg["sizes"], cfg["mps"])
o = orc.Oracle(cfg["tables"], mps_sizes=cfg["sizes"])
for kk in (0, 3):
    kind = np.ones(1, np.uint8); off = np.asarray([0, 1], np.int32); key = np.asarray([kk], np.int32); qty = np.asarray([1], np.int32)
    req, mask = pkg.synth.csr_to_dense(off, key, qty, 13)
    exp, ms, ps = o.plan(None, cfg["mps"], kind, req, mask, gate=False, sched=False)
    got = p.plan(kind, off, key, qty, gate=False, sched=False)
    print("key", kk, "exp", exp["repart_node"], "got", got["repart_node"])
    used, free, fm = p.get_mps_state()
    r = p.score_mps(off, key, qty)
    for g in range(8):
        print("  gpu", g, "used", used[g], "orig", cfg["mps"]["free"][g], "got", free[g], hex(fm[g]), "exp", ps["free"][g], hex(ps["free_mask"][g]), "score", r["out_free"][0, g], hex(r["out_free_mask"][0, g]))
