"""Oversized streaming sweep: a cluster far larger than L2 scored against a handful of
demand rows, so every state byte really comes from HBM once per launch.  Reports the DRAM rate the MIG
scorer reaches in that regime (it is ALU-bound: see DESIGN.md).  usage: stream_sweep.py [nodes] [demands]"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import walkai_nos_b200 as pkg

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
D = int(sys.argv[2]) if len(sys.argv) > 2 else 16
G = 8
names = pkg.synth.MIG_PROFILES
tables = pkg.synth.yaml_tables()
model_names, geom_off, geom_counts = pkg.synth.flatten_tables(tables, names)
mid = model_names.index("NVIDIA-H100-80GB")
rng = np.random.default_rng(0x6E6F73)
tg = N * G
tab = geom_counts[geom_off[mid]:geom_off[mid + 1]]                 # [12][P]
counts = tab[rng.integers(0, len(tab), size=tg)]                    # every GPU: a uniformly drawn allowed geometry
used = rng.binomial(counts, 0.5).astype(np.int32)                   # every device used with probability 0.5
free = (counts - used).astype(np.int32)
snap = dict(n_nodes=N, gpu_off=np.arange(0, tg + 1, G, dtype=np.int32), valid=np.ones(N, np.uint8),
            model=np.full(tg, mid, np.int32), used=used, free=free)
p = pkg.Planner(0)
p.set_mig_tables(names, geom_off, geom_counts)
p.set_mig_snapshot(snap)
h100 = [names.index(x) for x in ["1g.10gb", "2g.20gb", "3g.20gb", "3g.40gb", "4g.40gb", "7g.79gb"]]
key = np.asarray([h100[i % 6] for i in range(D)], np.int32)
qty = np.asarray([1 + (i // 6) % 2 for i in range(D)], np.int32)
off = np.arange(D + 1, dtype=np.int32)
dev = torch.device("cuda", 0)
t = lambda a: torch.from_numpy(a).to(dev)
d_off, d_key, d_qty = t(off), t(key), t(qty)
first = torch.empty(D, dtype=torch.int32, device=dev)
st = torch.cuda.current_stream().cuda_stream
run = lambda: p.score_mig_dev(D, d_off.data_ptr(), d_key.data_ptr(), d_qty.data_ptr(), first.data_ptr(), 0, 0, st)
for _ in range(3):
    run()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
reps = 10
ev[0].record()
for _ in range(reps):
    run()
ev[1].record()
torch.cuda.synchronize()
ms = ev[0].elapsed_time(ev[1]) / reps
# what one launch has to read per node and per pass over the node list (one pass per 4 rows of a profile on the
# streaming kernel): the snapshot-time rank table (gs * 2cw words = 192 B), the profile's free|cur column (gs u16),
# the model bytes, the GPU count
passes = sum((int((key == k).sum()) + 3) // 4 for k in set(key.tolist()))
state_bytes = passes * N * (G * 6 * 4 + G * 2 + G + 1)
evals = D * p.mig_evals_per_demand()
peak = 6542.1
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
print(json.dumps({"nodes": N, "gpus": tg, "demands": D, "stream_rows_knob": os.environ.get("NOSPLAN_STREAM_ROWS", "default"), "ms": ms, "evals_per_s": evals / (ms * 1e-3),
                  "state_bytes_streamed": state_bytes, "dram_GBps": state_bytes / (ms * 1e-3) / 1e9,
                  "frac_of_hbm_peak": state_bytes / (ms * 1e-3) / 1e9 / peak,
                  "survey_8d_GBps": evals * 68 / 12 / (ms * 1e-3) / 1e9, "first_node": first.cpu().numpy()[:6].tolist()}))
