"""Per-launch hardware counters of the two dominant kernels, kept next to the source they were measured on.

  python tools/ncu_counters.py add <capture.ncu-rep> --config 5 --nodes 4096 --pods 50000 --world 1
      reads the kernels of one `ncu --set full` capture of bench.py for that workload (rank 0's shard when
      world > 1) and records instructions, DRAM bytes and duration per launch in profiles/kernel_counters.json,
      stamped with the hash of csrc/ at that moment.
  csrc_sha()  the hash bench.py and tests/test_profiles_fresh.py compare against: counters measured on other
              kernels than the ones in the tree are not quoted."""
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATH = os.path.join(ROOT, "profiles", "kernel_counters.json")
KERNELS = ("score_both_kernel", "mig_score16_kernel", "mps_score1_kernel", "mig_stream_kernel")
METRICS = {"smsp__inst_executed.sum": "inst_executed", "dram__bytes_read.sum": "dram_bytes_read",
           "dram__bytes_write.sum": "dram_bytes_write", "gpu__time_duration.sum": "duration_ns",
           "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
           "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
           "launch__grid_size": "grid", "launch__registers_per_thread": "registers"}
SCALE = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1, "us": 1e3, "ns": 1, "ms": 1e6, "msecond": 1e6, "usecond": 1e3,
         "nsecond": 1, "second": 1e9}


def csrc_sha() -> str:
    d = os.path.join(ROOT, "walkai-nos_b200", "csrc")
    h = hashlib.sha1()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh")):
            h.update(f.encode())
            h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()[:16]


def load():
    if not os.path.exists(PATH):
        return {"csrc_sha": None, "entries": []}
    return json.load(open(PATH))


def lookup(config, nodes, pods, world):
    """Counters of (workload, shard of rank 0) if they were measured on the kernels now in the tree."""
    db = load()
    if db.get("csrc_sha") != csrc_sha():
        return None
    for e in db["entries"]:
        if (e["config"], e["nodes"], e["pods"], e["world"]) == (config, nodes, pods, world):
            return e
    return None


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd", choices=["add", "sha"])
    ap.add_argument("rep", nargs="?")
    ap.add_argument("--config", type=int, default=5)
    ap.add_argument("--nodes", type=int, required=False)
    ap.add_argument("--pods", type=int, required=False)
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--note", default="")
    a = ap.parse_args()
    if a.cmd == "sha":
        print(csrc_sha())
        return
    if a.rep.endswith(".csv"):  # `ncu -i x.ncu-rep --page raw --csv` made where the capture ran
        raw = open(a.rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    H, units = rr[0], rr[1]
    ki = H.index("Kernel Name")
    entry = {"config": a.config, "nodes": a.nodes, "pods": a.pods, "world": a.world, "rank": 0, "capture": os.path.basename(a.rep),
             "note": a.note, "kernels": {}}
    for r in rr[2:]:
        name = next((k for k in KERNELS if k in r[ki]), None)
        if name is None or name in entry["kernels"]:
            continue
        rec = {"name": r[ki]}
        for m, short in METRICS.items():
            if m in H:
                v = float(r[H.index(m)].replace(",", ""))
                rec[short] = v * SCALE.get(units[H.index(m)], 1)
        entry["kernels"][name] = rec
    db = load()
    sha = csrc_sha()
    if db.get("csrc_sha") != sha:
        db = {"csrc_sha": sha, "entries": []}  # counters of older kernels are not kept beside newer ones
    db["entries"] = [e for e in db["entries"] if (e["config"], e["nodes"], e["pods"], e["world"]) !=
                     (a.config, a.nodes, a.pods, a.world)] + [entry]
    json.dump(db, open(PATH, "w"), indent=1)
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()
