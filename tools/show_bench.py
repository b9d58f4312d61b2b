"""Prints the timing fields of bench.py JSON lines: python tools/show_bench.py file.json ..."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(f) if l.startswith("{")][-1])
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    r = d.get("roofline", {})
    print(f"{f}: n={d['n_gpus']} value={d['value']:.4g} ms={d['ms_per_step']:.4f} launch_in_step={r.get('scoring_launch_ms_in_step', 0):.4f} "
          f"mig_alone={r.get('kernel_ms', 0):.4f} mps_alone={r.get('mps', {}).get('kernel_ms', 0):.4f} frac={r.get('frac')} "
          f"e2e_ms={d['e2e'].get('ms_per_step', 0):.4f} e2e={d['e2e']['value']:.4g} plan_p50={d.get('planner_latency_ms_p50')} "
          f"gate_off={d.get('planner_latency_gate_off_ms_p50')} absent={d.get('planner_latency_absent_profile_ms_p50')} parity={d.get('sharded_parity')}")
