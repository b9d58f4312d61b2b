"""Summarises ncu outputs into profiles/: launch list (per-kernel share) and key metrics of a full capture.
usage: python tools/ncu_summary.py <launches.csv> <prof.ncu-rep> <out.md>"""
import collections, csv, subprocess, sys

launches, rep, out = sys.argv[1:4]
lines = []
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
hdr = [i for i, r in enumerate(rows) if r[0] == 'ID'][0]
H = rows[hdr]; rows = rows[hdr + 1:]
ki, vi = H.index('Kernel Name'), H.index('Metric Value')
agg = collections.OrderedDict()
for r in rows:
    agg.setdefault(r[ki][:70], []).append(float(r[vi].replace(',', '')))
tot = sum(sum(v) for v in agg.values())
lines.append("## Launch list (ncu --metrics gpu__time_duration.sum --clock-control none; cold-cache, serialised: compare shares)\n")
lines.append("| kernel | launches | avg us | share |\n|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    lines.append(f"| `{k}` | {len(v)} | {sum(v)/len(v)/1e3:.1f} | {sum(v)/tot*100:.1f}% |")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
H = rr[0]; units = rr[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_bytes.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__dynamic_shared_memory_per_block',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'launch__waves_per_multiprocessor',
        'smsp__average_warp_latency_issue_stalled_short_scoreboard.pct', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio']
lines.append("\n## Full capture (ncu --set full --clock-control none), per launch\n")
for r in rr[2:]:
    lines.append(f"### `{r[H.index('Kernel Name')]}`\n")
    lines.append("| metric | value | unit |\n|---|---|---|")
    for w in want:
        if w in H:
            lines.append(f"| {w} | {r[H.index(w)]} | {units[H.index(w)]} |")
    lines.append("")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
