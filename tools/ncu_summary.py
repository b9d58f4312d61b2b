"""Summarises ncu outputs into profiles/: launch list (per-kernel share) and key metrics of a full capture.
usage: python tools/ncu_summary.py <launches.csv> <prof.ncu-rep | raw-page.csv> <out.md> [W:<raw.csv>:<launches.csv> ...]
The optional W:... arguments append the same kernels on rank 0's shard of a W-rank run (tools/prof_bench.sh <tag> --shard-of W)."""
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
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
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
shards = [a.split(":") for a in sys.argv[4:]]
if shards:
    lines.append("\n## The same command on rank 0's shard of a W-rank run (`bench.py --shard-of W`, one GPU; `tools/prof_bench.sh`)\n")
    lines.append("| shard | kernel | duration (ncu) | warp-instructions | issue active | ALU pipe | DRAM read | grid | waves |\n|---|---|---|---|---|---|---|---|---|")
    tails = []
    for w, raw_csv, launches_csv in [("1", rep, launches)] + shards:
        if not raw_csv.endswith(".csv"):
            continue
        q = list(csv.reader(open(raw_csv).read().splitlines()))
        Hq, seen = q[0], set()
        col = lambda r, k: r[Hq.index(k)]
        for r in q[2:]:
            name = col(r, "Kernel Name").split("<")[0].replace("void ", "")
            if name in seen:
                continue
            seen.add(name)
            lines.append(f"| 1/{w} | `{name}` | {float(col(r, 'gpu__time_duration.sum')):.1f} us | {float(col(r, 'smsp__inst_executed.sum')):.4g} | "
                         f"{float(col(r, 'smsp__issue_active.avg.pct_of_peak_sustained_active')):.1f} % | "
                         f"{float(col(r, 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active')):.1f} % | "
                         f"{float(col(r, 'dram__bytes_read.sum')):.2f} {q[1][Hq.index('dram__bytes_read.sum')]} | {col(r, 'launch__grid_size')} | "
                         f"{col(r, 'launch__waves_per_multiprocessor')} |")
        if w != "1":
            rows2 = [r for r in csv.reader(open(launches_csv)) if len(r) > 5]
            h2 = [i for i, r in enumerate(rows2) if r[0] == 'ID'][0]
            H2 = rows2[h2]
            agg2 = collections.OrderedDict()
            for r in rows2[h2 + 1:]:
                agg2.setdefault(r[H2.index('Kernel Name')][:60], []).append(float(r[H2.index('Metric Value')].replace(',', '')))
            tot2 = sum(sum(v) for v in agg2.values())
            tails.append(f"* 1/{w}: " + "; ".join(f"`{k}` x{len(v)} avg {sum(v) / len(v) / 1e3:.1f} us ({sum(v) / tot2 * 100:.0f} %)"
                                                  for k, v in sorted(agg2.items(), key=lambda kv: -sum(kv[1]))[:6]))
    lines.append("\nLaunch lists of the shard runs (`ncu --metrics gpu__time_duration.sum`, cold-cache and serialised: compare shares):\n")
    lines += tails
MARK = "<!-- kept by tools/ncu_summary.py: everything below this line is hand-written -->"
import os
if os.path.exists(out) and MARK in open(out).read():
    lines.append("\n" + MARK + open(out).read().split(MARK, 1)[1].rstrip("\n"))
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
