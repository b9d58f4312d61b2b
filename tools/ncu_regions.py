"""Where the scoring launch's warps wait, by region of the kernel: aggregates `ncu -i X.ncu-rep --page source --csv
--print-source cuda,sass` (a capture made with --import-source on, tools/prof_source.sh) over the source lines of
nosplan_score16.cuh that make up the prologue, scoring loop and epilogue of the MIG and MPS bodies.
usage: python tools/ncu_regions.py <label>=<cuda_sass.csv> ..."""
import collections
import csv
import sys

REGIONS = [("MIG prologue (work item, rows, candidate rows, tile load, barrier)", 379, 418), ("MIG scoring loop", 419, 498),
           ("MIG epilogue (barrier, row winners out, push)", 499, 509), ("MPS prologue (rows, tile load, barrier)", 585, 631),
           ("MPS scoring loop", 632, 684), ("MPS epilogue", 685, 695)]


DIV_SMALL_LINE = 866  # nosplan_kernels.cuh


def num(x):
    try:
        return int(x)
    except ValueError:
        return 0


def load(path):
    out, cur, hdr = [], None, None
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif r[0] not in ("", "Function Name") and hdr:
            out.append((cur, int(r[0]), dict(zip(hdr, r))))
    return out, hdr


for arg in sys.argv[1:]:
    label, path = arg.split("=", 1)
    rows, hdr = load(path)
    tot_s = sum(num(d["# Samples"]) for _, _, d in rows)
    tot_i = sum(num(d["Instructions Executed"]) for _, _, d in rows)
    print(f"\n### {label}: {tot_s} stall samples, {tot_i:.4g} warp instructions\n")
    bar = sum(num(d.get("stall_barrier", 0)) for _, _, d in rows)
    lsb = sum(num(d.get("stall_long_sb", 0)) for _, _, d in rows)
    print(f"Of all samples: waiting at a CTA barrier {100 * bar / tot_s:.1f} %, waiting for global memory (long scoreboard) {100 * lsb / tot_s:.1f} %.\n")
    print("| region | samples | instructions | samples per 1000 instructions | top stall reasons |\n|---|---|---|---|---|")
    agg = collections.OrderedDict((r[0], [0, 0, collections.Counter()]) for r in REGIONS)
    other = [0, 0, collections.Counter()]
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for f, ln, d in rows:
        tgt = other
        if f == "nosplan_score16.cuh":
            for name, a, b in REGIONS:
                if a <= ln <= b:
                    tgt = agg[name]
        elif f == "device_functions.hpp":  # the DPX wrappers (__viaddmin_s16x2_relu, __vimax3_s16x2): only the MIG loop uses them
            tgt = agg["MIG scoring loop"]
        elif f == "nosplan_kernels.cuh" and ln == DIV_SMALL_LINE:  # div_small: only the MPS closed form uses it here
            tgt = agg["MPS scoring loop"]
        tgt[0] += num(d["# Samples"])
        tgt[1] += num(d["Instructions Executed"])
        for c in stall_cols:
            tgt[2][c] += num(d[c])
    for name, v in list(agg.items()) + [("other (kernel entry, byte_of, min / max and atomicMin wrappers, push_row / push_finish)", other)]:
        top = ", ".join(f"{k[6:]} {100 * c / max(v[0], 1):.0f} %" for k, c in v[2].most_common(4))
        print(f"| {name} | {100 * v[0] / tot_s:.1f} % | {100 * v[1] / tot_i:.1f} % | {1000 * v[0] / max(v[1], 1):.2f} | {top} |")
