#!/usr/bin/env python
"""Summarise an `ncu --set full` report (+ optional launch list) into profiles/<tag>_ncu_summary.md
and profiles/traffic.json.   usage: ncu_summary.py <tag> [prof.ncu-rep] [launches.csv]"""
import collections, csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
rep = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "prof.ncu-rep")
launches = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "launches.csv")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H, U = rows[0], rows[1]
col = lambda n: H.index(n)
MUL = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
M = [("gpu__time_duration.sum", "time us"), ("launch__registers_per_thread", "regs"),
     ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
     ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
     ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu pipe %"),
     ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
     ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
     ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
     ("l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum", "RED sectors"),
     ("l1tex__t_sectors_pipe_lsu_mem_global_op_atom.sum", "ATOM sectors"),
     ("smsp__inst_executed.sum", "warp inst"),
     ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb"),
     ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall lg_throttle"),
     ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
     ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall not_selected")]
seen, traffic, issue, lines = set(), {}, {}, []
lines.append(f"# {tag}: ncu --set full --clock-control none (one launch per kernel, cold-ish cache, serialised)\n")
lines.append("| kernel | " + " | ".join(n for _, n in M) + " | dram MB (r+w) |")
lines.append("|---|" + "---|" * (len(M) + 1))
for r in rows[2:]:
    k = r[col("Kernel Name")].split("(")[0].replace("wgs::", "")
    if k in seen:
        continue
    seen.add(k)
    b = sum(float(r[col(n)]) * MUL[U[col(n)]] for n in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    traffic[k.replace("_kernel", "")] = int(b)
    try:
        issue[k.replace("_kernel", "")] = round(float(r[col("smsp__issue_active.avg.pct_of_peak_sustained_active")]), 2)
    except Exception:
        pass
    vals = []
    for n, _ in M:
        try:
            vals.append(f"{float(r[col(n)]):.4g}")
        except Exception:
            vals.append("-")
    lines.append(f"| {k} | " + " | ".join(vals) + f" | {b / 1e6:.2f} |")
if "tile_sort_small" in traffic:
    traffic["tile_sort"] = traffic["tile_sort_small"]
if os.path.exists(launches):
    lr = [r for r in csv.reader(open(launches)) if len(r) > 5]
    h = [i for i, r in enumerate(lr) if r[0] == "ID"][0]
    LH, lr = lr[h], lr[h + 1:]
    agg = collections.OrderedDict()
    for r in lr:
        a = agg.setdefault(r[LH.index("Kernel Name")][:60], [0, 0.0])
        a[0] += 1
        a[1] += float(r[LH.index("Metric Value")].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    lines.append(f"\n## launch list (gpu__time_duration.sum, {os.path.basename(launches)}): share of device time\n")
    lines.append("| kernel | launches | avg us | share % |\n|---|---|---|---|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append(f"| {k} | {a[0]} | {a[1] / a[0] / 1e3:.1f} | {100 * a[1] / tot:.1f} |")
traffic["_issue_active_pct"] = issue
traffic["_source"] = f"{tag}: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch (see profiles/{tag}_ncu_summary.md)"
os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.md"), "w").write("\n".join(lines) + "\n")
json.dump(traffic, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print("\n".join(lines))
