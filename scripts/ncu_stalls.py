#!/usr/bin/env python
"""Warp-state breakdown of the kernels in an ncu --set full report: average warps per issue slot by stall reason
(smsp__average_warps_issue_stalled_*_per_issue_active), sorted.    usage: ncu_stalls.py prof.ncu-rep [kernel regex]"""
import csv, io, re, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
H = rows[0]
rx = re.compile(sys.argv[2] if len(sys.argv) > 2 else ".")
stall = [(i, h) for i, h in enumerate(H) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")
         and "not_issued" not in h]
extra = ["gpu__time_duration.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
         "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
         "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
         "sm__cycles_active.avg", "sm__cycles_active.max", "smsp__warps_eligible.avg.per_cycle_active",
         "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum"]
seen = set()
for r in rows[2:]:
    k = r[H.index("Kernel Name")].split("(")[0]
    if not rx.search(k) or k in seen:
        continue
    seen.add(k)
    print("==", k)
    for e in extra:
        if e in H:
            print(f"   {e}: {r[H.index(e)]}")
    vals = sorted(((float(r[i] or 0), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]) for i, h in stall), reverse=True)
    print("   stalls (warps per issue): " + ", ".join(f"{n} {v:.2f}" for v, n in vals if v >= 0.02))
