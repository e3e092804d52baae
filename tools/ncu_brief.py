#!/usr/bin/env python
"""The handful of ncu metrics the kernel notes quote: python tools/ncu_brief.py rep.ncu-rep"""
import csv, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__sass_inst_executed_op_local_ld.sum", "smsp__sass_inst_executed_op_local_st.sum",
        "launch__grid_size", "launch__block_size", "sm__warps_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:90])
    for i, n in enumerate(h):
        if n in want:
            print("  %-72s %s %s" % (n, r[i], u[i]))
    st = [(float(r[i] or 0), n) for i, n in enumerate(h) if n.startswith("smsp__average_warps_issue_stalled_") and n.endswith("_per_issue_active.ratio")]
    print("  stalls per issue:", ", ".join("%s %.2f" % (n[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], v) for v, n in sorted(st, reverse=True)[:7]))
