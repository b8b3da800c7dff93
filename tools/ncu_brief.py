#!/usr/bin/env python
"""Brief of one `ncu --set full` capture (raw csv page): duration, DRAM bytes, pipe activity, issue, stalls.
usage: ncu_brief.py file_raw.csv "<how it was captured>" """
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h, u, v = rows[0], rows[1], rows[2]
d = {n: (v[i], u[i]) for i, n in enumerate(h)}
print(sys.argv[2] if len(sys.argv) > 2 else "")
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "gpu__time_duration.sum",
        "sm__cycles_elapsed.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed"]
for k in want:
    if k in d:
        print(f"{k:78s} {d[k][1]:16s} {d[k][0]}")
print("--- warp stall reasons (warps per issue-active cycle, >= 0.2)")
for n in h:
    if "issue_stalled" in n and n.endswith("per_issue_active.ratio"):
        try:
            x = float(d[n][0])
        except ValueError:
            continue
        if x >= 0.2:
            print(f"{n.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', ''):30s} {x:6.2f}")
