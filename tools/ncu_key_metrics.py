"""Summarise an `ncu --set full` report (raw page, CSV) into the handful of metrics DESIGN.md quotes.

    ncu -i report.ncu-rep --page raw --csv > raw.csv
    python tools/ncu_key_metrics.py raw.csv "header line" > profiles/<name>.txt
"""
import csv
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.per_cycle_active",
    "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
]

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
print("#", sys.argv[2] if len(sys.argv) > 2 else "")
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    print()
    print("##", r[ix["Kernel Name"]])
    for k in KEYS:
        if k in ix:
            print(f"{k:<86} {r[ix[k]]:>16} {units[ix[k]]}")
    stalls = []
    for h, i in ix.items():
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            try:
                stalls.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
            except ValueError:
                pass
    for v, name in sorted(stalls, reverse=True)[:6]:
        print(f"   stall/issue {name:<28} {v:.3f}")
