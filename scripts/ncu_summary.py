"""Condense an .ncu-rep (read with `ncu -i rep --page raw --csv`) into the per-kernel table kept under profiles/.
usage: ncu -i x.ncu-rep --page raw --csv | python scripts/ncu_summary.py > profiles/x.csv"""
import csv
import sys

rows = list(csv.reader(sys.stdin))
hdr, units, data = rows[0], rows[1], rows[2:]
want = [
    "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sectors_srcunit_tex_op_read.sum",
    "lts__t_sectors_srcunit_tex_op_write.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
]
idx = {h: i for i, h in enumerate(hdr)}
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [r[idx["Kernel Name"]].split("(")[0].split("::")[-1] for r in data])
for m in want[1:]:
    if m in idx:
        w.writerow([m, units[idx[m]]] + [r[idx[m]] for r in data])
