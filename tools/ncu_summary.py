"""Key metrics per kernel from one or more `ncu --set full` reports, as one CSV (the profiles/*_summary.csv files).
    python tools/ncu_summary.py a.ncu-rep b.ncu-rep ... > profiles/rNN_ncu_full_xxx_summary.csv"""
import csv, subprocess, sys

COLS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
w = csv.writer(sys.stdout)
first = True
for rep in sys.argv[1:]:
    rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
    h = rows[0]
    idx = [h.index(c) if c in h else -1 for c in COLS]
    if first:
        w.writerow(COLS); w.writerow([rows[1][i] if i >= 0 else "" for i in idx]); first = False
    for r in rows[2:]:
        w.writerow([r[i] if i >= 0 else "" for i in idx])
