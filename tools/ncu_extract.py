"""profiles/rNN_ncu_<kernel>.json from one `ncu --set full` report: the numbers bench.py quotes in `roofline.traffic` / `roofline.issue`.
    python tools/ncu_extract.py gpurun_out/r02_ncu_extend_pool24.ncu-rep k_extend RAYS_IN_LAUNCH COMMIT "how it was captured" > profiles/r02_ncu_extend.json
RAYS_IN_LAUNCH: the rays the captured launch traced (the kernel does not count per launch; from the run's stats)."""
import csv
import json
import subprocess
import sys

rep, kernel, rays, commit, how = sys.argv[1], sys.argv[2], float(sys.argv[3]), sys.argv[4], sys.argv[5]
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()))
h = rows[0]
r = next(x for x in rows[2:] if kernel in x[h.index("Kernel Name")])
g = lambda name: float(r[h.index(name)].replace(",", "")) if name in h and r[h.index(name)] not in ("", "n/a") else None
unit = lambda name: rows[1][h.index(name)] if name in h else ""
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
rd = g("dram__bytes_read.sum") * scale.get(unit("dram__bytes_read.sum"), 1.0)
wr = g("dram__bytes_write.sum") * scale.get(unit("dram__bytes_write.sum"), 1.0)
dur = g("gpu__time_duration.sum")
dur_ms = dur * {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0, "second": 1e3}.get(unit("gpu__time_duration.sum"), 1e-6)
print(json.dumps({
    "kernel": kernel, "commit": commit, "how": how,
    "dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes_per_launch": rd + wr, "rays_in_captured_launch": rays,
    "dram_bytes_per_ray": (rd + wr) / rays, "duration_ms_under_ncu": dur_ms,
    "lanes_per_inst": g("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "warps_active_pct": g("sm__warps_active.avg.pct_of_peak_sustained_active"),
    "alu_pipe_pct": g("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    "fma_pipe_pct": g("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    "xu_pipe_pct": g("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    "l1_hit_pct": g("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": g("lts__t_sector_hit_rate.pct"),
    "registers": g("launch__registers_per_thread"), "inst_executed": g("smsp__inst_executed.sum"),
}, indent=1))
