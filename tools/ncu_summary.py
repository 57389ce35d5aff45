"""Print the handful of ncu raw metrics we track from a .ncu-rep (development aid)."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__average_warps_issue_stalled", "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
        "sm__throughput.avg.pct", "l1tex__throughput.avg.pct", "lts__throughput.avg.pct"]
for i, name in enumerate(h):
    if any(name.startswith(w) or w == name for w in want):
        vals = [r[i] for r in rows[2:]]
        if any(v not in ("0", "") for v in vals):
            print(f"{name} [{rows[1][i]}]: {vals}")
