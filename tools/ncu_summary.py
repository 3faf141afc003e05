"""Prints the headline metrics and the hottest source lines of an .ncu-rep (first kernel in it)."""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
        "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed"]
for i, h in enumerate(hdr):
    if h in want or (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio")
                     and float(vals[i] or 0) > 0.2) or h == "smsp__average_warp_latency_per_inst_issued.ratio":
        print("%-90s %-14s %s" % (h, units[i], vals[i]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
lines = [r for r in rows[hi + 1:] if len(r) > 8 and r[2] == "-" and r[4].isdigit()]
tot = sum(int(r[4]) for r in lines)
print("total samples", tot)
sidx = {n: hdr.index(n) for n in hdr if n.startswith("stall_") and "Not Issued" not in n}
lines.sort(key=lambda r: -int(r[4]))
for r in lines[:top]:
    st = {k[6:]: int(r[i]) for k, i in sidx.items() if i < len(r) and r[i].isdigit() and int(r[i]) > 0}
    t3 = sorted(st.items(), key=lambda x: -x[1])[:3]
    print(r[0].rjust(4), r[4].rjust(6), r[7].rjust(9), r[1][:84].ljust(84), t3)
