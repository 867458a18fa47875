"""One line per captured launch of an ncu report: python tools/ncu_summary.py <rep>"""
import csv, io, subprocess, sys
M = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__grid_size",
     "lts__t_bytes.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv", "--metrics", ",".join(M)], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h = rows[0]
cols = [h.index("Kernel Name")] + [h.index(m) for m in M if m in h]
print(",".join(["id"] + [h[c] for c in cols]))
print(",".join([""] + [rows[1][c] for c in cols]))
for i, r in enumerate(rows[2:]):
    print(",".join([str(i)] + [r[c][:48].replace(",", ";") for c in cols]))
