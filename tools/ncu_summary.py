"""Summarise an ncu report (--set full) as `metric unit | kernel A | kernel B ...` for the metrics the roofline reading uses.
usage: python tools/ncu_summary.py report.ncu-rep > profiles/<tag>_full_summary.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
names, units, data = rows[hdr], rows[hdr + 1], rows[hdr + 2:]
keep = ("Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput",
        "gpu__dram_throughput", "l1tex__t_sector_hit_rate", "l1tex__throughput", "lts__t_sector_hit_rate", "lts__throughput",
        "sm__throughput", "sm__warps_active", "smsp__issue_active", "sm__inst_executed_pipe_alu", "sm__inst_executed_pipe_fma.",
        "sm__inst_executed_pipe_fp64", "sm__inst_executed_pipe_lsu", "smsp__average_warps_issue_stalled", "launch__registers_per_thread",
        "launch__occupancy_limit", "sm__maximum_warps_per_active_cycle_pct", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed",
        "launch__shared_mem_per_block", "local_load", "local_store", "smsp__inst_executed_op_local")
print("# columns: metric unit | " + " | ".join(r[names.index("Kernel Name")][:40] for r in data))
for j, n in enumerate(names):
    if any(n.startswith(k) or k in n for k in keep):
        print(f"{n} {units[j]} | " + " | ".join(r[j] for r in data))
