"""Transpose `ncu -i X.ncu-rep --page raw --csv` into metric rows x launch columns, keeping the metrics the roofline
discussion uses.   usage: python tools/ncu_summary.py raw.csv > summary.csv"""
import csv
import sys

KEEP = ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "smsp__issue_active.avg.pct", "smsp__inst_executed.sum", "issue_stalled", "sm__pipe_tensor", "sm__inst_executed_pipe_alu.avg.pct",
        "sm__inst_executed_pipe_fma.avg.pct", "sm__inst_executed_pipe_lsu.avg.pct", "sm__inst_executed_pipe_xu.avg.pct",
        "sm__inst_executed_pipe_tensor", "lts__throughput", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__warps_active.avg.pct", "l1tex__throughput", "sm__throughput")
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [f"launch{i}" for i in range(len(data))])
for c, (h, u) in enumerate(zip(hdr, units)):
    if any(k in h for k in KEEP):
        w.writerow([h, u] + [d[c] for d in data])
