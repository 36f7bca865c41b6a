"""Print selected metrics of selected kernels from an `ncu --page raw --csv` export.  Usage: ncu_pick.py file.csv kernel-substring [metric-substring ...]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
ki = hdr.index("Kernel Name")
want = sys.argv[3:] or ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct", "sm__throughput.avg.pct",
                         "sm__inst_executed.sum", "smsp__inst_executed.sum", "issue_active.avg.pct", "pipe_fma", "pipe_alu", "pipe_fmaheavy", "pipe_fp32", "pipe_lsu", "pipe_xu",
                         "launch__registers", "achieved_occupancy", "warps_active.avg.pct", "l1tex__data_bank_conflicts", "lts__t_bytes.sum", "icc", "stall"]
seen = set()
for r in rows[2:]:
    if sys.argv[2] in r[ki] and r[ki] not in seen:
        seen.add(r[ki])
        print("==", r[ki][:100])
        for h, u, v in zip(hdr, units, r):
            if any(s in h for s in want):
                print(f"  {h:90s} {u:14s} {v}")
