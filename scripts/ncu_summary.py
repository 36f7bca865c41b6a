"""Summarise an `ncu --page raw --csv` export (one section per distinct kernel, first launch of each) into a markdown file.
Usage: python scripts/ncu_summary.py raw.csv out.md "title" ["command line that produced the report"]"""
import csv
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'smsp__inst_executed.sum',
        'sm__issue_active.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_static',
        'sm__icc_request_hit_rate.pct', 'smsp__warps_eligible.avg.per_cycle_active',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio']

src, dst, title = sys.argv[1], sys.argv[2], sys.argv[3]
cmd = sys.argv[4] if len(sys.argv) > 4 else ""
rows = list(csv.reader(open(src)))
hdr, units = rows[0], rows[1]
ki = hdr.index('Kernel Name')
out = [f"# {title}\n"]
if cmd:
    out.append(f"Command: `{cmd}` (the .ncu-rep stays in gpurun_out/, untracked; this file is its summary).")
out.append("Durations under ncu are cold-cache and serialised: use them for shares and per-kernel diagnosis, not as bench values.\n")
seen = set()
for r in rows[2:]:
    short = r[ki].split('(')[0].replace('void ', '').replace('unnamed>::', '').replace('lrb::<', '')
    if short in seen:
        continue
    seen.add(short)
    out.append(f"## `{short}`\n\n| metric | unit | value |\n|---|---|---:|")
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            out.append(f"| `{w}` | {units[i]} | {r[i]} |")
    out.append("")
open(dst, "w").write("\n".join(out))
print(dst, len(seen), "kernels")
