"""Per-kernel summary of an ncu launch list that carries time, warp instructions and DRAM / L2 traffic:
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_op_read.sum,
lts__t_sectors_op_write.sum,lts__t_sectors_op_atom.sum,lts__t_sectors_op_red.sum --csv ...
Usage: python scripts/summarize_launches_dram.py launches.csv out.md "title" frames [--timeline]"""
import collections
import csv
import sys

src, dst, title, frames = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
timeline = "--timeline" in sys.argv
rows = [r for r in csv.reader(open(src)) if len(r) == 15 and r[0] != "ID"]
agg = collections.defaultdict(lambda: collections.defaultdict(float))
cnt = collections.Counter(); seen = set(); grid = {}; seq = []


def short(name):
    return name.split("(")[0].replace("unnamed>::", "").replace("void ", "").replace("lrb::<", "").split("<")[0]


for r in rows:
    name, metric, unit = short(r[4]), r[12], r[13]
    v = float(r[14].replace(",", ""))
    if metric == "gpu__time_duration.sum":
        v = v / 1000 if unit == "ns" else (v * 1000 if unit == "ms" else v)
        seq.append((name, v, r[8]))
    if "bytes" in metric:
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    agg[name][metric] += v
    if r[0] not in seen:
        seen.add(r[0]); cnt[name] += 1; grid[name] = (r[8], r[7])
tot = collections.defaultdict(float)
for k, v in agg.items():
    if k != "k_mt_stream":
        for m, x in v.items():
            tot[m] += x
T = tot["gpu__time_duration.sum"]
out = [f"# {title}\n", f"Source: `{src}` (per-launch values under ncu are cold-cache and serialised: compare SHARES, not absolutes; the one-off "
       "`k_mt_stream` is excluded from the totals).\n",
       "| kernel | launches | total µs | share | M warp instr | DRAM read MB | DRAM write MB | L2 read MB | L2 write+atomic MB | grid | block |",
       "|---|---:|---:|---:|---:|---:|---:|---:|---:|---|---|"]
for k, v in sorted(agg.items(), key=lambda kv: -kv[1]["gpu__time_duration.sum"]):
    l2r = v["lts__t_sectors_op_read.sum"] * 32 / 1e6
    l2w = (v["lts__t_sectors_op_write.sum"] + v["lts__t_sectors_op_atom.sum"] + v["lts__t_sectors_op_red.sum"]) * 32 / 1e6
    share = "one-off" if k == "k_mt_stream" else f"{100 * v['gpu__time_duration.sum'] / T:.1f} %"
    out.append(f"| `{k}` | {cnt[k]} | {v['gpu__time_duration.sum']:.1f} | {share} | {v['smsp__inst_executed.sum'] / 1e6:.1f} | "
               f"{v['dram__bytes_read.sum'] / 1e6:.1f} | {v['dram__bytes_write.sum'] / 1e6:.1f} | {l2r:.1f} | {l2w:.1f} | {grid[k][0]} | {grid[k][1]} |")
out.append(f"\nTotals (without k_mt_stream): {T:.0f} µs, {tot['smsp__inst_executed.sum'] / 1e6:.0f} M warp instructions, DRAM "
           f"{tot['dram__bytes_read.sum'] / 1e6:.0f} MB read + {tot['dram__bytes_write.sum'] / 1e6:.0f} MB written over {frames} frame(s) = "
           f"**{tot['smsp__inst_executed.sum'] / 1e6 / frames:.1f} M warp instructions and {(tot['dram__bytes_read.sum'] + tot['dram__bytes_write.sum']) / 1e6 / frames:.0f} MB of DRAM traffic per frame**.")
if timeline:
    out.append("\n## Launch by launch\n\n| # | kernel | µs | grid |\n|---:|---|---:|---|")
    for i, (n, us, g) in enumerate(seq):
        out.append(f"| {i} | `{n}` | {us:.1f} | {g} |")
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out[:40]))
