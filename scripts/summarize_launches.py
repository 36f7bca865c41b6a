"""Turn an `ncu --metrics gpu__time_duration.sum --csv` launch list into a per-kernel summary (profiles/*.md)."""
import collections
import csv
import re
import sys

src, dst, title = sys.argv[1], sys.argv[2], sys.argv[3]
lines = [l for l in open(src) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    m = re.search(r"\b(k_[a-z0-9_]+)", row["Kernel Name"])
    name = m.group(1) if m else re.sub(r"\(.*", "", row["Kernel Name"])
    v = float(row["Metric Value"].replace(",", ""))
    v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
    a = agg.setdefault(name, [0, 0.0, row["Grid Size"], row["Block Size"]])
    a[0] += 1; a[1] += v
ours = {k: v for k, v in agg.items() if k.startswith("k_")}
tot = sum(v[1] for k, v in ours.items() if k != "k_mt_stream")
with open(dst, "w") as f:
    f.write(f"# {title}\n\nSource: `{src}` (ncu --metrics gpu__time_duration.sum --clock-control none; per-launch times are cold-cache and\nserialised — compare SHARES, not absolutes). Shares exclude the one-off `k_mt_stream` (uniform stream, once per seed).\n\n")
    f.write("| kernel | launches | total us | avg us | share of step | grid | block |\n|---|---:|---:|---:|---:|---|---|\n")
    for k, (c, t, g, b) in sorted(ours.items(), key=lambda kv: -kv[1][1]):
        share = "one-off" if k == "k_mt_stream" else f"{100 * t / tot:.1f} %"
        f.write(f"| `{k}` | {c} | {t:.1f} | {t / c:.1f} | {share} | {g} | {b} |\n")
    f.write(f"\nTotal of this repo's kernels (without k_mt_stream): {tot:.1f} us over the profiled launches.\n")
print(open(dst).read())
