"""SASS opcode histogram per kernel of the built library (evidence for TMA / packed-add / no-FMA-contraction claims).
Usage: python scripts/sass_histogram.py [lib.so] > profiles/<name>.md"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "librectify_b200/lib/librectify_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern = None
hist = collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("lrb::(anonymous namespace)::", "").replace("(anonymous namespace)::", "").replace("void ", "")
        name = re.sub(r"\(lrb::.*|\(unsigned.*|\(int.*|\(float.*|\(CUtensorMap.*|\(.*$", "", name)
        kern = name
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z0-9_]+)(\.[A-Z0-9_.]+)?", line)
    if m and kern:
        op = m.group(1); full = op + (m.group(2) or "")
        hist[kern][op] += 1
        if op in ("UTMALDG", "SYNCS", "LDG", "STG", "REDG", "ATOMG", "MUFU"):
            hist[kern][full.split(" ")[0]] += 0
        if full.startswith("UTMALDG") or full.startswith("SYNCS"):
            hist[kern][full] += 1
cols = ["UTMALDG", "SYNCS", "FADD2", "FADD", "FMUL", "FFMA", "FFMA2", "FMNMX3", "FMNMX", "MUFU", "PRMT", "VOTE", "SHFL", "BREV", "LDG", "STG", "LDS", "STS", "REDG", "ATOMG"]
print("# SASS opcode histogram of the built library (`cuobjdump -sass librectify_b200/lib/librectify_b200.so`, sm_100a)\n")
print("Static instruction counts per kernel.  `UTMALDG` = TMA tensor loads (`cp.async.bulk.tensor`), `SYNCS` = mbarrier operations, `FADD2` = packed `add.rn.f32x2`.")
print("All translation units are compiled with `-fmad=false`: the `FFMA` that remain are the compiler's own IEEE division / square-root")
print("sequences (`__fdiv_rn`, `__fsqrt_rn`) and the FFMA micro-benchmark; no `FFMA2` and no FFMA in the tap sums of K1.\n")
print("| kernel | total | " + " | ".join(cols) + " |")
print("|---|---:|" + "---:|" * len(cols))
for k, h in hist.items():
    tot = sum(v for op, v in h.items() if "." not in op)
    print(f"| `{k}` | {tot} | " + " | ".join(str(h.get(c, 0)) for c in cols) + " |")
