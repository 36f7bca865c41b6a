python -m pytest tests -m gpu -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py 2>/dev/null > gpurun_out/bench_final.json
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_final.json").read().strip().splitlines()[-1])
print(round(d["value"]), round(d["e2e"]["value"]), round(d["e2e_u8"]["value"]), round(d["roofline"]["frac"],3), round(d["roofline"]["k2_nms_GBps"]), d["gpu_launches"], d["clocks"])
PY
