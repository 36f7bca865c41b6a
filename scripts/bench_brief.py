"""Short view of a bench.py JSON line: python scripts/bench_brief.py gpurun_out/bench.json"""
import json
import sys

d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value %.0f  e2e %.0f  u8 %.0f  launches %s  clocks %s" % (d["value"], d["e2e"]["value"], d.get("e2e_u8", {}).get("value", 0),
                                                                  d.get("gpu_launches"), d.get("clocks")))
for k in ("stage_ms_per_frame", "stage_ms_per_frame_alone", "stage_share_of_step"):
    if k in d:
        print(k, {a: round(b, 4) for a, b in d[k].items()})
rs = d.get("roofline_stages", {})
for k, v in rs.items():
    print(k, "frac", round(v.get("frac", 0), 3), "alone", round(v.get("alone", {}).get("frac", 0), 3) if isinstance(v.get("alone"), dict) else None)
c2 = d.get("config2_single_frame", {})
print("latency p50", c2.get("p50_ms_device_resident"), {a: round(b, 3) for a, b in c2.get("stage_ms", {}).items()}, "kernels", c2.get("kernels_per_frame"))
for k in ("config4_4k_clutter", "config5_12mp_mixed"):
    if k in d:
        print(k, round(d[k]["images_per_s"], 1), "img/s", round(d[k]["latency_ms_p50"], 2), "ms")
print("dropin", {k: round(v["images_per_s"]) for k, v in d.get("dropin", {}).items() if isinstance(v, dict)})
print("parity", d.get("parity_sample", {}).get("segments_and_groups_byte_identical_pct"), "device==host", d.get("device_vs_host_results_identical"))
