#!/usr/bin/env python
"""Benchmark of the librectify hot path (detect + VP grouping + transform) on B200 — driver contract in the task brief.

  python bench.py --gpus N --steps K --warmup W            this repo's CUDA path (one process per GPU under torchrun)
  python bench.py --impl reference --steps K --warmup W    the reference's own CPU code (oracle/_ref, OpenMP, all cores)

Metric (BASELINE.json): 1080p images/s through find_line_segment_groups + compute_rectification_transform.
Workload = BASELINE config 3 as written: ONE batch of 4096 distinct synthetic 1920x1080 Manhattan-scene frames (SURVEY.md §8d
generator, seeds 0..4095, min_length 19.2, H+V RECTIFY), sharded in contiguous blocks over the N ranks (frame i -> rank
floor(i*N/4096): strong scaling, 4096/N frames per GPU), no data-path collective.  A step = one pass of the hot path over the
whole batch; `value` = 4096 * K / max-over-ranks time.
  value : frames already resident in HBM (device pointers handed to the C ABI, lrb_pool_rectify_device); results read back and
          the corner solve on the host are inside the timed region (they are part of the path).
  e2e   : the same call with HOST (pinned) frames, lrb_pool_rectify_host — H2D copies and result read-back inside the timed region.
  roofline : per stage, all from CUDA events inside the timed region (and, beside them, from one engine running alone);
          the headline entry is the kernel group with the largest share of the step.
Extra legs (outside the timed region; extra keys): single-frame latency with a per-stage breakdown (config 2), configs 4 and 5,
config 1 (doc/image.jpg fixture), the drop-in symbols from pageable memory, the CPU baselines and a parity sample vs the oracle.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
MIN_LENGTH = 19.2
METRIC = "1080p images/sec (detect+VP+transform)"
K1_BYTES_PER_PX = 17.0          # 4 B read (image) + 8 B (dx, dy) + 4 B (mag) + 1 B (bin) written     
K2_BYTES_PER_PX = 6.0           # 4 B (mag) + 1 B (bin) read, 1 B (mask8) written
K8_FLOP_PER_PAIR = 16.0         # flops of the reference's formula per (hypothesis, segment) pair
# DRAM traffic of K1 / K2 from the tracked ncu capture profiles/r2_ncu_full_stencils_seeds.md (dram__bytes_read.sum + dram__bytes_write.sum
# of one 32-frame launch: K1 265.4 + 802.4 MB, K2 332.6 + 61.2 MB); constants of that capture, not measured in this run
K1_NCU_DRAM_BYTES_PER_PX = (265.444e6 + 802.417e6) / (32 * 1920 * 1080)
K2_NCU_DRAM_BYTES_PER_PX = (332.646e6 + 61.236e6) / (32 * 1920 * 1080)
# warp instructions the flood kernels (prune + sub-warp + full-tile + verify, all rounds) issue per 1080p frame: smsp__inst_executed.sum
# over every flood launch of one 32-frame group, from the tracked launch list named in FLOOD_INSTR_SOURCE
FLOOD_WARP_INSTR_PER_FRAME = 42.6e6
# DRAM bytes of the flood kernels per 1080p frame from the same launch list (dram__bytes_read.sum + dram__bytes_write.sum: sub-warp pass
# 147 MB, full-tile pass 81 MB, verify 22 MB, prune 5 MB) and what a flood would have to read at least: every plane it looks at once
# (dx, dy 8 B/px, mask | visited 2 B/px) plus one owner stamp written and read per tentative pixel (~0.45 M pixels x 8 B)
FLOOD_DRAM_BYTES_PER_FRAME = 255e6
FLOOD_ALGORITHMIC_BYTES_PER_FRAME = 1920 * 1080 * 10 + 0.45e6 * 8
FLOOD_INSTR_SOURCE = "profiles/r2d_launches_32frames.md (ncu smsp__inst_executed.sum, 1364 M per 32-frame group: full-tile 776.3 + sub-warp 439.9 + verify 136.2 + prune 12.0)"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (profiling recipe's clocks line)"""

    def __init__(self, device):
        self.rows = []
        self.stop = threading.Event()
        self.device = device
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop.wait(0.2)

    def __enter__(self):
        self.t.start(); return self

    def __exit__(self, *a):
        self.stop.set(); self.t.join(timeout=6)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def make_frames(seeds, device, w=W, h=H, n_lines=40, pin=False):
    """[len(seeds), h, w] f32 on `device`, one synthetic Manhattan scene per seed (librectify_b200/synth.py)"""
    import torch
    from librectify_b200.synth import manhattan_frame
    out = torch.empty((len(seeds), h, w), dtype=torch.float32, device=device, pin_memory=pin)
    for i, s in enumerate(seeds):
        out[i] = manhattan_frame(int(s), w, h, n_lines=n_lines, device=device)
    return out


def mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return 0


# ---- CPU arms: the reference's own CPU implementation (compiled from /root/reference into oracle/_ref) -----------------------
def _ref_impl():
    from oracle.bind import Ref, ref_available
    if ref_available("ship"):
        return Ref("ship"), "reference"
    from oracle.bind import Oracle      # the compiled reference did not travel: the restatement (scalar, one core at a time)
    return Oracle(), "port"


def _cpu_pass(impl, frames, threads, cfg, w, h, min_length):
    t0 = time.perf_counter()
    for f in frames:
        segs = impl.find_line_segment_groups(f, min_length, num_threads=threads, seed=0)
        if segs is not None:
            impl.compute_rectification_transform(segs, w, h, cfg)
    return time.perf_counter() - t0


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    from oracle.bind import RectificationConfig
    impl, kind = _ref_impl()
    cores = os.cpu_count() or 1
    threads = cores if kind == "reference" else 1
    per_step = args.ref_frames
    frames = make_frames(range(per_step), "cpu").numpy()
    cfg = RectificationConfig.default()
    for _ in range(args.warmup):
        _cpu_pass(impl, frames, threads, cfg, W, H, MIN_LENGTH)
    dt = sum(_cpu_pass(impl, frames, threads, cfg, W, H, MIN_LENGTH) for _ in range(args.steps))
    value = per_step * args.steps / dt
    sample = f"{per_step} synthetic 1920x1080 frames per step (seeds 0..{per_step - 1} of the 4096), {args.steps} steps, num_threads={threads}"
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "BASELINE config 3 frames (synthetic 1920x1080 Manhattan scenes, min_length 19.2, H+V RECTIFY); each step is a "
                                   f"bounded sample of {per_step} frames of the 4096-frame batch — the metric is per image, so the sample size "
                                   "does not enter it (the GPU arm runs the whole batch per step)",
                       "frames_per_step": per_step, "threads": threads,
                       "note": "the reference parallelises inside one image with OpenMP; its flood is serial and its sampler sits in an omp critical"},
            "cpu_baseline": {"value": value, "unit": "images/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def cpu_baselines(frames, n_omp=24, n_serial=6, per_thread=2):
    """the reference's CPU path on the box's host cores, on bounded samples of the step's frames (rank 0, N=1):
    OpenMP with all cores (the reference's own mode), serial (num_threads=-1), `cores` single-thread callers on independent
    frames (the fair throughput baseline), and the corner gap between the shipped-flags and the canonical build."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle.bind import RectificationConfig, Ref, ref_available, transform_to_dict
    out = {}
    try:
        impl, kind = _ref_impl()
        cores = os.cpu_count() or 1
        cfg = RectificationConfig.default()
        threads = cores if kind == "reference" else 1
        impl.find_line_segment_groups(frames[0], MIN_LENGTH, num_threads=threads, seed=0)       # warm-up
        n = min(n_omp, len(frames))
        dt = _cpu_pass(impl, frames[:n], threads, cfg, W, H, MIN_LENGTH)
        out = {"value": n / dt, "unit": "images/s", "cores": threads, "kind": kind,
               "sample": f"{n} of the step's frames (seeds 0..{n - 1}), num_threads={threads}, {dt:.1f} s"}
        n = min(n_serial, len(frames))
        dt = _cpu_pass(impl, frames[:n], -1, cfg, W, H, MIN_LENGTH)
        out["serial"] = {"value": n / dt, "unit": "images/s", "cores": 1, "sample": f"{n} frames, num_threads=-1, {dt:.1f} s"}
        n = min(cores * per_thread, len(frames))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(max_workers=cores) as ex:       # ctypes releases the GIL; the reference is re-entrant
            list(ex.map(lambda f: _cpu_pass(impl, [f], -1, cfg, W, H, MIN_LENGTH), frames[:n]))
        dt = time.perf_counter() - t0
        out["independent_callers"] = {"value": n / dt, "unit": "images/s", "cores": cores,
                                      "sample": f"{n} frames, {cores} concurrent single-thread callers (num_threads=-1 each), {dt:.1f} s"}
        if kind == "reference" and ref_available("canon"):
            canon = Ref("canon"); worst = 0.0; m = min(4, len(frames))
            for f in frames[:m]:
                a = impl.find_line_segment_groups(f, MIN_LENGTH, num_threads=-1, seed=0)
                b = canon.find_line_segment_groups(f, MIN_LENGTH, num_threads=-1, seed=0)
                if a is None or b is None:
                    continue
                ta = transform_to_dict(impl.compute_rectification_transform(a, W, H, cfg)); tb = transform_to_dict(canon.compute_rectification_transform(b, W, H, cfg))
                for c in ("top_left", "top_right", "bottom_left", "bottom_right"):
                    d = np.abs(np.asarray(ta[c][:2]) - np.asarray(tb[c][:2]))
                    worst = max(worst, float(np.nanmax(d)) if np.all(np.isfinite(d)) else worst)
            out["ship_vs_canon_corner_gap_px"] = {"value": worst, "sample": f"{m} frames, same seed, shipped flags (-ffast-math) vs strict-IEEE build: the reference's own noise floor"}
    except Exception as ex:  # the baseline is a reported number, never a reason to lose the GPU line
        out.setdefault("value", None); out.setdefault("unit", "images/s"); out.setdefault("cores", 0); out.setdefault("kind", "unavailable")
        out["error"] = repr(ex)[:200]
    return out


# ---- this repo's arm ----------------------------------------------------------------------------------------------------
def stage_table(prof, n_frames):
    return {k: v / max(n_frames, 1) for k, v in prof["ms"].items()}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from librectify_b200 import api
    from librectify_b200.runner import Runner, shard_bounds

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the librectify_b200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    total = args.frames
    lo, hi = shard_bounds(total, world, rank)                                 # frame i -> rank floor(i * world / total)
    B = hi - lo
    t_setup = time.perf_counter()
    frames_dev = make_frames(range(lo, hi), dev)                              # [B, H, W] f32 in HBM: B * 8.3 MB >> L2 (126 MB)
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t_setup
    runner = Runner(local, args.engines, args.slots)
    cfg = api.make_config()                                                  # H+V RECTIFY (reference defaults)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        """K steps bracketed by barrier + synchronize; device clock via CUDA events; max over ranks"""
        barrier()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        fn(steps)                                                            # K steps, streamed (no join between steps inside a rank)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        ms = max(e0.elapsed_time(e1), 0.0)
        # the C ABI call is host-synchronous on the engines' own streams: the events bracket it in time; take the larger of
        # the event interval and the host wall clock so that neither can under-report
        sec = max(ms / 1e3, wall)
        t = torch.tensor([sec], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    out = {}

    def step_dev(steps=1):
        out["tf"], out["info"] = runner.rectify(frames_dev, MIN_LENGTH, cfg, steps=steps)

    warm = max(args.warmup, 3)
    step_dev(warm)
    runner.profile(True)
    with ClockSampler(local) as clk:
        sec = timed(step_dev, args.steps)
    prof = runner.profile_read()
    runner.profile(False)

    # ---- e2e: the same batch from pinned host memory (bounded only if the host cannot hold it) -----------------------------
    need = B * W * H * 4
    Be = B if mem_available_bytes() > 3 * need + (16 << 30) else max(args.slots * args.engines, min(B, int((mem_available_bytes() // 4) // (W * H * 4))))
    frames_host = torch.empty((Be, H, W), dtype=torch.float32, pin_memory=True)
    frames_host.copy_(frames_dev[:Be])
    torch.cuda.synchronize()
    host_np = frames_host.numpy()

    def step_host(steps=1):
        out["tf_h"], out["info_h"] = runner.rectify(host_np, MIN_LENGTH, cfg, steps=steps)

    step_host(1)
    runner.profile(True)
    sec_e2e = timed(step_host, args.steps)
    prof_e2e = runner.profile_read()
    runner.profile(False)
    same = bool(np.array_equal(out["tf"][:Be].view(np.uint8), out["tf_h"].view(np.uint8)))

    # the same frames quantised to 8 bits, for the device-ingest entry point (lrb_pool_rectify_host_u8):
    # 1 byte per pixel over PCIe instead of 4.  Reported as an extra; the headline e2e stays the float call.
    Bu = min(Be, 1024)
    frames_u8 = torch.empty((Bu, H, W), dtype=torch.uint8, pin_memory=True)
    frames_u8.copy_((frames_dev[:Bu] * 256.0).clamp_(0, 255).to(torch.uint8))
    u8_np = frames_u8.numpy()
    runner.rectify(u8_np, MIN_LENGTH, cfg)
    sec_e2e_u8 = timed(lambda steps: runner.rectify(u8_np, MIN_LENGTH, cfg, steps=steps), args.steps)

    info = out["info"]
    n_total = total * args.steps
    value = n_total / sec
    e2e = (world * Be * args.steps) / sec_e2e if Be != B else n_total / sec_e2e
    peak, peak_src = measured_peaks()
    line = {
        "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"BASELINE config 3: one batch of {total} distinct synthetic 1920x1080 Manhattan-scene frames (seeds 0..{total - 1}), "
                               f"min_length {MIN_LENGTH}, H+V RECTIFY, detect+VP+transform; sharded contiguously over {world} GPU(s) "
                               f"({B} frames on rank 0), one pass = one step",
                   "frames_per_step": total, "frames_per_gpu": B, "engines_per_gpu": args.engines, "slots_per_engine": args.slots,
                   "l2": f"inputs {B * W * H * 4 / 1e9:.1f} GB per GPU per step > 126 MB L2 (no flush needed)",
                   "parallelism": f"frames sharded over {world} GPU(s), no collective", "setup_s_generating_frames": round(setup_s, 1)},
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": int(prof_e2e["h2d_bytes"] // args.steps),
                "d2h_bytes_per_step": int(prof_e2e["d2h_bytes"] // args.steps),
                "frames_per_step_per_gpu": Be, "api": "lrb_pool_rectify_host (pinned host frames; H2D and read-back inside the timed region)",
                "bytes_source": "counted by the engines from the copies they issue (lrb_profile_read_bytes)"},
        "e2e_u8": {"value": world * Bu * args.steps / sec_e2e_u8, "unit": "images/s", "h2d_bytes_per_step": int(Bu * W * H),
                   "frames_per_step_per_gpu": Bu,
                   "note": "first frames of the batch quantised to 8-bit gray, lrb_pool_rectify_host_u8 (gray/256 on the device); extra, not the headline"},
        "gpu_launches": int(prof["kernels"]),
        "scoring_pairs_per_frame": float(np.mean(info["n_obs"].sum(axis=1)) * 10000),
        "frames_with_lines": int((info["n_lines"] > 0).sum()), "flood_fallbacks": int(info["flood_fallback"].sum()),
        "mean_flood_rounds": float(np.mean(info["flood_rounds"])), "max_flood_rounds": int(np.max(info["flood_rounds"])), "mean_seeds": float(np.mean(info["n_seeds"])),
        "device_vs_host_results_identical": same,
        "clocks": clk.summary(),
    }

    # ---- per-stage rooflines -------------------------------------------------------------------------------------------------
    groups = max(prof["groups"], 1)
    fpl = B * args.steps / groups                                             # frames per launch group
    st_bench = stage_table(prof, B * args.steps)                              # ms per frame, 4 engines time-slicing the SMs
    rank0_extra = rank == 0 and not args.quick
    alone = None; fp32 = None
    probe = api.Engine(local, args.slots)
    try:
        fp32 = probe.ubench_fp32()
        nprobe = min(B, 4 * args.slots) // args.slots * args.slots
        if nprobe >= args.slots:
            probe.rectify(frames_dev[:args.slots], MIN_LENGTH, cfg)
            probe.profile(True)
            for i0 in range(0, nprobe, args.slots):
                probe.rectify(frames_dev[i0:i0 + args.slots], MIN_LENGTH, cfg)
            alone = stage_table(probe.profile_read(), nprobe)
            probe.profile(False)
    finally:
        probe.close()
    px = W * H
    pairs = line["scoring_pairs_per_frame"]

    def hbm_entry(name, stage, bpp, traffic=None):
        e = {"kernel": name, "bound": "hbm", "unit": "GB/s", "peak": peak, "peak_source": peak_src, "algorithmic_bytes_per_px": bpp,
             "frames_per_launch": fpl}
        ms_f = st_bench[stage]
        e["achieved"] = bpp * px / (ms_f * 1e-3) / 1e9 if ms_f > 0 else None
        e["frac"] = e["achieved"] / peak if e["achieved"] else None
        e["ms_per_launch"] = ms_f * fpl
        if alone and alone[stage] > 0:
            e["alone"] = {"achieved": bpp * px / (alone[stage] * 1e-3) / 1e9, "frac": bpp * px / (alone[stage] * 1e-3) / 1e9 / peak,
                          "ms_per_launch": alone[stage] * args.slots, "note": "one engine, nothing else on the GPU"}
        e["traffic"] = traffic
        return e

    k1 = hbm_entry("k_grad5x5_bin (K1)", "grad", K1_BYTES_PER_PX, K1_NCU_DRAM_BYTES_PER_PX * px * fpl)
    k1["traffic_source"] = "constant from the tracked ncu capture profiles/r2_ncu_full_stencils_seeds.md (16.1 B/px), scaled to this launch size; not measured in this run"
    k2 = hbm_entry("k_nms5x5_seeds (K2)", "nms", K2_BYTES_PER_PX, K2_NCU_DRAM_BYTES_PER_PX * px * fpl)
    k2["traffic_source"] = "constant from the same capture (5.9 B/px)"
    stages = {"K1_grad": k1, "K2_nms": k2}
    if fp32:
        # "score" = CUDA events around the k_score launches of the <= 4 solves (a sub-interval of the RANSAC stage)
        flops = K8_FLOP_PER_PAIR * pairs
        k8 = {"kernel": "k_score (K8: 10^4 hypotheses x n segments scoring grid)", "bound": "fp32", "unit": "TFLOP/s",
              "peak": fp32["ffma"], "peak_fma_free": fp32["mul_add"],
              "peak_source": "measured in this run (lrb_ubench_fp32: FFMA chains; FMA-free FMUL/FADD chains — this library may not contract)",
              "algorithmic_flop_per_pair": K8_FLOP_PER_PAIR, "pairs_per_frame": pairs}
        if st_bench["score"] > 0:
            k8["achieved"] = flops / (st_bench["score"] * 1e-3) / 1e12; k8["frac"] = k8["achieved"] / fp32["ffma"]
            k8["frac_of_fma_free_peak"] = k8["achieved"] / fp32["mul_add"]; k8["ms_per_frame"] = st_bench["score"]
        if alone and alone["score"] > 0:
            a = flops / (alone["score"] * 1e-3) / 1e12
            k8["alone"] = {"achieved": a, "frac": a / fp32["ffma"], "frac_of_fma_free_peak": a / fp32["mul_add"]}
        stages["K8_score"] = k8
    sm_mhz = line["clocks"].get("sm_mhz") or 1965.0
    flood = {"kernel": "k_flood_* (K4: speculative region growing)", "bound": "latency / instruction issue, and scattered-sector DRAM traffic in the sub-warp pass (no single HBM or tensor roofline "
             "applies: data-dependent frontier expansion; hbm_view gives the memory side)", "unit": "fraction of issue slots",
             "ms_per_frame": st_bench["flood"], "ms_per_frame_alone": alone["flood"] if alone else None,
             "warp_instr_per_frame": FLOOD_WARP_INSTR_PER_FRAME, "instr_source": FLOOD_INSTR_SOURCE}
    if FLOOD_WARP_INSTR_PER_FRAME and alone and alone["flood"] > 0:
        slots_per_s = 148 * 4 * sm_mhz * 1e6
        flood["achieved"] = FLOOD_WARP_INSTR_PER_FRAME / (alone["flood"] * 1e-3) / slots_per_s
        flood["peak"] = 1.0; flood["frac"] = flood["achieved"]
        # the same stage seen from the memory side: the sub-warp pass runs at 3.3 TB/s of scattered sectors in a bulk round
        # (profiles/r2d_ncu_full_flood_bulk_round.md); over the whole stage the floods move ~10x their algorithmic bytes
        flood["traffic"] = FLOOD_DRAM_BYTES_PER_FRAME
        flood["hbm_view"] = {"unit": "GB/s", "peak": peak, "dram_bytes_per_frame": FLOOD_DRAM_BYTES_PER_FRAME,
                             "algorithmic_bytes_per_frame": FLOOD_ALGORITHMIC_BYTES_PER_FRAME,
                             "dram_GBps_alone": FLOOD_DRAM_BYTES_PER_FRAME / (alone["flood"] * 1e-3) / 1e9,
                             "frac_of_peak_alone": FLOOD_DRAM_BYTES_PER_FRAME / (alone["flood"] * 1e-3) / 1e9 / peak,
                             "algorithmic_GBps_alone": FLOOD_ALGORITHMIC_BYTES_PER_FRAME / (alone["flood"] * 1e-3) / 1e9,
                             "traffic_source": "constant from profiles/r2d_launches_32frames.md, per frame, not measured in this run"}
        flood["note"] = ("issue-slot fraction of the flood stage with ONE engine on the GPU = warp instructions per frame / (148 SMs x 4 schedulers x "
                         "SM clock x stage time per frame); with four engines the stage intervals overlap other engines' kernels, so no in-bench "
                         "fraction is given")
    stages["K4_flood"] = flood
    stage_sum = sum(v for k, v in st_bench.items() if k in api.STAGE_TIMERS)
    share = {k: v / max(stage_sum, 1e-12) for k, v in st_bench.items() if k in api.STAGE_TIMERS}
    top = max(share, key=share.get)
    head = dict({"flood": flood, "grad": k1, "nms": k2}.get(top, stages.get("K8_score", k1)))
    head["share_of_step"] = share[top]; head["stage"] = top
    head.setdefault("achieved", None); head.setdefault("peak", None); head.setdefault("frac", None); head.setdefault("traffic", None)
    line["roofline"] = head
    line["roofline_stages"] = stages
    line["stage_ms_per_frame"] = st_bench
    line["stage_ms_per_frame_alone"] = alone
    line["stage_share_of_step"] = share

    # ---- config 2: single-frame latency (one frame, one engine, device-resident), per-stage breakdown --------------------------
    one = api.Engine(local, 1)
    f1 = make_frames([1], dev) if (lo > 1 or hi <= 1) else frames_dev[1 - lo:2 - lo]
    lat = []
    for i in range(8):
        one.rectify(f1[0], MIN_LENGTH, cfg)
    for i in range(40):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        one.rectify(f1[0], MIN_LENGTH, cfg)
        lat.append(1e3 * (time.perf_counter() - t0))
    one.profile(True)
    for i in range(10):
        one.rectify(f1[0], MIN_LENGTH, cfg)
    p1 = one.profile_read(); one.profile(False)
    lat_host = []
    f1h = f1[0].cpu().numpy()
    for i in range(25):
        t0 = time.perf_counter(); one.rectify(f1h, MIN_LENGTH, cfg); lat_host.append(1e3 * (time.perf_counter() - t0))
    one.close()
    line["latency_ms_p50_single_frame"] = float(np.median(lat))
    line["config2_single_frame"] = {"frame": "seed 1, 1920x1080, H+V RECTIFY", "p50_ms_device_resident": float(np.median(lat)),
                                    "p90_ms_device_resident": float(np.percentile(lat, 90)),
                                    "p50_ms_from_pageable_host_memory": float(np.median(lat_host[5:])),
                                    "stage_ms": {k: v / 10 for k, v in p1["ms"].items()}, "kernels_per_frame": p1["kernels"] / 10}

    if rank0_extra:
        line.update(extra_legs(api, Runner, dev, local, frames_dev, out, world))
    runner.close()
    if world > 1:
        dist.destroy_process_group()
    return line if rank == 0 else None


def extra_legs(api, Runner, dev, local, frames_dev, out, world):
    """configs 1, 4, 5, the drop-in symbols, the CPU baselines and the parity sample (rank 0; outside every timed region)"""
    import torch
    ex = {}
    cfg = api.make_config()
    B = frames_dev.shape[0]

    def throughput(frames, ml, cfgs, engines, slots, reps=3):
        r = Runner(local, engines, slots)
        r.rectify(frames, ml, cfgs)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        tf, info = r.rectify(frames, ml, cfgs, steps=reps)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
        e = r.engines[0]; lat = []
        for i in range(6):
            t0 = time.perf_counter(); e.rectify(frames[i % len(frames)], ml, cfgs); lat.append(1e3 * (time.perf_counter() - t0))
        r.close()
        return reps * len(frames) / dt, float(np.median(lat[1:])), info

    try:    # config 4: 3840x2160 dense clutter, min_length 8
        f4 = make_frames(range(3, 11), dev, 3840, 2160, n_lines=400)
        ips, lat, info = throughput(f4, 8.0, cfg, 4, 2)
        ex["config4_4k_clutter"] = {"images_per_s": ips, "latency_ms_p50": lat, "frames": 8, "min_length": 8.0,
                                    "mean_filtered_segments": float(np.mean(info["n_lines"])),
                                    "scoring_pairs_per_frame": float(np.mean(info["n_obs"].sum(axis=1)) * 10000), "runner": "4 engines x 2 slots"}
        del f4
    except Exception as e:
        ex["config4_4k_clutter"] = {"error": repr(e)[:200]}
    try:    # config 5: 4000x3000, rotate-only and mixed strategies cycled per frame
        f5 = make_frames(range(60, 68), dev, 4000, 3000)
        strategies = [(api.ROTATE_H, api.ROTATE_H), (api.ROTATE_V, api.ROTATE_V), (api.RECTIFY, api.ROTATE_V), (api.KEEP, api.RECTIFY)]
        cfgs = np.concatenate([api.make_config(h_strategy=hs, v_strategy=vs, horizontal_vp_min_distance=2.0) for hs, vs in strategies])
        ips, lat, info = throughput(f5, 40.0, cfgs, 4, 2)
        ex["config5_12mp_mixed"] = {"images_per_s": ips, "latency_ms_p50": lat, "frames": 8, "min_length": 40.0,
                                    "mean_filtered_segments": float(np.mean(info["n_lines"])), "runner": "4 engines x 2 slots"}
        del f5
    except Exception as e:
        ex["config5_12mp_mixed"] = {"error": repr(e)[:200]}
    torch.cuda.empty_cache()

    # the drop-in symbols themselves: find_line_segment_groups + compute_rectification_transform per frame, pageable host memory
    try:
        from concurrent.futures import ThreadPoolExecutor
        n = min(B, 64)
        pageable = frames_dev[:n].cpu().numpy()

        def call(f):
            segs = api.find_line_segment_groups(f, MIN_LENGTH)
            if segs is not None:
                api.compute_rectification_transform(segs, W, H, cfg)
        call(pageable[0])
        res = {}
        for nt in (1, 8):
            with ThreadPoolExecutor(max_workers=nt) as tp:
                list(tp.map(call, pageable[:nt]))                 # engines of the pool come up
                t0 = time.perf_counter(); list(tp.map(call, pageable)); dt = time.perf_counter() - t0
            res[f"{nt}_caller_threads"] = {"images_per_s": n / dt, "ms_per_call": 1e3 * dt * nt / n}
        res["sample"] = f"{n} frames from pageable memory, one call per frame (reference src/interface.cpp:35-80, 122-208)"
        ex["dropin"] = res
    except Exception as e:
        ex["dropin"] = {"error": repr(e)[:200]}

    # config 1: the bundled doc/image.jpg (decoded fixture tests/golden/doc_image.npz), CPU reference as BASELINE says + this repo
    try:
        from oracle.bind import RectificationConfig
        g = dict(np.load(os.path.join(ROOT, "tests", "golden", "doc_image.npz"), allow_pickle=False))
        img = g["gray_u8"].astype(np.float32) / np.float32(256.0)
        ml = float(g["min_length"]); h1, w1 = img.shape
        impl, kind = _ref_impl()
        cores = os.cpu_count() or 1
        c1 = RectificationConfig(40.0, 1.5, 2, 2.0, 2)
        t_ref = []
        for i in range(4):
            t0 = time.perf_counter()
            s = impl.find_line_segment_groups(img, ml, num_threads=cores if kind == "reference" else -1, seed=0)
            impl.compute_rectification_transform(s, w1, h1, c1)
            t_ref.append(1e3 * (time.perf_counter() - t0))
        cfg1 = api.make_config(horizontal_vp_min_distance=2.0)
        t_gpu = []
        for i in range(12):
            t0 = time.perf_counter()
            s = api.find_line_segment_groups(img, ml); api.compute_rectification_transform(s, w1, h1, cfg1)
            t_gpu.append(1e3 * (time.perf_counter() - t0))
        ex["config1_doc_image"] = {"size": f"{w1}x{h1}", "min_length": ml, "cpu_reference_ms_p50": float(np.median(t_ref[1:])), "cpu_kind": kind,
                                   "cpu_threads": cores, "this_repo_dropin_ms_p50": float(np.median(t_gpu[2:])),
                                   "note": "BASELINE config 1 is the reference's CPU case; the GPU figure is the drop-in call from pageable memory"}
    except Exception as e:
        ex["config1_doc_image"] = {"error": repr(e)[:200]}

    host_frames = frames_dev[:48].cpu().numpy()
    if world == 1:
        ex["cpu_baseline"] = cpu_baselines(list(host_frames))
    # parity sample: the first 32 frames of the batch through the CPU oracle, compared with what the timed path produced
    try:
        from oracle.population import compare, oracle_results
        n = min(32, B)
        eng = api.Engine(local, 16)
        lines = []
        for i0 in range(0, n, 16):
            lines += eng.detect(frames_dev[i0:i0 + 16], MIN_LENGTH)[1]
        eng.close()
        rep = compare(lines, out["tf"][:n], oracle_results(host_frames[:n], MIN_LENGTH))
        rep["sample"] = f"seeds 0..{n - 1} of the batch; oracle = oracle/rectify_oracle.cpp (CPU restatement, pinned against the compiled reference), outside the timed region"
        ex["parity_sample"] = rep
    except Exception as e:
        ex["parity_sample"] = {"error": repr(e)[:200]}
    return ex


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=4096, help="frames of the batch (BASELINE config 3: 4096), sharded over the ranks")
    ap.add_argument("--engines", type=int, default=4, help="engines (streams + workspaces) per GPU")
    ap.add_argument("--slots", type=int, default=64, help="frame slots per engine (frames per launch group)")
    ap.add_argument("--ref-frames", type=int, default=4, help="frames per step of the reference arm (bounded CPU sample)")
    ap.add_argument("--quick", action="store_true", help="skip the extra legs (configs 1/4/5, drop-in, CPU baselines, parity sample)")
    args = ap.parse_args()
    # torchrun exports OMP_NUM_THREADS=1, which would throttle the reference's OpenMP path (threading.h:24-31 clamps to
    # omp_get_max_threads()): the CPU arms use all host cores, so say so before any OpenMP runtime starts
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    # the contract is ONE JSON line on stdout: libraries that print there (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        line = run_reference(args) if args.impl == "reference" else run_b200(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
