"""Stage-by-stage comparison of the CUDA path with the oracle; prints everything instead of stopping at the first
mismatch (one GPU call should tell as much as possible).  Usage: python scripts/gpu_diag.py [case ...]"""
import os
import sys
import time
import traceback

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from librectify_b200 import api  # noqa: E402
from oracle.bind import Oracle  # noqa: E402
from tests.util import labels_equal_up_to_relabelling, seeds_equal_modulo_ties  # noqa: E402


def load_case(name):
    if name.startswith("synth:"):
        from librectify_b200.synth import manhattan_frame
        _, seed, w, h = name.split(":")
        img = manhattan_frame(int(seed), int(w), int(h)).numpy()
        return img, max(int(w), int(h)) / 100.0
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
    img = g["image"] if "image" in g else g["gray_u8"].astype(np.float32) / np.float32(256.0)
    return img, float(g["min_length"])


def cmp(name, a, b, exact=True):
    a = np.asarray(a); b = np.asarray(b)
    if a.shape != b.shape:
        print(f"  {name:14s} SHAPE {a.shape} vs {b.shape}")
        return False
    if a.dtype.names:
        ok = a.tobytes() == b.tobytes()
        nbad = sum(int(np.sum(a[f] != b[f])) for f in a.dtype.names)
        print(f"  {name:14s} {'OK ' if ok else 'BAD'} n={len(a)} differing fields={nbad}")
        if not ok:
            for f in a.dtype.names:
                d = np.nonzero(a[f] != b[f])[0]
                if len(d):
                    print(f"      {f}: {len(d)} differ, first idx {d[:5]}, gpu {a[f][d[:3]]}, ref {b[f][d[:3]]}, max abs {np.max(np.abs(a[f][d].astype(np.float64) - b[f][d])):.3g}")
        return ok
    ne = a != b
    n = int(ne.sum())
    extra = ""
    if n and a.dtype.kind == "f":
        extra = f" max abs diff {np.max(np.abs(a.astype(np.float64) - b)):.3g}"
    if n:
        idx = np.argwhere(ne)[:4]
        extra += f" first {idx.tolist()} gpu {a[tuple(idx[0])]} ref {b[tuple(idx[0])]}"
    print(f"  {name:14s} {'OK ' if n == 0 else 'BAD'} mismatches={n}/{a.size}{extra}")
    return n == 0


def run_case(eng, orc, name):
    img, min_length = load_case(name)
    h, w = img.shape
    print(f"== {name} {w}x{h} min_length={min_length}")
    t0 = time.time()
    info, lines = eng.detect(img, min_length)
    t1 = time.time()
    info = info[0]; lines = lines[0]
    print("  info", {k: (info[k].tolist()) for k in info.dtype.names}, f"wall {1e3 * (t1 - t0):.1f} ms")
    st = orc.detect_stages(img, want_grad8=True)
    ok = True
    for k in ("dx", "dy", "mag"):
        ok &= cmp(k, eng.fetch(k), st[k])
    ok &= cmp("bin", eng.fetch("bin").astype(np.int32), st["bin"])
    mask_ref = np.zeros((h, w), np.uint8)
    for b in range(8):
        mask_ref |= ((st["grad8"][b] > 0) | False).astype(np.uint8) << b
    # grad8 > 0 is a subset of the dilated mask (energy can be 0 inside the mask): compare as implication + where dx/dy != 0
    m8 = eng.fetch("mask8")
    sub = (mask_ref & ~m8) != 0
    print(f"  mask8          {'OK ' if not sub.any() else 'BAD'} ref-bits missing on gpu: {int(sub.sum())}")
    seeds = eng.fetch("seeds")
    s_ok = seeds_equal_modulo_ties(seeds, st["seeds"], st["mag"])
    print(f"  seeds          {'OK ' if s_ok else 'BAD'} gpu {len(seeds)} ref {len(st['seeds'])} exact-order={np.array_equal(seeds, st['seeds'])}")
    ok &= s_ok
    labels = eng.fetch("labels")
    l_ok = labels_equal_up_to_relabelling(labels, st["labels"])
    print(f"  labels         {'OK ' if l_ok else 'BAD'} px gpu {int((labels >= 0).sum())} ref {int((st['labels'] >= 0).sum())} exact={np.array_equal(labels, st['labels'])}")
    ok &= l_ok
    cs, rs = eng.fetch("comp_sizes"), eng.fetch("raw_segs")
    if not np.array_equal(seeds, st["seeds"]) and len(rs) == len(st["segs"]):
        # seeds of exactly equal magnitude may be processed in either order (std::sort is unstable): compare as multisets
        ka = np.lexsort((rs["y1"], rs["x1"])); kb = np.lexsort((st["segs"]["y1"], st["segs"]["x1"]))
        ok &= cmp("comp_sizes(set)", cs[ka], st["comp_sizes"][kb])
        ok &= cmp("raw_segs(set)", rs[ka], st["segs"][kb])
    else:
        ok &= cmp("comp_sizes", cs, st["comp_sizes"])
        ok &= cmp("raw_segs", rs, st["segs"])
    ref = orc.find_line_segment_groups(img, min_length, seed=0)
    if ref is None:
        print("  oracle returns None; gpu n_lines", info["n_lines"])
        return ok and info["n_lines"] == 0
    geo = [f for f in api.SEG_DTYPE.names if f != "group_id"]
    if len(ref) == len(lines):
        ok &= cmp("lines(geom)", lines[geo], ref[geo])
        m = orc.pencil_model(ref)
        ok &= cmp("model_h", eng.fetch("model_h"), m["h"])
        ok &= cmp("model_anchor", eng.fetch("model_anchor"), m["anchor"])
        ok &= cmp("model_dir", eng.fetch("model_dir"), m["dir"])
        ok &= cmp("model_len", eng.fetch("model_len"), m["len"])
        bb = eng.fetch("model_bbox")
        ok &= cmp("model_bbox", bb, np.concatenate([m["bb"], m["center"], [m["scale"]]]).astype(np.float32))
        # RANSAC trace
        import ctypes as C
        T = api.RANSAC_ITER
        n = len(ref)
        samples = np.zeros((4, T, 2), np.int32); scores = np.zeros((4, T), np.float32); n_obs = np.zeros(4, np.int32)
        winner = np.zeros(4, np.int32); best_h = np.zeros((4, 3), np.float32); refined = np.zeros((4, 3), np.float32)
        groups = np.zeros(n, np.int32)
        orc.set_seed(0)
        ks = orc.lib.orc_ransac_trace(C.c_void_p(ref.ctypes.data), n, T, samples.ctypes.data_as(C.c_void_p), scores.ctypes.data_as(C.c_void_p),
                                      n_obs.ctypes.data_as(C.c_void_p), winner.ctypes.data_as(C.c_void_p), best_h.ctypes.data_as(C.c_void_p),
                                      refined.ctypes.data_as(C.c_void_p), groups.ctypes.data_as(C.c_void_p))
        print(f"  oracle solves {ks} n_obs {n_obs.tolist()} winner {winner.tolist()}")
        tr = eng.ransac_trace()
        print(f"  gpu    solves {info['n_solves']} n_obs {info['n_obs'].tolist()} winner {tr['winner'].tolist()}")
        ok &= cmp("n_obs", info["n_obs"][:ks], n_obs[:ks])
        ok &= cmp("winner", tr["winner"][:ks], winner[:ks])
        ok &= cmp("best_h", tr["best_h"][:ks], best_h[:ks])
        ok &= cmp("refined", tr["refined"][:ks], refined[:ks])
        gs = eng.fetch("scores")
        ok &= cmp("scores", gs[:ks], scores[:ks])
        ok &= cmp("pairs(last)", eng.fetch("pairs"), samples[ks - 1])
        ok &= cmp("group_id", lines["group_id"], ref["group_id"])
        ok &= cmp("lines(all)", lines, ref)
    else:
        print(f"  lines          BAD count gpu {len(lines)} ref {len(ref)}")
        ok = False
    return ok


def main():
    cases = sys.argv[1:] or ["synth_320x240", "synth_640x360", "doc_image", "synth:1:1920:1080"]
    orc = Oracle()
    eng = api.Engine(0, 1)
    # sampler known answers first
    for n, cnt in ((10, 5), (940, 10000), (1570, 10000)):
        p, d = eng.sample_pairs(n, cnt)
        rp, rd, _ = orc.choice_knuth(0, n, cnt)
        print(f"sampler n={n}: pairs {'OK' if np.array_equal(p, rp) else 'BAD'} draws gpu {d} ref {rd}; head {p[:3].tolist()}")
    strm = eng.fetch("stream", count=4)
    print("stream head", strm.tolist(), "expect floats", (strm / 2.0**32).tolist())
    res = {}
    for c in cases:
        try:
            res[c] = run_case(eng, orc, c)
        except Exception:
            traceback.print_exc()
            res[c] = False
    print("SUMMARY", res)
    prof_eng = api.Engine(0, 1)
    img, ml = load_case("synth:1:1920:1080")
    prof_eng.detect(img, ml)
    prof_eng.profile(True)
    for _ in range(5):
        prof_eng.detect(img, ml)
    print("profile (5 frames, ms total)", prof_eng.profile_read())


if __name__ == "__main__":
    main()
