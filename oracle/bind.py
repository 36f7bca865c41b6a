"""TEST INFRASTRUCTURE ONLY — ctypes bindings for the two CPU checkers:

* ``Ref``    : the compiled, unmodified reference (oracle/_ref/librectify_{canon,ship}.so, built by oracle/Makefile
               from /root/reference; see oracle/ref_harness.cpp for the stage window).
* ``Oracle`` : the plain-C++ restatement (oracle/rectify_oracle.cpp -> oracle/_build/librectify_oracle.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
The product package (librectify_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
BUILD_DIR = os.path.join(HERE, "_build")


class LineSegment(C.Structure):
    _fields_ = [("x1", C.c_float), ("y1", C.c_float), ("x2", C.c_float), ("y2", C.c_float),
                ("weight", C.c_float), ("err", C.c_float), ("group_id", C.c_int)]


class Point(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class ImageTransform(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("top_left", Point), ("top_right", Point),
                ("bottom_left", Point), ("bottom_right", Point), ("horizontal_vp", Point), ("vertical_vp", Point)]


class RectificationConfig(C.Structure):
    _fields_ = [("vertical_vp_angular_tolerance", C.c_float), ("vertical_vp_min_distance", C.c_float),
                ("v_strategy", C.c_int), ("horizontal_vp_min_distance", C.c_float), ("h_strategy", C.c_int)]

    @classmethod
    def default(cls):
        return cls(40.0, 1.5, 2, 1.5, 2)


ROTATE_H, ROTATE_V, RECTIFY, KEEP = 0, 1, 2, 3

SEG_DTYPE = np.dtype([("x1", "<f4"), ("y1", "<f4"), ("x2", "<f4"), ("y2", "<f4"),
                      ("weight", "<f4"), ("err", "<f4"), ("group_id", "<i4")])
assert SEG_DTYPE.itemsize == C.sizeof(LineSegment) == 28


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t)) if a is not None else None


def _fp(a):
    return _p(a, C.c_float)


def _ip(a):
    return _p(a, C.c_int)


def transform_to_dict(t: ImageTransform) -> dict:
    g = lambda p: (p.x, p.y, p.z)
    return dict(width=t.width, height=t.height, top_left=g(t.top_left), top_right=g(t.top_right),
                bottom_left=g(t.bottom_left), bottom_right=g(t.bottom_right),
                horizontal_vp=g(t.horizontal_vp), vertical_vp=g(t.vertical_vp))


def make_all(target="all"):
    subprocess.check_call(["make", "-s", "-C", HERE, "-j8", target])


class _StageLib:
    """Shared wrapper: both checkers export the same stage window, with prefix 'ref_' resp. 'orc_'."""

    prefix = ""

    def __init__(self, path):
        self.path = path
        self.lib = C.CDLL(path)
        self._declare()

    def _f(self, name):
        return getattr(self.lib, self.prefix + name)

    def _declare(self):
        f = self._f
        f("set_seed").argtypes = [C.c_uint]
        f("set_seed").restype = None
        f("choice_knuth").restype = C.c_longlong
        f("choice_knuth").argtypes = [C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_uint)]
        f("sizeof").restype = C.c_int
        self.lib.find_line_segment_groups.restype = C.POINTER(LineSegment)
        self.lib.find_line_segment_groups.argtypes = [C.POINTER(C.c_float), C.c_int, C.c_int, C.c_int, C.c_float,
                                                      C.c_bool, C.c_int, C.POINTER(C.c_int)]
        self.lib.release_line_segments.argtypes = [C.POINTER(C.POINTER(LineSegment))]
        self.lib.release_line_segments.restype = None

    # ---- public API -------------------------------------------------------------------------------------------
    def set_seed(self, s):
        self._f("set_seed")(C.c_uint(s))

    def find_line_segment_groups(self, img: np.ndarray, min_length: float, refine=False, num_threads=-1, seed=0,
                                 stride=None, flipped=False):
        """img: float32 [H, W] (C-contiguous unless stride given). Returns structured array (SEG_DTYPE) or None."""
        assert img.dtype == np.float32
        h, w = img.shape
        if stride is None:
            assert img.flags.c_contiguous
            stride = w
        self.set_seed(seed)
        n = C.c_int(-7)
        base = img.ctypes.data_as(C.POINTER(C.c_float))
        if flipped:  # pass pointer to LAST row and negative stride (reference src/librectify.h:15-18)
            base = C.cast(C.c_void_p(img.ctypes.data + 4 * stride * (h - 1)), C.POINTER(C.c_float))
            stride = -stride
        ptr = self.lib.find_line_segment_groups(base, w, h, stride, C.c_float(min_length), bool(refine),
                                                int(num_threads), C.byref(n))
        if not ptr:
            assert n.value == 0
            return None
        out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_byte)), shape=(n.value * 28,)).view(SEG_DTYPE).copy()
        pp = C.cast(ptr, C.POINTER(LineSegment))
        self.lib.release_line_segments(C.byref(pp))
        assert not pp
        return out

    def compute_rectification_transform(self, segs: np.ndarray, w: int, h: int, cfg: RectificationConfig = None):
        cfg = cfg or RectificationConfig.default()
        segs = np.ascontiguousarray(segs)
        out = ImageTransform()
        self._f("compute_rectification_transform")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), C.c_int(w),
                                                   C.c_int(h), C.byref(cfg), C.byref(out))
        return out

    def compute_rectification_transform_from_vp(self, w, h, vph, vpv):
        out = ImageTransform()
        a, b = Point(*vph), Point(*vpv)
        self._f("compute_rectification_transform_from_vp")(C.c_int(w), C.c_int(h), C.byref(a), C.byref(b),
                                                           C.byref(out))
        return out

    def fit_vanishing_point(self, segs, group):
        segs = np.ascontiguousarray(segs)
        out = Point()
        self._f("fit_vanishing_point")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), C.c_int(group), C.byref(out))
        return (out.x, out.y, out.z)

    def assign_to_group(self, segs, new_segs, tol_deg):
        segs = np.ascontiguousarray(segs)
        new_segs = np.ascontiguousarray(new_segs).copy()
        self.lib.assign_to_group(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), C.c_void_p(new_segs.ctypes.data),
                                 C.c_int(len(new_segs)), C.c_float(tol_deg))
        return new_segs

    # ---- stage window -----------------------------------------------------------------------------------------
    def gauss_deriv_kernel(self, dir_x: bool):
        out = np.zeros(25, np.float32)
        self._f("gauss_deriv_kernel")(C.c_int(int(dir_x)), _fp(out))
        return out.reshape(5, 5)

    def detect_stages(self, img: np.ndarray, num_threads=-1, want_grad8=True, cap_seeds=None, cap_comp=None):
        assert img.dtype == np.float32 and img.flags.c_contiguous
        h, w = img.shape
        cap_seeds = cap_seeds or (h * w // 4 + 16)
        cap_comp = cap_comp or (h * w // 6 + 16)
        dx = np.zeros((h, w), np.float32); dy = np.zeros_like(dx); mag = np.zeros_like(dx)
        bins = np.zeros((h, w), np.int32)
        grad8 = np.zeros((8, h, w), np.float32) if want_grad8 else None
        seeds = np.zeros((cap_seeds, 2), np.int32)
        labels = np.zeros((h, w), np.int32)
        sizes = np.zeros(cap_comp, np.int32)
        segs = np.zeros(cap_comp, SEG_DTYPE)
        ns, nc = C.c_int(0), C.c_int(0)
        self._f("detect_stages")(_fp(img), C.c_int(w), C.c_int(h), C.c_int(num_threads), _fp(dx), _fp(dy), _fp(mag),
                                 _ip(bins), _fp(grad8), _ip(seeds), C.c_int(cap_seeds), C.byref(ns), _ip(labels),
                                 _ip(sizes), C.c_void_p(segs.ctypes.data), C.c_int(cap_comp), C.byref(nc))
        assert ns.value <= cap_seeds and nc.value <= cap_comp
        return dict(dx=dx, dy=dy, mag=mag, bin=bins, grad8=grad8, seeds=seeds[:ns.value].copy(), labels=labels,
                    comp_sizes=sizes[:nc.value].copy(), segs=segs[:nc.value].copy())

    def pencil_model(self, segs):
        segs = np.ascontiguousarray(segs); n = len(segs)
        bb = np.zeros(4, np.float32); c = np.zeros(2, np.float32); s = C.c_float(0)
        h = np.zeros((n, 3), np.float32); a = np.zeros((n, 2), np.float32); d = np.zeros((n, 2), np.float32)
        ln = np.zeros(n, np.float32)
        self._f("pencil_model")(C.c_void_p(segs.ctypes.data), C.c_int(n), _fp(bb), _fp(c), C.byref(s), _fp(h), _fp(a),
                                _fp(d), _fp(ln))
        return dict(bb=bb, center=c, scale=np.float32(s.value), h=h, anchor=a, dir=d, len=ln)

    def score_pairs(self, segs, idx, pairs, tol):
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32)
        pairs = np.ascontiguousarray(pairs, np.int32); k = len(pairs)
        check = np.zeros(k, np.int32); hyp = np.zeros((k, 3), np.float32); score = np.zeros(k, np.float32)
        err = np.zeros(len(idx), np.float32)
        self._f("score_pairs")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), _ip(pairs),
                               C.c_int(k), C.c_float(tol), _ip(check), _fp(hyp), _fp(score), _fp(err))
        return dict(check=check, hyp=hyp, score=score, err_last=err)

    def errors(self, segs, idx, hyp):
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32)
        hyp = np.ascontiguousarray(hyp, np.float32); err = np.zeros(len(idx), np.float32)
        self._f("errors")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), _fp(hyp), _fp(err))
        return err

    def fit_optimal(self, segs, idx):
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32); out = np.zeros(3, np.float32)
        self._f("fit_optimal")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), _fp(out))
        return out

    def get_weights(self, segs, idx):
        """LinePencilModel::get_weights (reference src/line_pencil.cpp:47-86): Hough-voted inclination^4 per listed segment"""
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32); out = np.zeros(len(idx), np.float32)
        self._f("get_weights")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), _fp(out))
        return out

    def direct_solve(self, segs, idx):
        """DirectEstimator::solve (reference src/estimator.h:82-96)"""
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32); out = np.zeros(3, np.float32)
        self._f("direct_solve")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), _fp(out))
        return out

    def prosac_solve(self, segs, idx, tol, seed=0, trace=False):
        """PROSAC_Estimator<LinePencilModel>::solve (reference src/prosac.h:104-299) on lines ``idx`` of ``segs``"""
        segs = np.ascontiguousarray(segs); idx = np.ascontiguousarray(idx, np.int32)
        self.set_seed(seed)
        out = np.zeros(3, np.float32); tr = np.zeros(31, np.int32)
        self._f("prosac_solve")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), _ip(idx), C.c_int(len(idx)), C.c_float(tol), _fp(out), _ip(tr))
        if trace:
            return out, dict(iterations=int(tr[0]), n_final=int(tr[1]), n_star=int(tr[2]), I_N_best=int(tr[3]),
                             samples=tr[4:22].reshape(9, 2).copy(), I_N=tr[22:31].copy())
        return out

    def choice_knuth(self, seed, n, count):
        pairs = np.zeros((count, 2), np.int32); nxt = C.c_uint(0)
        draws = self._f("choice_knuth")(C.c_uint(seed), C.c_int(n), C.c_int(count), _ip(pairs), C.byref(nxt))
        return pairs, int(draws), int(nxt.value)

    def estimate_line_pencils(self, segs, seed=0, num_threads=-1):
        segs = np.ascontiguousarray(segs).copy()
        self.set_seed(seed)
        self._f("estimate_line_pencils")(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), C.c_int(num_threads))
        return segs

    def refine(self, segs):
        """postprocess_lines_segments (reference src/line_detector.cpp:340-444) on raw segments; returns the merged list"""
        segs = np.ascontiguousarray(segs, SEG_DTYPE)
        out = np.zeros(len(segs), SEG_DTYPE)
        f = self._f("refine"); f.restype = C.c_int
        n = f(C.c_void_p(segs.ctypes.data), C.c_int(len(segs)), C.c_void_p(out.ctypes.data))
        return out[:n].copy()

    def fit_line_parameters(self, X, w):
        X = np.ascontiguousarray(X, np.float32); w = np.ascontiguousarray(w, np.float32)
        out = np.zeros(1, SEG_DTYPE)
        self._f("fit_line_parameters")(_fp(X), _fp(w), C.c_int(len(w)), C.c_void_p(out.ctypes.data))
        return out[0]

    def sizeof(self, which):
        return self._f("sizeof")(C.c_int(which))


class Ref(_StageLib):
    prefix = "ref_"

    def __init__(self, flavour="canon"):
        path = os.path.join(REF_DIR, f"librectify_{flavour}.so")
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        super().__init__(path)
        if flavour == "canon":
            self.lib.ref_set_deterministic_malloc()


class Oracle(_StageLib):
    prefix = "orc_"

    def __init__(self):
        path = os.path.join(BUILD_DIR, "librectify_oracle.so")
        if not os.path.exists(path):
            make_all("oracle")
        super().__init__(path)

    def std_sort_order(self, vals, depth_override=-1):
        """order in which std::sort with find_peaks' comparator (reference src/filter.cpp:188) leaves ``vals``"""
        v = np.ascontiguousarray(vals, np.float32)
        order = np.zeros(len(v), np.int32)
        self.lib.orc_std_sort_order(_fp(v), C.c_int(len(v)), C.c_int(depth_override), _ip(order))
        return order


def ref_available(flavour="canon"):
    return os.path.exists(os.path.join(REF_DIR, f"librectify_{flavour}.so"))
