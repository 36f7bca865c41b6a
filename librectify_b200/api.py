"""Python host-side mirror of the C ABI (include/librectify.h + include/librectify_b200.h) over ctypes.

Names and argument meaning follow the reference's C API (reference src/librectify.h:111-222): ``find_line_segment_groups``
takes a float32 grayscale image and returns the grouped segments, ``compute_rectification_transform`` turns them into the
warped image corners.  ``Engine`` exposes the batched entry points for frames resident in device memory (torch CUDA
tensors are passed by ``data_ptr()`` — torch is plumbing here, the work happens in librectify_b200.so).

The shared library is the product; this module never computes anything itself and raises if the library is missing or
no CUDA device is present (there is no CPU fallback).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "librectify_b200.so")

ROTATE_H, ROTATE_V, RECTIFY, KEEP = 0, 1, 2, 3
RANSAC_ITER = 10000

SEG_DTYPE = np.dtype([("x1", "<f4"), ("y1", "<f4"), ("x2", "<f4"), ("y2", "<f4"),
                      ("weight", "<f4"), ("err", "<f4"), ("group_id", "<i4")])
TF_DTYPE = np.dtype([("width", "<i4"), ("height", "<i4"), ("top_left", "<f4", 3), ("top_right", "<f4", 3),
                     ("bottom_left", "<f4", 3), ("bottom_right", "<f4", 3), ("horizontal_vp", "<f4", 3),
                     ("vertical_vp", "<f4", 3)])
CFG_DTYPE = np.dtype([("vertical_vp_angular_tolerance", "<f4"), ("vertical_vp_min_distance", "<f4"), ("v_strategy", "<i4"),
                      ("horizontal_vp_min_distance", "<f4"), ("h_strategy", "<i4")])
INFO_DTYPE = np.dtype([("n_lines", "<i4"), ("n_seeds", "<i4"), ("n_components", "<i4"), ("n_pixels", "<i4"),
                       ("flood_rounds", "<i4"), ("flood_fallback", "<i4"), ("n_solves", "<i4"), ("n_obs", "<i4", 4),
                       ("max_magnitude", "<f4"), ("floods_small", "<i4"), ("floods_mid", "<i4"), ("floods_escalated", "<i4"), ("n_tied", "<i4"), ("replay_steps", "<i4")])
assert SEG_DTYPE.itemsize == 28 and TF_DTYPE.itemsize == 80 and CFG_DTYPE.itemsize == 20 and INFO_DTYPE.itemsize == 68

STAGES = dict(dx=0, dy=1, mag=2, bin=3, mask8=4, seeds=5, labels=6, comp_sizes=7, raw_segs=8, model_h=9, model_anchor=10,
              model_dir=11, model_len=12, model_bbox=13, pairs=14, scores=15, ransac=16, stream=17, refined_segs=18)
TIMERS = ("ingest", "grad", "nms", "seeds", "flood", "fit", "model", "ransac", "readback", "score")
STAGE_TIMERS = TIMERS[:9]            # "score" is the scoring grid alone, a sub-interval of "ransac"


class LibrectifyError(RuntimeError):
    pass


_lib = None


def load_library(path: str | None = None):
    """dlopen the in-tree shared object (built by librectify_b200/build.py). No fallback if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise LibrectifyError(f"{p} not found: build it with `python -m librectify_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(p)
    vp, ip, fp, ll = C.c_void_p, C.c_int, C.c_float, C.c_longlong
    lib.lrb_last_error.restype = C.c_char_p
    lib.lrb_version.restype = C.c_char_p
    lib.lrb_last_status.restype = C.c_int
    lib.lrb_fetch_lines.argtypes = [vp, ip, vp, ip]
    lib.lrb_set_seed.argtypes = [C.c_uint]; lib.lrb_set_seed.restype = None
    lib.lrb_engine_set_refine.argtypes = [vp, ip]
    for name in ("lrb_detect_host_u8", "lrb_detect_device_u8"):
        getattr(lib, name).argtypes = [vp, vp, ip, ip, ip, ip, ll, ll, fp, vp, vp, ip]
    for name in ("lrb_rectify_host_u8", "lrb_rectify_device_u8"):
        getattr(lib, name).argtypes = [vp, vp, ip, ip, ip, ip, ll, ll, fp, vp, ip, vp, vp]
    lib.lrb_engine_create.argtypes = [ip, ip, C.POINTER(vp)]
    lib.lrb_engine_destroy.argtypes = [vp]; lib.lrb_engine_destroy.restype = None
    lib.lrb_engine_slots.argtypes = [vp]
    for name in ("lrb_detect_device", "lrb_detect_host"):
        getattr(lib, name).argtypes = [vp, vp, ip, ip, ip, ll, ll, fp, vp, vp, ip]
    for name in ("lrb_rectify_device", "lrb_rectify_host"):
        getattr(lib, name).argtypes = [vp, vp, ip, ip, ip, ll, ll, fp, vp, ip, vp, vp]
    lib.lrb_pool_create.argtypes = [vp, ip, ip, ip, C.POINTER(vp)]
    lib.lrb_pool_destroy.argtypes = [vp]; lib.lrb_pool_destroy.restype = None
    lib.lrb_pool_engines.argtypes = [vp]
    lib.lrb_pool_engine.argtypes = [vp, ip]; lib.lrb_pool_engine.restype = vp
    lib.lrb_pool_rectify_host.argtypes = [vp, vp, ip, ip, ip, ll, ll, fp, vp, ip, vp, vp, ip]
    lib.lrb_pool_rectify_host_u8.argtypes = [vp, vp, ip, ip, ip, ip, ll, ll, fp, vp, ip, vp, vp, ip]
    lib.lrb_pool_rectify_device.argtypes = [vp, vp, vp, ip, ip, ll, ll, fp, vp, ip, vp, vp, ip]
    lib.lrb_compute_transform.argtypes = [vp, ip, ip, ip, vp, vp]; lib.lrb_compute_transform.restype = None
    lib.lrb_compute_transform_from_vp.argtypes = [ip, ip, vp, vp, vp]; lib.lrb_compute_transform_from_vp.restype = None
    lib.lrb_fit_vanishing_point.argtypes = [vp, ip, ip, vp]; lib.lrb_fit_vanishing_point.restype = None
    lib.lrb_assign_to_group.argtypes = [vp, ip, vp, ip, fp]; lib.lrb_assign_to_group.restype = None
    lib.lrb_prescale_size.argtypes = [ip, ip, ip, C.POINTER(ip), C.POINTER(ip), C.POINTER(fp)]
    lib.lrb_prescale_area_f32.argtypes = [vp, vp, ip, ip, ip, ll, ip, vp, ip, ll]
    lib.lrb_homography_from_corners.argtypes = [vp, fp, vp, C.POINTER(ip), C.POINTER(ip)]
    lib.lrb_warp_perspective_u8.argtypes = [vp, vp, ip, ip, ip, ip, ll, vp, ip, vp, ip, ip, ip, ll]
    lib.lrb_autorectify_u8.argtypes = [vp, vp, ip, ip, ip, ll, fp, ip, vp, fp, vp, vp, ip, C.POINTER(ip), vp, C.c_size_t, C.POINTER(ip), C.POINTER(ip)]
    lib.lrb_direct_estimate.argtypes = [vp, vp, ip, vp, ip, vp, vp, vp]
    lib.lrb_prosac_estimate.argtypes = [vp, vp, ip, vp, ip, fp, vp, vp]
    lib.lrb_fit_vanishing_point_device.argtypes = [vp, vp, ip, ip, vp]
    lib.lrb_assign_to_group_device.argtypes = [vp, vp, ip, vp, ip, fp]
    lib.lrb_fit_vanishing_point_noalloc.argtypes = [ip, vp, ip, ip, vp, vp]
    lib.lrb_stage_fetch.argtypes = [vp, ip, ip, vp, C.c_size_t]; lib.lrb_stage_fetch.restype = ll
    lib.lrb_profile_enable.argtypes = [vp, ip]; lib.lrb_profile_enable.restype = None
    lib.lrb_profile_reset.argtypes = [vp]; lib.lrb_profile_reset.restype = None
    lib.lrb_profile_read.argtypes = [vp, vp, vp, vp]; lib.lrb_profile_read.restype = None
    lib.lrb_profile_read_bytes.argtypes = [vp, vp, vp]; lib.lrb_profile_read_bytes.restype = None
    lib.lrb_ubench_fp32.argtypes = [vp, vp, vp]
    lib.lrb_sort_replay.argtypes = [vp, vp, ip, ip, vp]; lib.lrb_sort_replay.restype = ll
    lib.lrb_sample_pairs.argtypes = [vp, ip, ip, vp]; lib.lrb_sample_pairs.restype = ll
    lib.lrb_sample_pairs_parallel.argtypes = [vp, ip, ip, vp]; lib.lrb_sample_pairs_parallel.restype = ll
    # the drop-in symbols (reference names)
    lib.find_line_segment_groups.restype = vp
    lib.find_line_segment_groups.argtypes = [vp, ip, ip, ip, fp, C.c_bool, ip, C.POINTER(ip)]
    lib.release_line_segments.argtypes = [C.POINTER(vp)]; lib.release_line_segments.restype = None
    if path is None:
        _lib = lib
    return lib


def _err(lib) -> str:
    return (lib.lrb_last_error() or b"").decode()


def make_config(vertical_vp_angular_tolerance=40.0, vertical_vp_min_distance=1.5, v_strategy=RECTIFY,
                horizontal_vp_min_distance=1.5, h_strategy=RECTIFY) -> np.ndarray:
    """RectificationConfig with the reference's defaults (reference src/librectify.h:137-150)."""
    c = np.zeros(1, CFG_DTYPE)
    c[0] = (vertical_vp_angular_tolerance, vertical_vp_min_distance, v_strategy, horizontal_vp_min_distance, h_strategy)
    return c


def set_seed(seed: int) -> None:
    load_library().lrb_set_seed(C.c_uint(seed))


# ---- the reference's functions ----------------------------------------------------------------------------------------
def find_line_segment_groups(image: np.ndarray, min_length: float, refine: bool = False, num_threads: int = -1,
                             stride: int | None = None, flipped: bool = False):
    """reference src/interface.cpp:35-80 through the drop-in symbol.  image: float32 [H, W] host array.
    Returns a SEG_DTYPE array, or None where the reference returns nullptr."""
    lib = load_library()
    assert image.dtype == np.float32 and image.ndim == 2
    h, w = image.shape
    if stride is None:
        assert image.flags.c_contiguous
        stride = w
    base = image.ctypes.data
    if flipped:
        base += 4 * stride * (h - 1); stride = -stride
    n = C.c_int(-1)
    ptr = lib.find_line_segment_groups(C.c_void_p(base), w, h, stride, C.c_float(min_length), bool(refine), int(num_threads), C.byref(n))
    if not ptr:
        if n.value != 0:
            raise LibrectifyError("null result with n_lines != 0")
        if lib.lrb_last_status() != 0:          # a failure, not the reference's "no segments" nullptr
            raise LibrectifyError(_err(lib))
        return None
    out = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_byte)), shape=(n.value * 28,)).view(SEG_DTYPE).copy()
    p = C.c_void_p(ptr)
    lib.release_line_segments(C.byref(p))
    assert not p.value
    return out


def compute_rectification_transform(lines: np.ndarray, width: int, height: int, cfg: np.ndarray | None = None) -> np.ndarray:
    """reference src/interface.cpp:122-208 (host).  Returns a TF_DTYPE scalar array."""
    lib = load_library()
    cfg = make_config() if cfg is None else cfg
    lines = np.ascontiguousarray(lines, SEG_DTYPE)
    out = np.zeros(1, TF_DTYPE)
    lib.lrb_compute_transform(lines.ctypes.data, len(lines), width, height, cfg.ctypes.data, out.ctypes.data)
    return out[0]


def compute_rectification_transform_from_vp(width, height, vp_h, vp_v) -> np.ndarray:
    lib = load_library()
    a = np.asarray(vp_h, np.float32); b = np.asarray(vp_v, np.float32)
    out = np.zeros(1, TF_DTYPE)
    lib.lrb_compute_transform_from_vp(width, height, a.ctypes.data, b.ctypes.data, out.ctypes.data)
    return out[0]


def fit_vanishing_point(lines: np.ndarray, group: int):
    lib = load_library()
    lines = np.ascontiguousarray(lines, SEG_DTYPE)
    out = np.zeros(3, np.float32)
    lib.lrb_fit_vanishing_point(lines.ctypes.data, len(lines), group, out.ctypes.data)
    return tuple(float(v) for v in out)


def assign_to_group(lines: np.ndarray, new_lines: np.ndarray, angular_tolerance: float) -> np.ndarray:
    lib = load_library()
    lines = np.ascontiguousarray(lines, SEG_DTYPE)
    new = np.ascontiguousarray(new_lines, SEG_DTYPE).copy()
    lib.lrb_assign_to_group(lines.ctypes.data, len(lines), new.ctypes.data, len(new), C.c_float(angular_tolerance))
    return new


def tform_vec(t) -> np.ndarray:
    """flat float64 vector in the golden-fixture layout (tests/util.py)."""
    return np.array([t["width"], t["height"], *t["top_left"], *t["top_right"], *t["bottom_left"], *t["bottom_right"],
                     *t["horizontal_vp"], *t["vertical_vp"]], dtype=np.float64)


# ---- the engine (batched C ABI) ------------------------------------------------------------------------------------------
class Engine:
    """One CUDA device + stream + workspace of ``slots`` frames (include/librectify_b200.h)."""

    def __init__(self, device: int = 0, slots: int = 1, _borrowed=None):
        self.lib = load_library()
        self.owned = _borrowed is None
        if self.owned:
            h = C.c_void_p()
            rc = self.lib.lrb_engine_create(device, slots, C.byref(h))
            if rc != 0:
                raise LibrectifyError(_err(self.lib))
        else:
            h = C.c_void_p(_borrowed)              # an engine that belongs to a Pool
        self.h = h
        self.device = device
        self.slots = slots
        self.w = self.h_ = 0

    def close(self):
        if getattr(self, "h", None):
            if self.owned:
                self.lib.lrb_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise LibrectifyError(_err(self.lib))

    def set_refine(self, on: bool = True):
        """the reference's ``refine`` flag (src/interface.cpp:56-59) for the following detect / rectify calls"""
        self._check(self.lib.lrb_engine_set_refine(self.h, 1 if on else 0))

    @staticmethod
    def _frames_args(frames):
        """(pointer, n, h, w, pitch, frame_stride, on_device, keepalive) for a numpy array or a torch tensor [n,h,w] / [h,w]."""
        if isinstance(frames, np.ndarray):
            a = frames if frames.ndim == 3 else frames[None]
            assert a.dtype == np.float32 and a.strides[2] == 4 and a.strides[1] % 4 == 0
            fs = a.strides[0] // 4 if a.shape[0] > 1 else a.shape[1] * (a.strides[1] // 4)
            return a.ctypes.data, a.shape[0], a.shape[1], a.shape[2], a.strides[1] // 4, fs, False, a
        import torch
        t = frames if frames.dim() == 3 else frames[None]
        assert t.dtype == torch.float32 and t.stride(2) == 1
        fs = t.stride(0) if t.shape[0] > 1 else t.shape[1] * t.stride(1)
        return t.data_ptr(), t.shape[0], t.shape[1], t.shape[2], t.stride(1), fs, t.is_cuda, t

    def detect(self, frames, min_length: float, lines_cap: int = 4096):
        """find_line_segment_groups over a batch. Returns (info[n] INFO_DTYPE, list of SEG_DTYPE arrays)."""
        ptr, n, h, w, pitch, fs, dev, keep = self._frames_args(frames)
        info = np.zeros(n, INFO_DTYPE)
        lines = np.zeros((n, lines_cap), SEG_DTYPE)
        fn = self.lib.lrb_detect_device if dev else self.lib.lrb_detect_host
        self._check(fn(self.h, ptr, n, w, h, pitch, fs, C.c_float(min_length), info.ctypes.data, lines.ctypes.data, lines_cap))
        self.w, self.h_ = w, h
        if int(info["n_lines"].max(initial=0)) > lines_cap:
            if n > self.slots:                       # only the last launch group is still on the device
                return self.detect(frames, min_length, int(info["n_lines"].max()))
            out = []
            for i in range(n):                       # long lists come straight from device memory (no second pass)
                k = int(info["n_lines"][i])
                if k <= lines_cap:
                    out.append(lines[i, :k].copy()); continue
                buf = np.zeros(k, SEG_DTYPE)
                got = self.lib.lrb_fetch_lines(self.h, i, buf.ctypes.data, k)
                if got != k:
                    raise LibrectifyError(_err(self.lib))
                out.append(buf)
            return info, out
        return info, [lines[i, :info["n_lines"][i]].copy() for i in range(n)]

    def rectify(self, frames, min_length: float, cfg: np.ndarray | None = None):
        """detect + compute_rectification_transform per frame. Returns (transforms[n] TF_DTYPE, info[n])."""
        ptr, n, h, w, pitch, fs, dev, keep = self._frames_args(frames)
        cfg = make_config() if cfg is None else np.ascontiguousarray(cfg, CFG_DTYPE)
        info = np.zeros(n, INFO_DTYPE)
        tf = np.zeros(n, TF_DTYPE)
        fn = self.lib.lrb_rectify_device if dev else self.lib.lrb_rectify_host
        self._check(fn(self.h, ptr, n, w, h, pitch, fs, C.c_float(min_length), cfg.ctypes.data, len(cfg), tf.ctypes.data, info.ctypes.data))
        self.w, self.h_ = w, h
        return tf, info

    # 8-bit frames (gray [n,h,w] or BGR [n,h,w,3]; numpy = host, torch CUDA tensor = device) ----------------------------
    @staticmethod
    def _frames_args_u8(frames):
        if isinstance(frames, np.ndarray):
            a = frames
            if a.ndim == 2 or (a.ndim == 3 and a.shape[-1] == 3 and a.strides[-1] == 1 and a.strides[-2] == 3):
                a = a[None]
            ch = 3 if a.ndim == 4 else 1
            assert a.dtype == np.uint8 and a.strides[-1] == 1 and (ch == 1 or a.strides[2] == 3)
            fs = a.strides[0] if a.shape[0] > 1 else a.shape[1] * a.strides[1]
            return a.ctypes.data, a.shape[0], a.shape[1], a.shape[2], ch, a.strides[1], fs, False, a
        import torch
        t = frames
        ch = 3 if t.dim() == 4 else 1
        assert t.dtype == torch.uint8 and t.is_contiguous() and t.dim() in (3, 4)
        return t.data_ptr(), t.shape[0], t.shape[1], t.shape[2], ch, t.stride(1), t.stride(0), t.is_cuda, t

    def detect_u8(self, frames, min_length: float, lines_cap: int = 4096):
        ptr, n, h, w, ch, pitch, fs, dev, keep = self._frames_args_u8(frames)
        info = np.zeros(n, INFO_DTYPE)
        lines = np.zeros((n, lines_cap), SEG_DTYPE)
        fn = self.lib.lrb_detect_device_u8 if dev else self.lib.lrb_detect_host_u8
        self._check(fn(self.h, ptr, n, w, h, ch, pitch, fs, C.c_float(min_length), info.ctypes.data, lines.ctypes.data, lines_cap))
        self.w, self.h_ = w, h
        if int(info["n_lines"].max(initial=0)) > lines_cap:
            return self.detect_u8(frames, min_length, int(info["n_lines"].max()))
        return info, [lines[i, :info["n_lines"][i]].copy() for i in range(n)]

    def rectify_u8(self, frames, min_length: float, cfg: np.ndarray | None = None):
        ptr, n, h, w, ch, pitch, fs, dev, keep = self._frames_args_u8(frames)
        cfg = make_config() if cfg is None else np.ascontiguousarray(cfg, CFG_DTYPE)
        info = np.zeros(n, INFO_DTYPE)
        tf = np.zeros(n, TF_DTYPE)
        fn = self.lib.lrb_rectify_device_u8 if dev else self.lib.lrb_rectify_host_u8
        self._check(fn(self.h, ptr, n, w, h, ch, pitch, fs, C.c_float(min_length), cfg.ctypes.data, len(cfg), tf.ctypes.data, info.ctypes.data))
        self.w, self.h_ = w, h
        return tf, info

    # the demo's image operations either side of the path (reference src/autorectify.cpp) ---------------
    def prescale(self, image: np.ndarray, max_size: int):
        """prescale_image (src/autorectify.cpp:56-67): cv::resize(INTER_AREA) of a float frame; returns (frame, scale)."""
        a = np.ascontiguousarray(image, np.float32)
        h, w = a.shape
        dw, dh, sc = C.c_int(), C.c_int(), C.c_float()
        self._check(self.lib.lrb_prescale_size(w, h, int(max_size), C.byref(dw), C.byref(dh), C.byref(sc)))
        out = np.empty((dh.value, dw.value), np.float32)
        self._check(self.lib.lrb_prescale_area_f32(self.h, a.ctypes.data, 0, w, h, w, int(max_size), out.ctypes.data, 0, dw.value))
        return out, np.float32(sc.value)

    def homography_from_corners(self, tf, clip: float = 3.0):
        """homography_from_corners (src/autorectify.cpp:152-192): (H 3x3 float64, (dst_width, dst_height))."""
        t = np.ascontiguousarray(tf, TF_DTYPE).reshape(-1)[:1]
        H = np.zeros(9, np.float64)
        dw, dh = C.c_int(), C.c_int()
        self._check(self.lib.lrb_homography_from_corners(t.ctypes.data, C.c_float(clip), H.ctypes.data, C.byref(dw), C.byref(dh)))
        return H.reshape(3, 3), (dw.value, dh.value)

    def warp_perspective(self, image: np.ndarray, H: np.ndarray, dsize, inverse_map: bool = False) -> np.ndarray:
        """cv::warpPerspective(image, H, dsize) (INTER_LINEAR, constant 0 border) of an 8-bit gray or BGR image on the device."""
        a = np.ascontiguousarray(image, np.uint8)
        ch = 1 if a.ndim == 2 else a.shape[2]
        h, w = a.shape[:2]
        dw, dh = int(dsize[0]), int(dsize[1])
        Hc = np.ascontiguousarray(H, np.float64).reshape(9)
        out = np.empty((dh, dw) if a.ndim == 2 else (dh, dw, ch), np.uint8)
        self._check(self.lib.lrb_warp_perspective_u8(self.h, a.ctypes.data, 0, w, h, ch, w * ch, Hc.ctypes.data, 1 if inverse_map else 0,
                                                     out.ctypes.data, 0, dw, dh, dw * ch))
        return out

    def autorectify(self, image: np.ndarray, max_size: float = 1200, refine: bool = False, cfg: np.ndarray | None = None, clip: float = 3.0):
        """The flow of the reference's demo for one 8-bit gray / BGR image (src/autorectify.cpp:113-148,330-360), all image
        work on the device (lrb_autorectify_u8).  Returns (warped image or None, transform, lines)."""
        a = np.ascontiguousarray(image, np.uint8)
        ch = 1 if a.ndim == 2 else a.shape[2]
        h, w = a.shape[:2]
        cfg = make_config(horizontal_vp_min_distance=2.0) if cfg is None else np.ascontiguousarray(cfg, CFG_DTYPE)
        tf = np.zeros(1, TF_DTYPE)
        cap = 8192
        lines = np.zeros(cap, SEG_DTYPE)
        n, ww, wh = C.c_int(), C.c_int(), C.c_int()
        buf = np.empty(int(w * clip) * int(h * clip) * ch + 16, np.uint8)
        self._check(self.lib.lrb_autorectify_u8(self.h, a.ctypes.data, w, h, ch, w * ch, C.c_float(max_size), 1 if refine else 0,
                                                cfg.ctypes.data, C.c_float(clip), tf.ctypes.data, lines.ctypes.data, cap, C.byref(n),
                                                buf.ctypes.data, buf.size, C.byref(ww), C.byref(wh)))
        warped = None
        if ww.value > 0 and wh.value > 0:
            shape = (wh.value, ww.value) if a.ndim == 2 else (wh.value, ww.value, ch)
            warped = buf[: wh.value * ww.value * ch].reshape(shape).copy()
        return warped, tf[0], lines[: min(n.value, cap)].copy()

    def prosac_estimate(self, lines: np.ndarray, idx: np.ndarray | None, tol: float):
        """The reference's dormant PROSAC estimator (src/prosac.h:104-299) on the device. Returns (vanishing point[3], trace dict)."""
        ln = np.ascontiguousarray(lines, SEG_DTYPE)
        ix = None if idx is None else np.ascontiguousarray(idx, np.int32)
        vp = np.zeros(3, np.float32); tr = np.zeros(31, np.int32)
        self._check(self.lib.lrb_prosac_estimate(self.h, ln.ctypes.data, len(ln), None if ix is None else ix.ctypes.data,
                                                 0 if ix is None else len(ix), C.c_float(tol), vp.ctypes.data, tr.ctypes.data))
        return vp, dict(iterations=int(tr[0]), n_final=int(tr[1]), n_star=int(tr[2]), I_N_best=int(tr[3]),
                        samples=tr[4:22].reshape(9, 2).copy(), I_N=tr[22:31].copy())

    def fit_vanishing_point(self, lines: np.ndarray, group: int):
        """fit_vanishing_point (reference src/interface.cpp:211-215) computed on the device"""
        ln = np.ascontiguousarray(lines, SEG_DTYPE)
        out = np.zeros(3, np.float32)
        self._check(self.lib.lrb_fit_vanishing_point_device(self.h, ln.ctypes.data, len(ln), int(group), out.ctypes.data))
        return tuple(float(v) for v in out)

    def assign_to_group(self, lines: np.ndarray, new_lines: np.ndarray, angular_tolerance: float) -> np.ndarray:
        """assign_to_group (reference src/interface.cpp:218-265) computed on the device"""
        ln = np.ascontiguousarray(lines, SEG_DTYPE)
        new = np.ascontiguousarray(new_lines, SEG_DTYPE).copy()
        self._check(self.lib.lrb_assign_to_group_device(self.h, ln.ctypes.data, len(ln), new.ctypes.data, len(new), C.c_float(angular_tolerance)))
        return new

    def direct_estimate(self, lines: np.ndarray, idx: np.ndarray | None = None):
        """The reference's Hough-voted DirectEstimator (src/line_pencil.cpp:47-86, src/estimator.h:82-96).
        Returns (weights[len(idx)], (u, v) winning accumulator cell, vanishing point[3])."""
        ln = np.ascontiguousarray(lines, SEG_DTYPE)
        ix = None if idx is None else np.ascontiguousarray(idx, np.int32)
        m = len(ln) if ix is None else len(ix)
        w = np.zeros(m, np.float32); cell = np.zeros(2, np.int32); vp = np.zeros(3, np.float32)
        self._check(self.lib.lrb_direct_estimate(self.h, ln.ctypes.data, len(ln), None if ix is None else ix.ctypes.data, m,
                                                 w.ctypes.data, cell.ctypes.data, vp.ctypes.data))
        return w, (int(cell[0]), int(cell[1])), vp

    # stage window ------------------------------------------------------------------------------------------------------
    def fetch(self, stage: str, slot: int = 0, count: int | None = None) -> np.ndarray:
        w, h = self.w, self.h_
        kinds = dict(dx=(np.float32, (h, w)), dy=(np.float32, (h, w)), mag=(np.float32, (h, w)), bin=(np.uint8, (h, w)),
                     mask8=(np.uint8, (h, w)), labels=(np.int32, (h, w)), model_bbox=(np.float32, (7,)),
                     pairs=(np.int32, (RANSAC_ITER, 2)), scores=(np.float32, (4, RANSAC_ITER)), ransac=(np.int32, (28,)))
        if stage in kinds:
            dt, shape = kinds[stage]
            buf = np.zeros(shape, dt)
        elif stage == "stream":
            buf = np.zeros(count or 16, np.uint32)
        else:
            cap = count if count is not None else (w * h // 4 + 4096)
            dt = {"seeds": np.int32, "comp_sizes": np.int32, "raw_segs": SEG_DTYPE, "refined_segs": SEG_DTYPE}.get(stage, np.float32)
            mult = {"seeds": 2, "model_h": 3, "model_anchor": 2, "model_dir": 2}.get(stage, 1)
            buf = np.zeros(cap * mult, dt)
        got = self.lib.lrb_stage_fetch(self.h, slot, STAGES[stage], buf.ctypes.data, buf.nbytes)
        if got < 0:
            raise LibrectifyError(_err(self.lib))
        if stage in kinds:
            return buf
        out = buf[: got // buf.itemsize]
        if stage == "seeds":
            return out.reshape(-1, 2)
        if stage in ("model_h",):
            return out.reshape(-1, 3)
        if stage in ("model_anchor", "model_dir"):
            return out.reshape(-1, 2)
        return out

    def ransac_trace(self, slot: int = 0) -> dict:
        raw = self.fetch("ransac", slot)
        return dict(winner=raw[:4].copy(), best_h=raw[4:16].view(np.float32).reshape(4, 3).copy(),
                    refined=raw[16:28].view(np.float32).reshape(4, 3).copy())

    def sort_replay(self, vals: np.ndarray, depth_override: int = -1):
        """order in which libstdc++ std::sort (comparator: value descending) leaves ``vals`` — indices into vals; also the
        number of partition steps the device replayed (k3_replay.cu)."""
        v = np.ascontiguousarray(vals, np.float32)
        order = np.zeros(len(v), np.int32)
        steps = self.lib.lrb_sort_replay(self.h, v.ctypes.data, len(v), int(depth_override), order.ctypes.data)
        if steps < 0:
            raise LibrectifyError(_err(self.lib))
        return order, int(steps)

    def sample_pairs(self, n: int, count: int, parallel: bool = False):
        pairs = np.zeros((count, 2), np.int32)
        fn = self.lib.lrb_sample_pairs_parallel if parallel else self.lib.lrb_sample_pairs
        draws = fn(self.h, n, count, pairs.ctypes.data)
        if draws < 0:
            raise LibrectifyError(_err(self.lib))
        return pairs, int(draws)

    # profiling -----------------------------------------------------------------------------------------------------------
    def profile(self, on: bool = True):
        self.lib.lrb_profile_enable(self.h, int(on))
        self.lib.lrb_profile_reset(self.h)

    def profile_read(self) -> dict:
        ms = (C.c_double * len(TIMERS))()
        g, k = C.c_longlong(0), C.c_longlong(0)
        self.lib.lrb_profile_read(self.h, ms, C.byref(g), C.byref(k))
        h2d, d2h = C.c_longlong(0), C.c_longlong(0)
        self.lib.lrb_profile_read_bytes(self.h, C.byref(h2d), C.byref(d2h))
        return dict(ms={name: ms[i] for i, name in enumerate(TIMERS)}, groups=g.value, kernels=k.value,
                    h2d_bytes=h2d.value, d2h_bytes=d2h.value)

    def ubench_fp32(self) -> dict:
        """measured FP32-pipe peaks of this device in TFLOP/s: FFMA and FMA-free FMUL/FADD (csrc/ubench.cu)"""
        a, b = C.c_double(0), C.c_double(0)
        self._check(self.lib.lrb_ubench_fp32(self.h, C.byref(a), C.byref(b)))
        return dict(ffma=a.value, mul_add=b.value)


# ---- the multi-GPU batch entry ---------------------------------------------------------------------------------------------
class Pool:
    """lrb_pool_*: ``engines_per_device`` engines on each of ``devices``; a call shards its frames over devices and engines
    (frame i of B -> device floor(i * G / B)), one host thread per engine inside the library, results gathered in order."""

    def __init__(self, devices=(0,), engines_per_device: int = 4, slots: int = 32):
        self.lib = load_library()
        self.devices = [int(d) for d in devices]
        self.engines_per_device = engines_per_device
        self.slots = slots
        dv = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        if self.lib.lrb_pool_create(dv, len(self.devices), engines_per_device, slots, C.byref(h)) != 0:
            raise LibrectifyError(_err(self.lib))
        self.h = h
        n = self.lib.lrb_pool_engines(self.h)
        self.engines = [Engine(self.devices[i // engines_per_device], slots, _borrowed=self.lib.lrb_pool_engine(self.h, i)) for i in range(n)]

    def close(self):
        if getattr(self, "h", None):
            for e in self.engines:
                e.close()
            self.lib.lrb_pool_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise LibrectifyError(_err(self.lib))

    def rectify_host(self, frames: np.ndarray, min_length: float, cfg: np.ndarray | None = None, passes: int = 1):
        """frames: float32 [n,h,w] or uint8 [n,h,w] / [n,h,w,3] in host memory. Returns (transforms[n], info[n])."""
        cfg = make_config() if cfg is None else np.ascontiguousarray(cfg, CFG_DTYPE)
        if frames.dtype == np.uint8:
            ptr, n, h, w, ch, pitch, fs, _, keep = Engine._frames_args_u8(frames)
        else:
            ptr, n, h, w, pitch, fs, _, keep = Engine._frames_args(frames); ch = 0
        tf = np.zeros(n, TF_DTYPE); info = np.zeros(n, INFO_DTYPE)
        if ch:
            self._check(self.lib.lrb_pool_rectify_host_u8(self.h, ptr, n, w, h, ch, pitch, fs, C.c_float(min_length), cfg.ctypes.data, len(cfg),
                                                          tf.ctypes.data, info.ctypes.data, passes))
        else:
            self._check(self.lib.lrb_pool_rectify_host(self.h, ptr, n, w, h, pitch, fs, C.c_float(min_length), cfg.ctypes.data, len(cfg),
                                                       tf.ctypes.data, info.ctypes.data, passes))
        return tf, info

    def rectify_device(self, shards, min_length: float, cfg: np.ndarray | None = None, passes: int = 1):
        """shards: one torch CUDA tensor [n_d,h,w] per pool device (resident on that device), or a single tensor for a
        one-device pool. Returns (transforms, info) in shard order."""
        if not isinstance(shards, (list, tuple)):
            shards = [shards]
        assert len(shards) == len(self.devices)
        cfg = make_config() if cfg is None else np.ascontiguousarray(cfg, CFG_DTYPE)
        args = [Engine._frames_args(t) if t is not None and t.shape[0] > 0 else None for t in shards]
        first = next(a for a in args if a is not None)
        _, _, h, w, pitch, fs, _, _ = first
        ptrs = (C.c_void_p * len(shards))(*[a[0] if a else None for a in args])
        counts = (C.c_int * len(shards))(*[a[1] if a else 0 for a in args])
        for a in args:
            assert a is None or (a[6] and a[2:6] == first[2:6]), "all shards: CUDA tensors of one frame size and layout"
        n = sum(counts)
        tf = np.zeros(n, TF_DTYPE); info = np.zeros(n, INFO_DTYPE)
        self._check(self.lib.lrb_pool_rectify_device(self.h, ptrs, counts, w, h, pitch, fs, C.c_float(min_length), cfg.ctypes.data, len(cfg),
                                                     tf.ctypes.data, info.ctypes.data, passes))
        return tf, info
