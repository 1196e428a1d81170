"""oracle/ — CPU restatement of the reference's stereo hot path. TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package; the product (``depthestimation_b200``) never does.

* ``sgbm_oracle.c``  — cv2.StereoSGBM.compute as called at depthlib/stereo_core.py:63-75,231
* ``imgops_oracle.c`` — remap / gray / median / box / speckle / depth ops of rectify.py:185-186,
  postprocess.py:6-171 and stereo_core.py:234-272
* ``cv2_reference.py`` — the live cv2 binary called exactly as the reference calls it (pins the C
  restatement; also the timed CPU baseline when cv2 is importable)

Pinned: the C restatement is checked bit-exact against the live cv2 4.13.0 binary
(tests/test_oracle_vs_cv2.py, runs wherever cv2 imports) and against committed golden vectors
generated from the *reference's own* ``depthlib.stereo_core.StereoCore`` imported from
/root/reference (tools/make_golden.py -> tests/golden/*.npz).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "liboracle.so"
_SOURCES = ["sgbm_oracle.c", "imgops_oracle.c"]

MODES = {"sgbm": 0, "hh": 1, "sgbm_3way": 2, "hh4": 3}


def build(force: bool = False) -> Path:
    """Compile the C restatement with gcc into oracle/_build/liboracle.so."""
    srcs = [_HERE / s for s in _SOURCES]
    if not force and _LIB_PATH.exists() and all(
            _LIB_PATH.stat().st_mtime >= s.stat().st_mtime for s in srcs):
        return _LIB_PATH
    _LIB_PATH.parent.mkdir(exist_ok=True)
    cmd = ["gcc", "-O2", "-fno-fast-math", "-ffp-contract=off", "-Wall", "-shared", "-fPIC", "-o",
           str(_LIB_PATH)] + [str(s) for s in srcs] + ["-lm"]
    subprocess.run(cmd, check=True)
    return _LIB_PATH


class OrcSgbmParams(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "minD", "D", "bs", "P1", "P2", "disp12MaxDiff", "preFilterCap", "uniquenessRatio",
        "speckleWindowSize", "speckleRange", "mode")]


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(str(_LIB_PATH))
    return _lib


def _p(a: np.ndarray, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct))


def make_params(min_disp=0, num_disp=128, block_size=5, P1=None, P2=None, disp12_max_diff=1,
                prefilter_cap=31, uniqueness_ratio=10, speckle_window_size=50, speckle_range=2,
                sgbm_mode="sgbm_3way") -> OrcSgbmParams:
    """Same derivation as depthlib/stereo_core.py:44-75 (P1 = 8 bs^2, P2 = 32 bs^2)."""
    if P1 is None:
        P1 = 8 * block_size ** 2
    if P2 is None:
        P2 = 32 * block_size ** 2
    mode = MODES[sgbm_mode] if isinstance(sgbm_mode, str) else int(sgbm_mode)
    return OrcSgbmParams(min_disp, num_disp, block_size, P1, P2, disp12_max_diff, prefilter_cap,
                         uniqueness_ratio, speckle_window_size, speckle_range, mode)


def geometry(p: OrcSgbmParams, W: int):
    w1, mx, inv = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    lib().orc_sgbm_geometry(ctypes.byref(p), W, ctypes.byref(w1), ctypes.byref(mx), ctypes.byref(inv))
    return w1.value, mx.value, inv.value


def sgbm_compute(left: np.ndarray, right: np.ndarray, p: OrcSgbmParams, stages: bool = False):
    """cv2.StereoSGBM.compute restated. Returns disp16 (int16 HxW) or, with stages=True,
    a dict with disp, C, S ([H][W1][D] int16) and raw (pre-median disparity)."""
    left = np.ascontiguousarray(left, dtype=np.uint8)
    right = np.ascontiguousarray(right, dtype=np.uint8)
    H, W = left.shape
    disp = np.empty((H, W), np.int16)
    C = S = raw = None
    cp = sp = rp = None
    if stages:
        W1, _, _ = geometry(p, W)
        if W1 > 0:
            C = np.empty((H, W1, p.D), np.int16)
            S = np.empty((H, W1, p.D), np.int16)
            cp, sp = _p(C, ctypes.c_int16), _p(S, ctypes.c_int16)
        raw = np.empty((H, W), np.int16)
        rp = _p(raw, ctypes.c_int16)
    lib().orc_sgbm_compute(_p(left, ctypes.c_uint8), _p(right, ctypes.c_uint8), H, W,
                           ctypes.byref(p), _p(disp, ctypes.c_int16), cp, sp, rp)
    if stages:
        return dict(disp=disp, C=C, S=S, raw=raw)
    return disp


def prefilter(img: np.ndarray, ftzero: int):
    img = np.ascontiguousarray(img, dtype=np.uint8)
    H, W = img.shape
    g = np.empty_like(img)
    r = np.empty_like(img)
    lib().orc_prefilter(_p(img, ctypes.c_uint8), H, W, ftzero, _p(g, ctypes.c_uint8), _p(r, ctypes.c_uint8))
    return g, r


def median3_s16(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.int16)
    out = np.empty_like(a)
    lib().orc_median3_s16(_p(a, ctypes.c_int16), a.shape[0], a.shape[1], _p(out, ctypes.c_int16))
    return out


def filter_speckles(a: np.ndarray, new_val: int, max_size: int, max_diff: int) -> np.ndarray:
    out = np.ascontiguousarray(a, dtype=np.int16).copy()
    lib().orc_filter_speckles(_p(out, ctypes.c_int16), out.shape[0], out.shape[1], int(new_val),
                              int(max_size), int(max_diff))
    return out


def remap_u8(src: np.ndarray, map1: np.ndarray, map2: np.ndarray) -> np.ndarray:
    src = np.ascontiguousarray(src, dtype=np.uint8)
    map1 = np.ascontiguousarray(map1, dtype=np.float32)
    map2 = np.ascontiguousarray(map2, dtype=np.float32)
    H, W = map1.shape
    out = np.empty((H, W), np.uint8)
    lib().orc_remap_u8(_p(src, ctypes.c_uint8), src.shape[0], src.shape[1], _p(map1, ctypes.c_float),
                       _p(map2, ctypes.c_float), H, W, _p(out, ctypes.c_uint8))
    return out


def bgr2gray(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img, dtype=np.uint8)
    H, W, _ = img.shape
    out = np.empty((H, W), np.uint8)
    lib().orc_bgr2gray(_p(img, ctypes.c_uint8), H * W, _p(out, ctypes.c_uint8))
    return out


def resize_area(img: np.ndarray, dsize) -> np.ndarray:
    """cv2.resize(img, (dW, dH), interpolation=cv2.INTER_AREA) for uint8 [H,W] or [H,W,C], shrinking only
    (depthlib/input.py:39-43)."""
    img = np.ascontiguousarray(img, np.uint8)
    dW, dH = int(dsize[0]), int(dsize[1])
    cn = 1 if img.ndim == 2 else img.shape[2]
    out = np.empty((dH, dW) if img.ndim == 2 else (dH, dW, cn), np.uint8)
    fn = lib().orc_resize_area_u8
    fn.argtypes = [ctypes.POINTER(ctypes.c_uint8), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                   ctypes.POINTER(ctypes.c_uint8), ctypes.c_int, ctypes.c_int]
    fn.restype = ctypes.c_int
    rc = fn(_p(img, ctypes.c_uint8), img.shape[0], img.shape[1], cn, _p(out, ctypes.c_uint8), dH, dW)
    if rc:
        raise ValueError("resize_area: only shrinking is restated (dsize must not exceed the source size)")
    return out


def resize_linear(img: np.ndarray, dsize) -> np.ndarray:
    """cv2.resize(img, (dW, dH), interpolation=cv2.INTER_LINEAR) for uint8 [H,W] or [H,W,C] (depthlib/rectify.py:92-105)."""
    img = np.ascontiguousarray(img, np.uint8)
    dW, dH = int(dsize[0]), int(dsize[1])
    cn = 1 if img.ndim == 2 else img.shape[2]
    out = np.empty((dH, dW) if img.ndim == 2 else (dH, dW, cn), np.uint8)
    fn = lib().orc_resize_linear_u8
    fn.argtypes = [ctypes.POINTER(ctypes.c_uint8), ctypes.c_int, ctypes.c_int, ctypes.c_int,
                   ctypes.POINTER(ctypes.c_uint8), ctypes.c_int, ctypes.c_int]
    fn.restype = ctypes.c_int
    if fn(_p(img, ctypes.c_uint8), img.shape[0], img.shape[1], cn, _p(out, ctypes.c_uint8), dH, dW):
        raise ValueError("resize_linear: bad arguments")
    return out


def fill_nearest(disp: np.ndarray, mask=None, kernel_size: int = 5) -> np.ndarray:
    """fill_holes(disparity, mask, method='nearest', kernel_size) of depthlib/postprocess.py:72-118 (mask None: disparity <= 0)."""
    d = np.ascontiguousarray(disp, dtype=np.float32).copy()
    m = np.ascontiguousarray((d <= 0) if mask is None else np.asarray(mask, bool), dtype=np.uint8)
    fn = lib().orc_fill_nearest_f32
    fn.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_uint8), ctypes.c_int, ctypes.c_int, ctypes.c_int]
    fn.restype = ctypes.c_int
    if fn(_p(d, ctypes.c_float), _p(m, ctypes.c_uint8), d.shape[0], d.shape[1], int(kernel_size)):
        raise ValueError("fill_nearest: kernel_size must be in 1..63")
    return d


def median3_f32(a: np.ndarray) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    out = np.empty_like(a)
    lib().orc_median3_f32(_p(a, ctypes.c_float), a.shape[0], a.shape[1], _p(out, ctypes.c_float))
    return out


def postprocess(disp: np.ndarray, max_speckle_size=50, max_diff=1.0, outlier_threshold=3.0,
                apply_outlier_removal=True) -> np.ndarray:
    """postprocess.postprocess_disparity with hole filling off (postprocess.py:120-171)."""
    a = np.ascontiguousarray(disp, dtype=np.float32)
    out = np.empty_like(a)
    lib().orc_postprocess.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_int, ctypes.c_int,
                                      ctypes.c_int, ctypes.c_int, ctypes.c_float, ctypes.c_int,
                                      ctypes.POINTER(ctypes.c_float)]
    lib().orc_postprocess(_p(a, ctypes.c_float), a.shape[0], a.shape[1], int(max_speckle_size),
                          int(max_diff * 16), float(outlier_threshold), int(bool(apply_outlier_removal)),
                          _p(out, ctypes.c_float))
    return out


def disp_to_depth(disp: np.ndarray, f_pixels: float, baseline_m: float, doffs: float = 0.0,
                  eps: float = 1e-6, max_depth=None) -> np.ndarray:
    """StereoCore.disparity_to_depth (stereo_core.py:234-272)."""
    a = np.ascontiguousarray(disp, dtype=np.float32)
    out = np.empty_like(a)
    fn = lib().orc_disp_to_depth
    fn.argtypes = [ctypes.POINTER(ctypes.c_float), ctypes.c_size_t, ctypes.c_float, ctypes.c_float,
                   ctypes.c_float, ctypes.c_float, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
    fB = np.float32(float(f_pixels) * float(baseline_m))
    fn(_p(a, ctypes.c_float), a.size, fB, np.float32(doffs), np.float32(eps),
       np.float32(max_depth if max_depth is not None else 0.0), int(max_depth is not None),
       _p(out, ctypes.c_float))
    return out
