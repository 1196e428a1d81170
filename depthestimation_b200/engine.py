"""Python face of the C ABI: one ``StereoEngine`` = one ``B200sHandle`` (GPU + stream + scratch).

Plays the role ``cv2.StereoSGBM`` / ``cv2.remap`` / ``cv2.filterSpeckles`` ... play for the
reference's ``StereoCore`` (depthlib/stereo_core.py:44-75, 212-293): numpy in, numpy out, all
arithmetic on the GPU. Device-resident batch calls take integer device pointers (e.g.
``torch.Tensor.data_ptr()``) so that torch is only ever used for memory and streams.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib
from ._lib import B200sDepthParams, B200sFrameIO, B200sParams, B200sPostParams, MODES

__all__ = ["StereoEngine", "EngineError", "make_params", "make_post_params", "make_depth_params", "thread_engine",
           "pinned_empty"]


class EngineError(RuntimeError):
    pass


def make_params(min_disp=0, num_disp=128, block_size=5, P1=None, P2=None, disp12_max_diff=1, prefilter_cap=31,
                uniqueness_ratio=10, speckle_window_size=50, speckle_range=2, sgbm_mode="sgbm_3way") -> B200sParams:
    """Same derivation as StereoCore._build_sgbm (depthlib/stereo_core.py:44-75): P1 = 8 bs^2,
    P2 = 32 bs^2, unknown mode strings fall back to sgbm_3way."""
    if P1 is None:
        P1 = 8 * block_size ** 2
    if P2 is None:
        P2 = 32 * block_size ** 2
    mode = MODES.get(sgbm_mode, 2) if isinstance(sgbm_mode, str) else int(sgbm_mode)
    return B200sParams(int(min_disp), int(num_disp), int(block_size), int(P1), int(P2), int(disp12_max_diff),
                       int(prefilter_cap), int(uniqueness_ratio), int(speckle_window_size), int(speckle_range), mode)


def make_post_params(fast_mode=False, max_speckle_size=50, max_diff=1.0, outlier_threshold=3.0,
                     apply_outlier_removal=True) -> B200sPostParams:
    return B200sPostParams(int(bool(fast_mode)), int(max_speckle_size), int(max_diff * 16),
                           float(outlier_threshold), int(bool(apply_outlier_removal)))


def make_depth_params(f_pixels=None, baseline_m=None, doffs=0.0, eps=1e-6, max_depth=None) -> B200sDepthParams:
    if f_pixels is None or baseline_m is None:
        return B200sDepthParams(0, 0.0, 0.0, 0.0, 0.0, 0)
    fB = np.float32(float(f_pixels) * float(baseline_m))
    return B200sDepthParams(1, float(fB), float(np.float32(doffs)), float(np.float32(eps)),
                            float(np.float32(max_depth if max_depth is not None else 0.0)), int(max_depth is not None))


def _u8(a):
    a = np.asarray(a)
    if a.dtype != np.uint8:
        raise TypeError(f"expected uint8 image, got {a.dtype}")
    if a.ndim != 2:
        raise ValueError(f"expected a single-channel HxW image, got shape {a.shape}")
    if a.strides[1] != 1 or a.strides[0] < a.shape[1]:
        a = np.ascontiguousarray(a)
    return a


def _pair(left, right):
    left, right = _u8(left), _u8(right)
    if left.shape != right.shape:
        raise ValueError("left and right images must have the same size")
    if left.strides != right.strides:
        left, right = np.ascontiguousarray(left), np.ascontiguousarray(right)
    return left, right


def pinned_empty(shape, dtype) -> np.ndarray:
    """numpy array in page-locked host memory (b200s_host_alloc): the host side of the asynchronous batch calls.
    Freed when the array (and every view of it) is gone."""
    import weakref
    lib = _lib.load()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    ptr = C.c_void_p()
    if n == 0 or lib.b200s_host_alloc(n, C.byref(ptr)) != 0:
        raise EngineError(f"b200s_host_alloc({n}) failed: no usable CUDA device or out of page-locked memory")
    buf = (C.c_ubyte * n).from_address(ptr.value)
    weakref.finalize(buf, lib.b200s_host_free, C.c_void_p(ptr.value))
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


_tls = threading.local()


def thread_engine(device: int = 0) -> "StereoEngine":
    """One lazily created engine per (thread, device) for the stateless image ops (gray, INTER_AREA downscale) that
    the reference issues as free cv2 calls (input.py:39-43, threaded_stereo.py:41-46): a handle owns a stream and
    scratch, so it is not shared between threads."""
    cache = getattr(_tls, "engines", None)
    if cache is None:
        cache = _tls.engines = {}
    e = cache.get(device)
    if e is None:
        e = cache[device] = StereoEngine(device)
    return e


class StereoEngine:
    """Owns one B200sHandle on ``device``."""

    def __init__(self, device: int = 0, **sgbm_kw):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        self._inflight = {}
        self.params = make_params(**sgbm_kw)
        rc = self._lib.b200s_create(int(device), C.byref(self.params), C.byref(self._h))
        if rc == _lib.E_INVALID:
            raise ValueError("b200s_create: invalid matcher parameters "
                             "(num_disp in [1,256], odd block_size <= 11, mode in sgbm/hh/sgbm_3way/hh4)")
        if rc != 0:
            raise EngineError("b200s_create failed: no usable CUDA device (the engine has no CPU fallback)")
        self.device = int(device)
        self.maps_shape = None

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b200s_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc == 0:
            return
        msg = self._lib.b200s_last_error(self._h).decode("utf-8", "replace")
        if rc == _lib.E_INVALID:
            raise ValueError(f"{what}: {msg}")
        raise EngineError(f"{what}: {msg} (code {rc})")

    def configure(self, **sgbm_kw):
        p = make_params(**sgbm_kw)
        self._check(self._lib.b200s_configure(self._h, C.byref(p)), "b200s_configure")
        self.params = p

    def set_stream(self, cuda_stream):
        self._check(self._lib.b200s_set_stream(self._h, C.c_void_p(cuda_stream or 0)), "b200s_set_stream")

    def sync(self):
        self._check(self._lib.b200s_sync(self._h), "b200s_sync")

    def preferred_batch(self, H: int, W: int) -> int:
        """Frames per batch call that fill the GPU without a partial last wave (throughput hint)."""
        b = C.c_int()
        self._check(self._lib.b200s_preferred_batch(self._h, int(H), int(W), C.byref(b)), "b200s_preferred_batch")
        return int(b.value)

    def set_debug_flow(self, on=True):
        """Generic per-direction kernels that materialise the full S volume (for the S stage tap)."""
        self._lib.b200s_set_debug_flow(self._h, int(on))

    def enable_timing(self, on=True):
        self._lib.b200s_enable_timing(self._h, int(on))

    @property
    def last_ms(self) -> float:
        return float(self._lib.b200s_last_ms(self._h))

    @property
    def launch_count(self) -> int:
        return int(self._lib.b200s_launch_count(self._h))

    def geometry(self, W: int):
        w1, mx, dp, inv = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._lib.b200s_geometry(self._h, int(W), C.byref(w1), C.byref(mx), C.byref(dp), C.byref(inv))
        return dict(W1=w1.value, minX1=mx.value, Dp=dp.value, INV=inv.value)

    # -- host-buffer ops ----------------------------------------------------------------------
    def set_maps(self, map1_L, map2_L, map1_R, map2_R):
        ms = [np.ascontiguousarray(m, dtype=np.float32) for m in (map1_L, map2_L, map1_R, map2_R)]
        H, W = ms[0].shape
        if any(m.shape != (H, W) for m in ms):
            raise ValueError("rectification maps must share one HxW shape")
        self._check(self._lib.b200s_set_maps(self._h, *[m.ctypes.data for m in ms], H, W), "b200s_set_maps")
        self.maps_shape = (H, W)

    def build_maps(self, K, dist, R, P, size):
        """== cv2.initUndistortRectifyMap(K, dist, R, P, size, cv2.CV_32FC1) (rectify.py:68-73): (map1, map2) float32."""
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        d = np.zeros(0) if dist is None else np.ascontiguousarray(dist, dtype=np.float64).ravel()
        Rm = None if R is None else np.ascontiguousarray(R, dtype=np.float64).reshape(3, 3)
        Pm = None if P is None else np.ascontiguousarray(P, dtype=np.float64)
        if Pm is not None and Pm.shape == (3, 3):
            Pm = np.ascontiguousarray(np.hstack([Pm, np.zeros((3, 1))]))
        if Pm is not None and Pm.shape != (3, 4):
            raise ValueError("build_maps: P must be 3x3 or 3x4")
        W, H = int(size[0]), int(size[1])
        m1, m2 = np.empty((H, W), np.float32), np.empty((H, W), np.float32)
        self._check(self._lib.b200s_build_maps(
            self._h, K.ctypes.data, d.ctypes.data if d.size else None, int(d.size), Rm.ctypes.data if Rm is not None else None,
            Pm.ctypes.data if Pm is not None else None, H, W, m1.ctypes.data, m2.ctypes.data), "b200s_build_maps")
        return m1, m2

    def bgr2gray(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        if img.ndim != 3 or img.shape[2] != 3:
            raise ValueError("bgr2gray expects HxWx3 uint8")
        H, W, _ = img.shape
        out = np.empty((H, W), np.uint8)
        self._check(self._lib.b200s_bgr2gray(self._h, img.ctypes.data, H, W, img.strides[0], out.ctypes.data),
                    "b200s_bgr2gray")
        return out

    def resize_area(self, img, dsize):
        """== cv2.resize(img, (dW, dH), interpolation=cv2.INTER_AREA) for uint8 HxW or HxWx3, shrinking only."""
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim not in (2, 3) or (img.ndim == 3 and img.shape[2] not in (1, 3)):
            raise ValueError("resize_area expects a uint8 HxW or HxWx3 image")
        img = np.ascontiguousarray(img)
        cn = 1 if img.ndim == 2 else img.shape[2]
        dW, dH = int(dsize[0]), int(dsize[1])
        out = np.empty((dH, dW) if img.ndim == 2 else (dH, dW, cn), np.uint8)
        self._check(self._lib.b200s_resize_area_u8(self._h, img.ctypes.data, img.shape[0], img.shape[1], cn,
                                                   img.strides[0], out.ctypes.data, dH, dW), "b200s_resize_area_u8")
        return out

    def resize_linear(self, img, dsize):
        """== cv2.resize(img, (dW, dH), interpolation=cv2.INTER_LINEAR) for uint8 HxW or HxWx3 (rectify.py:92-105)."""
        img = np.asarray(img)
        if img.dtype != np.uint8 or img.ndim not in (2, 3) or (img.ndim == 3 and img.shape[2] not in (1, 3)):
            raise ValueError("resize_linear expects a uint8 HxW or HxWx3 image")
        img = np.ascontiguousarray(img)
        cn = 1 if img.ndim == 2 else img.shape[2]
        dW, dH = int(dsize[0]), int(dsize[1])
        out = np.empty((dH, dW) if img.ndim == 2 else (dH, dW, cn), np.uint8)
        self._check(self._lib.b200s_resize_linear_u8(self._h, img.ctypes.data, img.shape[0], img.shape[1], cn,
                                                     img.strides[0], out.ctypes.data, dH, dW), "b200s_resize_linear_u8")
        return out

    def remap(self, left, right):
        if self.maps_shape is None:
            raise EngineError("remap: rectification maps not set")
        left, right = _pair(left, right)
        H, W = self.maps_shape
        oL, oR = np.empty((H, W), np.uint8), np.empty((H, W), np.uint8)
        self._check(self._lib.b200s_remap_u8(self._h, left.ctypes.data, right.ctypes.data, left.shape[0],
                                             left.shape[1], left.strides[0], oL.ctypes.data, oR.ctypes.data),
                    "b200s_remap_u8")
        return oL, oR

    def compute(self, left, right) -> np.ndarray:
        """== cv2.StereoSGBM.compute(left, right): int16 HxW, 16x fixed point."""
        left, right = _pair(left, right)
        H, W = left.shape
        out = np.empty((H, W), np.int16)
        self._check(self._lib.b200s_sgbm_compute(self._h, left.ctypes.data, right.ctypes.data, H, W,
                                                 left.strides[0], out.ctypes.data), "b200s_sgbm_compute")
        return out

    def postprocess(self, disp, pp: B200sPostParams) -> np.ndarray:
        out = np.array(disp, dtype=np.float32, order="C", copy=True)
        if out.ndim != 2:
            raise ValueError("postprocess expects a 2-D disparity map")
        if out.size:
            self._check(self._lib.b200s_postprocess(self._h, out.ctypes.data, out.shape[0], out.shape[1],
                                                    C.byref(pp)), "b200s_postprocess")
        return out

    def fill_nearest(self, disp, mask=None, kernel_size: int = 5) -> np.ndarray:
        """== fill_holes(disp, mask, method='nearest', kernel_size) of depthlib/postprocess.py:72-118 (mask None: disp <= 0)."""
        out = np.array(disp, dtype=np.float32, order="C", copy=True)
        if out.ndim != 2:
            raise ValueError("fill_nearest expects a 2-D disparity map")
        m = None if mask is None else np.ascontiguousarray(np.asarray(mask, bool), dtype=np.uint8)
        if m is not None and m.shape != out.shape:
            raise ValueError("fill_nearest: mask and disparity shapes differ")
        if out.size:
            self._check(self._lib.b200s_fill_nearest(self._h, out.ctypes.data, None if m is None else m.ctypes.data,
                                                     out.shape[0], out.shape[1], int(kernel_size)), "b200s_fill_nearest")
        return out

    def disp_to_depth(self, disp, dp: B200sDepthParams) -> np.ndarray:
        a = np.ascontiguousarray(disp, dtype=np.float32)
        out = np.empty_like(a)
        if a.size:
            self._check(self._lib.b200s_disp_to_depth(self._h, a.ctypes.data, a.size, C.byref(dp), out.ctypes.data),
                        "b200s_disp_to_depth")
        return out

    def pipeline(self, left, right, pp: B200sPostParams, dp: B200sDepthParams, use_maps=False, want_rect=False):
        """rectify(optional) -> match -> crop -> post -> depth for one gray pair, one C-ABI call."""
        left, right = _pair(left, right)
        H, W = left.shape
        Wc = W - self.params.numDisparities
        if Wc <= 0:
            raise ValueError("num_disp must be smaller than the image width")
        disp = np.empty((H, Wc), np.float32)
        depth = np.empty((H, Wc), np.float32) if dp.enabled else None
        rl = np.empty((H, W), np.uint8) if want_rect else None
        rr = np.empty((H, W), np.uint8) if want_rect else None
        self._check(self._lib.b200s_pipeline(
            self._h, left.ctypes.data, right.ctypes.data, H, W, left.strides[0], int(bool(use_maps)), C.byref(pp),
            C.byref(dp), disp.ctypes.data, depth.ctypes.data if depth is not None else None,
            rl.ctypes.data if rl is not None else None, rr.ctypes.data if rr is not None else None), "b200s_pipeline")
        return disp, depth, rl, rr

    # -- device-resident batch ops (pointers are ints, e.g. torch.Tensor.data_ptr()) ---------------
    def compute_batch_dev(self, dL: int, dR: int, B: int, H: int, W: int, dDisp16: int):
        self._check(self._lib.b200s_sgbm_compute_batch_dev(self._h, dL, dR, B, H, W, dDisp16),
                    "b200s_sgbm_compute_batch_dev")

    def pipeline_batch_dev(self, dL: int, dR: int, B: int, H: int, W: int, pp, dp, dDisp: int, dDepth,
                           use_maps=False):
        self._check(self._lib.b200s_pipeline_batch_dev(self._h, dL, dR, B, H, W, int(bool(use_maps)), C.byref(pp),
                                                       C.byref(dp), dDisp, dDepth), "b200s_pipeline_batch_dev")

    def pipeline_batch(self, lefts, rights, pp, dp, use_maps=False, out_disp=None, out_depth=None, want_disp=True):
        """Host batch: lefts/rights uint8 [B,H,W] C-contiguous (pinned for best speed) -> float32 [B,H,W-D].
        want_disp=False: depth only (the disparity plane is not copied back; returns (None, depth))."""
        lefts, rights = np.asarray(lefts), np.asarray(rights)
        if lefts.dtype != np.uint8 or rights.dtype != np.uint8 or lefts.ndim != 3 or lefts.shape != rights.shape:
            raise ValueError("pipeline_batch expects two uint8 [B,H,W] arrays of equal shape")
        lefts, rights = np.ascontiguousarray(lefts), np.ascontiguousarray(rights)
        B, H, W = lefts.shape
        Wc = W - self.params.numDisparities
        if Wc <= 0:
            raise ValueError("num_disp must be smaller than the image width")
        disp = (out_disp if out_disp is not None else np.empty((B, H, Wc), np.float32)) if want_disp else None
        depth = (out_depth if out_depth is not None else np.empty((B, H, Wc), np.float32)) if dp.enabled else None
        self._check(self._lib.b200s_pipeline_batch(
            self._h, lefts.ctypes.data, rights.ctypes.data, B, H, W, int(bool(use_maps)), C.byref(pp), C.byref(dp),
            disp.ctypes.data if disp is not None else None, depth.ctypes.data if depth is not None else None),
            "b200s_pipeline_batch")
        return disp, depth

    def pipeline_batch_submit(self, lefts, rights, pp, dp, use_maps=False, out_disp=None, out_depth=None, want_disp=True):
        """Asynchronous pipeline_batch: enqueue H2D -> kernels -> D2H for one chunk and return a ticket at once.
        At most four chunks in flight; the arrays (pinned for real overlap) must not be touched until
        pipeline_batch_wait(ticket) returns (disp, depth)."""
        lefts, rights = np.asarray(lefts), np.asarray(rights)
        if lefts.dtype != np.uint8 or rights.dtype != np.uint8 or lefts.ndim != 3 or lefts.shape != rights.shape:
            raise ValueError("pipeline_batch expects two uint8 [B,H,W] arrays of equal shape")
        lefts, rights = np.ascontiguousarray(lefts), np.ascontiguousarray(rights)
        B, H, W = lefts.shape
        Wc = W - self.params.numDisparities
        if Wc <= 0:
            raise ValueError("num_disp must be smaller than the image width")
        disp = (out_disp if out_disp is not None else np.empty((B, H, Wc), np.float32)) if want_disp else None
        depth = (out_depth if out_depth is not None else np.empty((B, H, Wc), np.float32)) if dp.enabled else None
        ticket = C.c_int(0)
        self._check(self._lib.b200s_pipeline_batch_submit(
            self._h, lefts.ctypes.data, rights.ctypes.data, B, H, W, int(bool(use_maps)), C.byref(pp), C.byref(dp),
            disp.ctypes.data if disp is not None else None, depth.ctypes.data if depth is not None else None,
            C.byref(ticket)), "b200s_pipeline_batch_submit")
        self._inflight[ticket.value] = (lefts, rights, disp, depth, pp, dp)   # keep the buffers alive
        return ticket.value

    def pipeline_batch_wait(self, ticket: int):
        self._check(self._lib.b200s_pipeline_batch_wait(self._h, int(ticket)), "b200s_pipeline_batch_wait")
        held = self._inflight.pop(int(ticket))
        if isinstance(held, dict):
            return held["disp"], held["disp16"], held["depth"]
        _, _, disp, depth, _, _ = held
        return disp, depth

    def pipeline_frames_submit(self, lefts, rights, pp, dp, size=None, area=None, use_maps=False, want_disp=True,
                               want_disp16=False, want_depth=True, out_disp=None, out_disp16=None, out_depth=None):
        """The general asynchronous batch call (b200s_pipeline_frames_submit): ``lefts`` / ``rights`` are uint8
        [B,h,w] or [B,h,w,3] frames as captured; ``area`` = (W, H) of the INTER_AREA downscale (input.py:39-43), ``size``
        = (W, H) the matcher runs at (the calibration size; INTER_LINEAR fix-up when it differs, rectify.py:92-105);
        gray conversion, remap, matching, post-filters and depth follow on the device. Only the requested planes are
        copied back: float32 disparity, int16 x16 disparity, float32 depth. Returns a ticket for pipeline_batch_wait,
        which returns (disp, disp16, depth) with None for planes not requested."""
        lefts, rights = np.asarray(lefts), np.asarray(rights)
        if lefts.dtype != np.uint8 or rights.dtype != np.uint8 or lefts.ndim not in (3, 4) or lefts.shape != rights.shape \
                or (lefts.ndim == 4 and lefts.shape[3] not in (1, 3)):
            raise ValueError("pipeline_frames expects two uint8 [B,H,W] or [B,H,W,3] arrays of equal shape")
        lefts, rights = np.ascontiguousarray(lefts), np.ascontiguousarray(rights)
        B, sH, sW = lefts.shape[:3]
        cn = 1 if lefts.ndim == 3 else int(lefts.shape[3])
        aW, aH = (int(area[0]), int(area[1])) if area is not None else (sW, sH)
        W, H = (int(size[0]), int(size[1])) if size is not None else (aW, aH)
        Wc = W - self.params.numDisparities
        if Wc <= 0:
            raise ValueError("num_disp must be smaller than the image width")
        want_depth = bool(want_depth and dp.enabled)
        disp = (out_disp if out_disp is not None else np.empty((B, H, Wc), np.float32)) if want_disp else None
        disp16 = (out_disp16 if out_disp16 is not None else np.empty((B, H, Wc), np.int16)) if want_disp16 else None
        depth = (out_depth if out_depth is not None else np.empty((B, H, Wc), np.float32)) if want_depth else None
        for a, shp in ((disp, np.float32), (disp16, np.int16), (depth, np.float32)):
            if a is not None and (a.dtype != shp or a.shape != (B, H, Wc) or not a.flags.c_contiguous):
                raise ValueError("output buffers must be C-contiguous [B, H, W - num_disp] arrays of the plane's dtype")
        io = B200sFrameIO(lefts.ctypes.data, rights.ctypes.data, B, sH, sW, cn, aH, aW, H, W, int(bool(use_maps)),
                          disp.ctypes.data if disp is not None else None,
                          disp16.ctypes.data if disp16 is not None else None,
                          depth.ctypes.data if depth is not None else None)
        ticket = C.c_int(0)
        self._check(self._lib.b200s_pipeline_frames_submit(self._h, C.byref(io), C.byref(pp), C.byref(dp),
                                                           C.byref(ticket)), "b200s_pipeline_frames_submit")
        self._inflight[ticket.value] = dict(L=lefts, R=rights, disp=disp, disp16=disp16, depth=depth, pp=pp, dp=dp)
        return ticket.value

    def pipeline_frames(self, lefts, rights, pp, dp, **kw):
        return self.pipeline_batch_wait(self.pipeline_frames_submit(lefts, rights, pp, dp, **kw))

    @property
    def last_flow(self) -> int:
        """B200S_FLOW_* bits of the last matcher call (which kernels ran)."""
        return int(self._lib.b200s_last_flow(self._h))

    def stage_times(self, reset=False):
        """{stage: (device ms, kernel launches)} accumulated while timing is enabled."""
        n = len(_lib.STAGES)
        ms = (C.c_double * n)()
        ln = (C.c_longlong * n)()
        self._lib.b200s_stage_times(self._h, ms, ln, n)
        if reset:
            self._lib.b200s_reset_stage_times(self._h)
        return {name: (ms[i], int(ln[i])) for i, name in enumerate(_lib.STAGES)}

    # -- stage taps (tests) ----------------------------------------------------------------------------
    def debug_stage(self, which: str, H: int, W: int, b: int = 0) -> np.ndarray:
        g = self.geometry(W)
        D = self.params.numDisparities
        idx = {"C": 0, "S": 1, "raw": 2}[which]
        out = np.empty((H, W), np.int16) if idx == 2 else np.empty((H, g["W1"], D), np.int16)
        self._check(self._lib.b200s_debug_stage(self._h, idx, b, out.ctypes.data), "b200s_debug_stage")
        return out
