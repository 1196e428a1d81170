"""Stereo rectification — drop-in for depthlib/rectify.py of the reference.

Per frame (hot path): gray conversion and the two bilinear remaps run on the GPU (C ABI
``b200s_bgr2gray`` / ``b200s_remap_u8``; reference: rectify.py:108-119, 185-186).
Once per calibration (or per call when ``rectify_images`` gets no cache, rectify.py:209-235): ``cv2.stereoRectify`` (3x3
algebra) on the host, then the per-pixel ``cv2.initUndistortRectifyMap`` of rectify.py:68-73 on the GPU in double
precision (``b200s_build_maps``). The size fix-up (rectify.py:92-105) runs on the GPU too (``b200s_resize_linear_u8``).
"""
from __future__ import annotations

import warnings
from typing import Any, Dict, Optional, Tuple

import numpy as np

__all__ = ["rectify_images", "RectificationCache"]

def _engine(engine=None, device: int = 0):
    """The caller's engine, or the calling thread's own engine on ``device`` for the free functions (the reference's
    free cv2 calls are thread-safe; a handle owns a stream and scratch, so it is never shared between threads)."""
    if engine is not None:
        return engine
    from ..engine import thread_engine
    return thread_engine(device)


def _cv2():
    try:
        import cv2
        return cv2
    except Exception as exc:  # pragma: no cover
        raise RuntimeError("building rectification maps needs host cv2 (stereoRectify / initUndistortRectifyMap); "
                           "pass precomputed maps to StereoEngine.set_maps instead") from exc


def _key_part(arr):
    return None if arr is None else tuple(np.asarray(arr).flatten())


class RectificationCache:
    """Keeps the four CV_32FC1 maps of the last calibration (reference: rectify.py:14-87): same key,
    same dict object returned while the calibration is unchanged."""

    def __init__(self, engine=None):
        self._cache_key: Optional[tuple] = None
        self._maps: Optional[Dict[str, np.ndarray]] = None
        self.engine = engine     # StereoEngine that generates the maps (None: the calling thread's engine on device 0)

    def _make_key(self, cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                  dist_coeff_L, dist_coeff_R, rotation, translation, alpha) -> tuple:
        return (_key_part(cam_matrix_L), _key_part(cam_matrix_R), baseline, image_width, image_height,
                _key_part(dist_coeff_L), _key_part(dist_coeff_R), _key_part(rotation), _key_part(translation), alpha)

    def get_maps(self, cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                 dist_coeff_L=None, dist_coeff_R=None, rotation=None, translation=None,
                 alpha=0.0) -> Dict[str, np.ndarray]:
        key = self._make_key(cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                             dist_coeff_L, dist_coeff_R, rotation, translation, alpha)
        if self._cache_key == key and self._maps is not None:
            return self._maps
        self._maps = _build_maps(cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                                 dist_coeff_L, dist_coeff_R, rotation, translation, alpha, getattr(self, "engine", None))
        self._cache_key = key
        return self._maps

    def clear(self):
        self._cache_key = None
        self._maps = None


_global_cache = RectificationCache()


def _build_maps(cam_matrix_L, cam_matrix_R, baseline, image_width, image_height, dist_coeff_L, dist_coeff_R,
                rotation, translation, alpha, engine=None) -> Dict[str, np.ndarray]:
    """cv2.stereoRectify(CALIB_ZERO_DISPARITY) + 2x initUndistortRectifyMap(CV_32FC1); defaults as the
    reference: zero distortion, R = I, T = [-baseline, 0, 0] (rectify.py:52-73). stereoRectify is 3x3 algebra, once per
    calibration, and stays host cv2; the per-pixel part, initUndistortRectifyMap, runs on the GPU (b200s_build_maps)."""
    cv2 = _cv2()
    K_L = np.asarray(cam_matrix_L, dtype=np.float64)
    K_R = np.asarray(cam_matrix_R, dtype=np.float64)
    zero = np.zeros(5, dtype=np.float64)
    d_L = np.asarray(dist_coeff_L, dtype=np.float64) if dist_coeff_L is not None else zero
    d_R = np.asarray(dist_coeff_R, dtype=np.float64) if dist_coeff_R is not None else zero
    size = (image_width, image_height)
    R = np.asarray(rotation, dtype=np.float64) if rotation is not None else np.eye(3, dtype=np.float64)
    T = (np.asarray(translation, dtype=np.float64) if translation is not None
         else np.array([-baseline, 0.0, 0.0], dtype=np.float64))
    rect_L, rect_R, proj_L, proj_R, _Q, _roi_L, _roi_R = cv2.stereoRectify(
        K_L, d_L, K_R, d_R, size, R, T, flags=cv2.CALIB_ZERO_DISPARITY, alpha=alpha)
    eng = _engine(engine)
    map1_L, map2_L = eng.build_maps(K_L, d_L, rect_L, proj_L, size)
    map1_R, map2_R = eng.build_maps(K_R, d_R, rect_R, proj_R, size)
    return {'map1_L': map1_L, 'map2_L': map2_L, 'map1_R': map1_R, 'map2_R': map2_R}


def _ensure_image_size(image: np.ndarray, size: Tuple[int, int], engine=None) -> np.ndarray:
    """Warn and resize (INTER_LINEAR) when the image does not match the calibration size (rectify.py:92-105). Runs on
    the GPU (b200s_resize_linear_u8, bit-exact) for what the stereo path carries: uint8 gray / 3-channel images.
    There is no host fallback: any other dtype or layout raises."""
    height, width = image.shape[:2]
    if (width, height) == tuple(size):
        return image
    warnings.warn("Input image size %sx%s does not match calibration %sx%s; resizing for rectification." %
                  (width, height, size[0], size[1]), RuntimeWarning, stacklevel=2)
    if image.dtype != np.uint8 or not (image.ndim == 2 or (image.ndim == 3 and image.shape[2] in (1, 3))):
        raise ValueError("size fix-up runs on the GPU for uint8 HxW / HxWx1 / HxWx3 images only (got %s %s); "
                         "there is no CPU fallback" % (image.dtype, image.shape))
    out = _engine(engine).resize_linear(image, tuple(size))
    return out[:, :, 0] if (image.ndim == 3 and image.shape[2] == 1) else out   # cv2.resize drops a singleton channel


def _to_grayscale(image: np.ndarray, engine=None) -> np.ndarray:
    """Gray conversion with BGR weights on the channels as stored (GPU; reference: rectify.py:108-119)."""
    if image.ndim == 2:
        return image
    if image.ndim == 3 and image.shape[2] == 1:
        return image[:, :, 0]
    if image.ndim == 3 and image.shape[2] == 3:
        return _engine(engine).bgr2gray(image)
    raise ValueError("Unsupported image format for grayscale conversion")


def rectify_images(img_L: np.ndarray, img_R: np.ndarray, cam_matrix_L: np.ndarray, cam_matrix_R: np.ndarray,
                   baseline: float, image_width: int, image_height: int,
                   dist_coeff_L: Any = None, dist_coeff_R: Any = None, rotation: Any = None, translation: Any = None,
                   alpha: float = 0.0, cache: Optional[RectificationCache] = None, engine=None
                   ) -> Tuple[np.ndarray, np.ndarray]:
    """Rectify a stereo pair; returns two uint8 gray images of the calibration size (rectify.py:122-237).
    With ``cache`` the maps are reused across frames; without it they are rebuilt per call, as upstream."""
    if cache is not None:
        if engine is not None and getattr(cache, "engine", None) is None:
            cache.engine = engine           # the maps are generated on the caller's device
        maps = cache.get_maps(cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                              dist_coeff_L, dist_coeff_R, rotation, translation, alpha)
    else:
        maps = _build_maps(cam_matrix_L, cam_matrix_R, baseline, image_width, image_height,
                           dist_coeff_L, dist_coeff_R, rotation, translation, alpha, engine)
    size = (image_width, image_height)
    eng = _engine(engine)
    img_L = _ensure_image_size(img_L, size, eng)
    img_R = _ensure_image_size(img_R, size, eng)
    gray_L = np.ascontiguousarray(_to_grayscale(img_L, eng))
    gray_R = np.ascontiguousarray(_to_grayscale(img_R, eng))
    if getattr(eng, "_maps_id", None) != id(maps):
        eng.set_maps(maps['map1_L'], maps['map2_L'], maps['map1_R'], maps['map2_R'])
        eng._maps_id = id(maps)
        eng._maps_ref = maps  # keep the dict alive so its id cannot be reused
    return eng.remap(gray_L, gray_R)
