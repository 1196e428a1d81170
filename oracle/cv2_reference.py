"""oracle/cv2_reference.py — TEST INFRASTRUCTURE ONLY.

The live ``cv2`` binary driven with exactly the arguments depthlib uses, so the C restatement
(``sgbm_oracle.c`` / ``imgops_oracle.c``) can be pinned against the real arithmetic engine and so
``bench.py`` has a CPU baseline that is the reference's own code path:

* ``create_sgbm``            depthlib/stereo_core.py:44-75   (P1 = 8 bs^2, P2 = 32 bs^2, mode map)
* ``compute_disparity16``    depthlib/stereo_core.py:231
* ``rectify_pair``           depthlib/rectify.py:63-73,175-187 (maps once, remap per frame)
* ``postprocess_disparity``  depthlib/postprocess.py:6-70,120-171 (hole filling off)
* ``estimate_depth``         depthlib/stereo_core.py:125-200,234-293 — whole per-frame path

cv2 here is opencv-python-headless 4.13.0.92 (the reference pins opencv-python 4.12.0.88, which
cannot be installed offline); the version is reported with every result.
"""
from __future__ import annotations

import numpy as np

try:  # cv2 is present in this image; the import is guarded so callers can report its absence
    import cv2
    HAVE_CV2 = True
    CV2_VERSION = cv2.__version__
except Exception:  # pragma: no cover
    cv2 = None
    HAVE_CV2 = False
    CV2_VERSION = None

_MODE = {"sgbm": 0, "hh": 1, "sgbm_3way": 2, "hh4": 3}


def create_sgbm(min_disp=0, num_disp=128, block_size=5, disp12_max_diff=1, prefilter_cap=31,
                uniqueness_ratio=10, speckle_window_size=50, speckle_range=2, sgbm_mode="sgbm_3way",
                P1=None, P2=None):
    if P1 is None:
        P1 = 8 * block_size ** 2
    if P2 is None:
        P2 = 32 * block_size ** 2
    return cv2.StereoSGBM_create(
        minDisparity=min_disp, numDisparities=num_disp, blockSize=block_size, P1=P1, P2=P2,
        disp12MaxDiff=disp12_max_diff, preFilterCap=prefilter_cap, uniquenessRatio=uniqueness_ratio,
        speckleWindowSize=speckle_window_size, speckleRange=speckle_range,
        mode=_MODE.get(sgbm_mode, 2))


def compute_disparity16(left, right, **kw) -> np.ndarray:
    return create_sgbm(**kw).compute(left, right)


def build_maps(calib: dict, alpha: float = 1.0):
    K_L = np.asarray(calib["cam_matrix_L"], np.float64)
    K_R = np.asarray(calib["cam_matrix_R"], np.float64)
    z = np.zeros(5, np.float64)
    d_L = np.asarray(calib.get("dist_coeff_L"), np.float64) if calib.get("dist_coeff_L") is not None else z
    d_R = np.asarray(calib.get("dist_coeff_R"), np.float64) if calib.get("dist_coeff_R") is not None else z
    size = (int(calib["image_width"]), int(calib["image_height"]))
    R = np.asarray(calib["rotation"], np.float64) if calib.get("rotation") is not None else np.eye(3)
    T = (np.asarray(calib["translation"], np.float64) if calib.get("translation") is not None
         else np.array([-calib["baseline"], 0.0, 0.0]))
    rL, rR, pL, pR, _, _, _ = cv2.stereoRectify(K_L, d_L, K_R, d_R, size, R, T,
                                                flags=cv2.CALIB_ZERO_DISPARITY, alpha=alpha)
    m1L, m2L = cv2.initUndistortRectifyMap(K_L, d_L, rL, pL, size, cv2.CV_32FC1)
    m1R, m2R = cv2.initUndistortRectifyMap(K_R, d_R, rR, pR, size, cv2.CV_32FC1)
    return dict(map1_L=m1L, map2_L=m2L, map1_R=m1R, map2_R=m2R)


def to_gray(img):
    if img.ndim == 3 and img.shape[2] == 3:
        return cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    return img if img.ndim == 2 else img[:, :, 0]


def rectify_pair(left, right, maps):
    gl, gr = to_gray(left), to_gray(right)
    return (cv2.remap(gl, maps["map1_L"], maps["map2_L"], cv2.INTER_LINEAR),
            cv2.remap(gr, maps["map1_R"], maps["map2_R"], cv2.INTER_LINEAR))


def postprocess_disparity(disp, max_speckle_size=50, max_diff=1.0, outlier_threshold=3.0,
                          apply_outlier_removal=True):
    d16 = (disp.copy() * 16.0).astype(np.int16)
    cv2.filterSpeckles(d16, 0, max_speckle_size, int(max_diff * 16))
    res = d16.astype(np.float32) / 16.0
    if apply_outlier_removal:
        mean = cv2.boxFilter(res, -1, (5, 5))
        mean_sq = cv2.boxFilter(res ** 2, -1, (5, 5))
        std = np.sqrt(np.maximum(mean_sq - mean ** 2, 0))
        res[(np.abs(res - mean) > outlier_threshold * std) & (res > 0)] = 0
    return cv2.medianBlur(res.astype(np.float32), 3)


def disparity_to_depth(disp, f_pixels, baseline_m, doffs=0.0, eps=1e-6, max_depth=None):
    adj = disp + doffs
    Z = np.full_like(disp, np.inf, dtype=np.float32)
    ok = adj > eps
    Z[ok] = (f_pixels * baseline_m) / adj[ok]
    if max_depth is not None:
        Z[Z > max_depth] = max_depth
    return Z


class Cv2StereoPipeline:
    """Per-frame path of StereoCore.estimate_depth, stage by stage (one instance per thread —
    cv2.StereoSGBM keeps internal scratch)."""

    def __init__(self, sgbm_kw: dict, calib: dict | None = None, downscale: float = 1.0,
                 fast_mode: bool = False, focal_length=None, baseline=None, doffs=0.0, max_depth=None):
        self.kw = dict(sgbm_kw)
        self.sgbm = create_sgbm(**self.kw)
        self.maps = build_maps(calib, alpha=1.0) if calib is not None else None
        self.downscale, self.fast_mode = downscale, fast_mode
        self.f, self.B, self.doffs, self.max_depth = focal_length, baseline, doffs, max_depth

    def __call__(self, left, right, timers: dict | None = None):
        import time
        t0 = time.perf_counter()
        if self.maps is not None:
            left, right = rectify_pair(left, right, self.maps)
        else:
            left, right = to_gray(left), to_gray(right)
        t1 = time.perf_counter()
        disp = self.sgbm.compute(left, right).astype(np.float32) / 16.0
        t2 = time.perf_counter()
        disp = disp[:, self.kw.get("num_disp", 128):]
        if self.fast_mode:
            disp = cv2.medianBlur(disp.astype(np.float32), 3)
        else:
            disp = postprocess_disparity(disp, max_speckle_size=int(100 * self.downscale),
                                         max_diff=1.0, outlier_threshold=2.5)
        t3 = time.perf_counter()
        depth = None
        if self.f is not None and self.B is not None:
            depth = disparity_to_depth(disp, self.f, self.B, self.doffs,
                                       eps=self.kw.get("min_disp", 0), max_depth=self.max_depth)
        t4 = time.perf_counter()
        if timers is not None:
            for k, v in (("rectify", t1 - t0), ("match", t2 - t1), ("post", t3 - t2), ("depth", t4 - t3)):
                timers[k] = timers.get(k, 0.0) + v
        return disp, depth
