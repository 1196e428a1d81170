#!/usr/bin/env python
"""tools/make_golden.py — generate tests/golden/*.npz from the REFERENCE ITSELF.

Runs only in the build container (needs /root/reference and cv2). It imports the reference's own
``depthlib.stereo_core.StereoCore`` (with a throw-away 2-file matplotlib stub, because
``depthlib/__init__.py`` imports matplotlib transitively and matplotlib is not installed) and
records, for a handful of small seeded cases, everything the parity tests need:

  left/right inputs, rectification maps, rectified pair, raw int16 disparity of
  ``core.sgbm.compute`` (stereo_core.py:231), post-processed float disparity and depth of
  ``StereoCore.estimate_depth`` (stereo_core.py:274-293).

The fixtures travel to the GPU box (which has no /root/reference); tests/ compare both the
C oracle and the CUDA engine against them.

    python tools/make_golden.py            # rewrites tests/golden/*.npz
"""
from __future__ import annotations

import os
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference")


def _import_reference():
    stub = Path(tempfile.mkdtemp(prefix="mplstub_"))
    (stub / "matplotlib").mkdir()
    (stub / "matplotlib" / "__init__.py").write_text("")
    (stub / "matplotlib" / "pyplot.py").write_text("")
    sys.path.insert(0, str(stub))
    sys.path.insert(0, str(REF))
    import depthlib.stereo_core as sc  # noqa: E402  (the reference's module)
    assert str(REF) in sc.__file__, sc.__file__
    return sc


# name -> (H, W, pair-D, seed, configure kwargs, calibrated?, downscale, fast_mode)
CASES = {
    "sgbm_d64": (72, 200, 64, 1001, dict(num_disp=64, sgbm_mode="sgbm"), False, 1.0, False),
    "hh_d48_bs7": (64, 176, 48, 1002, dict(num_disp=48, block_size=7, sgbm_mode="hh", uniqueness_ratio=5,
                                           speckle_window_size=30, speckle_range=1), False, 1.0, False),
    "3way_default_d64": (96, 208, 64, 1003, dict(num_disp=64), False, 1.0, False),
    "3way_d44_tail": (80, 160, 44, 1004, dict(num_disp=44, block_size=3), False, 1.0, False),
    "hh4_d32_mind8": (70, 150, 32, 1005, dict(num_disp=32, min_disp=8, sgbm_mode="hh4",
                                              disp12_max_diff=2, prefilter_cap=63), False, 1.0, False),
    "sgbm_negmin": (60, 168, 40, 1006, dict(num_disp=40, min_disp=-8, sgbm_mode="sgbm", block_size=9,
                                            speckle_window_size=0), False, 1.0, False),
    "calib_sgbm_d64_depth": (96, 224, 64, 1007, dict(num_disp=64, sgbm_mode="sgbm", max_depth=50.0), True, 1.0, False),
    "calib_3way_fast": (96, 224, 48, 1008, dict(num_disp=48), True, 1.0, True),
    "calib_3way_ds07": (96, 224, 64, 1009, dict(num_disp=64, doffs=3.5), True, 0.7, False),
}


def main():
    sc = _import_reference()
    import cv2
    from depthestimation_b200.synth import make_pair, make_calibration
    out_dir = ROOT / "tests" / "golden"
    out_dir.mkdir(parents=True, exist_ok=True)
    for name, (H, W, D, seed, kw, calibrated, ds, fast) in CASES.items():
        L, R, _ = make_pair(H, W, D, seed)
        core = sc.StereoCore(downscale_factor=ds, fast_mode=fast)
        kw = dict(kw)
        rec = {}
        if calibrated:
            cal = make_calibration(H, W)
            kw.update(cam_matrix_L=cal["cam_matrix_L"], cam_matrix_R=cal["cam_matrix_R"],
                      dist_coeff_L=cal["dist_coeff_L"], dist_coeff_R=cal["dist_coeff_R"],
                      baseline=cal["baseline"], image_width=W, image_height=H,
                      focal_length=cal["focal_length"])
        else:
            kw.update(focal_length=0.8 * W, baseline=0.54)
        core.configure_sgbm(**kw)
        disp, depth = core.estimate_depth(L, R)
        raw16 = core.sgbm.compute(core.left_rectified, core.right_rectified)
        p = core.get_sgbm_params()
        rec.update(left=L, right=R, left_rect=core.left_rectified, right_rect=core.right_rectified,
                   disp16=raw16, disp=disp, depth=depth,
                   params=np.array([p["min_disp"], p["num_disp"], p["block_size"], p["disp12_max_diff"],
                                    p["prefilter_cap"], p["uniqueness_ratio"], p["speckle_window_size"],
                                    p["speckle_range"]], np.int32),
                   mode=np.array(p["sgbm_mode"]), downscale=np.float64(ds), fast=np.int32(fast),
                   focal_length=np.float64(p["focal_length"]), baseline=np.float64(p["baseline"]),
                   doffs=np.float64(p["doffs"]),
                   max_depth=np.float64(p["max_depth"] if p["max_depth"] is not None else np.nan),
                   cv2_version=np.array(cv2.__version__))
        if calibrated:
            maps = core._rect_cache.get_maps(p["cam_matrix_L"], p["cam_matrix_R"], p["baseline"], W, H,
                                             dist_coeff_L=p["dist_coeff_L"], dist_coeff_R=p["dist_coeff_R"],
                                             rotation=None, translation=None, alpha=1.0)
            rec.update(map1_L=maps["map1_L"], map2_L=maps["map2_L"], map1_R=maps["map1_R"], map2_R=maps["map2_R"])
        np.savez_compressed(out_dir / f"{name}.npz", **rec)
        print(f"{name}: disp16 valid={(raw16 > (p['min_disp'] - 1) * 16).mean():.2f} "
              f"disp {disp.shape} depth {None if depth is None else depth.shape}")


if __name__ == "__main__":
    main()
