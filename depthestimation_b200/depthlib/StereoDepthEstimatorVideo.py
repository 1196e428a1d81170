"""StereoDepthEstimatorVideo — same façade as depthlib/StereoDepthEstimatorVideo.py of the reference:
a generator that yields one depth map per stereo frame, with threaded capture, ``target_fps`` pacing and
``drop_frames``. Differences, both forced by the platform and documented in DESIGN.md:

* the HighGUI window / waitKey calls of the reference (:81-82, :111, :138) are left out — display is
  "unavailable on this box"; ``visualize_live=True`` raises when the first frame would be drawn
* ``devices=[...]`` (new, optional) shards chunks of frames across several B200s through
  ``sharding.ThreadedSharder`` and still yields in frame order
"""
from __future__ import annotations

import time
from typing import Optional, Sequence

import numpy as np

from ..sharding import ThreadedSharder
from .input import stereo_stream
from .rectify import _ensure_image_size
from .stereo_core import StereoCore
from .threaded_stereo import ThreadedStereoCapture
from .visualizations import visualize_depth_live, visualize_depth_live_gray


class StereoDepthEstimatorVideo:
    def __init__(self, left_source=None, right_source=None, downscale_factor=1.0, visualize_live=False,
                 saving_path=None, fast_mode=False, use_threading=True, target_fps=30, drop_frames=False,
                 visualize_gray=False, devices: Optional[Sequence[int]] = None, chunk: int = 8) -> None:
        self.left_source, self.right_source = left_source, right_source
        self.downscale_factor = downscale_factor
        self.visualize_live = visualize_live
        self.saving_path = saving_path          # accepted and unused, as upstream
        self.fast_mode = fast_mode
        self.use_threading = use_threading
        self.target_fps = target_fps
        self._frame_interval = 1.0 / target_fps if target_fps > 0 else 0
        self.drop_frames = drop_frames
        self.visualize_gray = visualize_gray
        self.devices = list(devices) if devices else None
        self.chunk = chunk
        self.core = StereoCore(downscale_factor=downscale_factor, fast_mode=fast_mode,
                               device=self.devices[0] if self.devices else 0)

    def configure_sgbm(self, **kwargs):
        self.core.configure_sgbm(**kwargs)

    # -- frame sources ----------------------------------------------------------------------------
    def _frames(self):
        if self.use_threading:
            cap = ThreadedStereoCapture(self.left_source, self.right_source, downscale_factor=self.downscale_factor,
                                        drop_frames=self.drop_frames)
            cap.start()
            try:
                while True:
                    pair = cap.read()
                    if pair is None:
                        return
                    yield pair
            finally:
                cap.stop()
        else:
            yield from stereo_stream(self.left_source, self.right_source, downscale_factor=self.downscale_factor)

    def _pace(self, t0):
        if self._frame_interval > 0:
            sleep = self._frame_interval - (time.time() - t0)
            if sleep > 0:
                time.sleep(sleep)
        return time.time()

    def _show(self, depth_m):
        if self.visualize_live:
            (visualize_depth_live_gray if self.visualize_gray else visualize_depth_live)(depth_m, self.target_fps)

    # -- the generator (reference: StereoDepthEstimatorVideo.py:69-148) ------------------------------
    def estimate_depth(self):
        if self.left_source is None or self.right_source is None:
            raise ValueError("Both left_source and right_source must be provided for video depth estimation.")
        # the reference re-applies the (already scaled) parameters here, which scales num_disp /
        # focal_length / doffs a second time when downscale_factor != 1 (:78) — kept for parity
        self.core.configure_sgbm(**self.core.get_sgbm_params())
        if self.devices and (len(self.devices) > 1 or self.chunk > 1):
            yield from self._estimate_sharded()
            return
        t0 = time.time()
        for left, right in self._frames():
            _disp, depth_m = self.core.estimate_depth(left, right)
            yield depth_m
            self._show(depth_m)
            t0 = self._pace(t0)

    def _estimate_sharded(self):
        """Chunks of frames round-robin over ``self.devices``; host gather in frame order."""
        from ..engine import StereoEngine
        core = self.core
        p = core.get_sgbm_params()
        use_maps = core._calibration()
        maps = None
        if use_maps:
            maps = core._rect_cache.get_maps(p['cam_matrix_L'], p['cam_matrix_R'], p['baseline'], p['image_width'],
                                             p['image_height'], p.get('dist_coeff_L'), p.get('dist_coeff_R'),
                                             p.get('rotation'), p.get('translation'), 1.0)
        pp, dp, kw = core._post_params(), core._depth_params(), core._matcher_kwargs()

        def make_worker(dev):
            eng = StereoEngine(dev, **kw)
            if maps is not None:
                eng.set_maps(maps['map1_L'], maps['map2_L'], maps['map1_R'], maps['map2_R'])
            return lambda L, R: eng.pipeline_batch(L, R, pp, dp, use_maps=use_maps)

        def gray_pairs():
            eng0 = core.sgbm.engine
            for left, right in self._frames():
                if use_maps:
                    size = (p['image_width'], p['image_height'])
                    left, right = _ensure_image_size(left, size), _ensure_image_size(right, size)
                if left.ndim == 3:
                    left = eng0.bgr2gray(left)
                if right.ndim == 3:
                    right = eng0.bgr2gray(right)
                yield left, right

        t0 = time.time()
        for _disp, depth_m in ThreadedSharder(self.devices, make_worker, self.chunk).run(gray_pairs()):
            yield depth_m
            self._show(depth_m)
            t0 = self._pace(t0)
