"""StereoDepthEstimator — same façade as depthlib/StereoDepthEstimator.py of the reference."""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np

from .input import load_stereo_pair
from .stereo_core import StereoCore
from .visualizations import visualize_depth, visualize_disparity


class StereoDepthEstimator:
    def __init__(self, left_source=None, right_source=None, downscale_factor=1.0, device: int = 0):
        if downscale_factor <= 0 or downscale_factor > 1.0:
            raise ValueError("downscale_factor must be between 0 and 1.")
        self.downscale_factor = downscale_factor
        self.core = StereoCore(downscale_factor=downscale_factor, device=device)
        self.left_source = None
        self.right_source = None
        if left_source is not None and right_source is not None:
            self.left_source, self.right_source = load_stereo_pair(left_source, right_source,
                                                                   downscale_factor=downscale_factor, device=device)
        self.sgbm = None
        self.disparity_map = None
        self.depth_map = None

    def configure_sgbm(self, **kwargs):
        self.core.configure_sgbm(**kwargs)

    def get_sgbm_params(self) -> Dict[str, int]:
        return self.core.get_sgbm_params()

    def estimate_depth(self) -> Tuple[np.ndarray, Optional[np.ndarray]]:
        if self.left_source is None or self.right_source is None:
            raise ValueError("Left and right sources must be provided for depth estimation.")
        self.disparity_map, self.depth_map = self.core.estimate_depth(self.left_source, self.right_source)
        return self.disparity_map, self.depth_map

    def visualize_results(self):
        if self.disparity_map is None:
            raise ValueError("Disparity map not computed. Call estimate_depth() first.")
        visualize_disparity(self.disparity_map, title='Disparity Map (Raw)', cmap='jet')
        if self.depth_map is None:
            raise ValueError("Depth map not computed. Call estimate_depth() with calibration data first.")
        visualize_depth(self.depth_map, title='Depth Map (Raw)', cmap='turbo_r')
