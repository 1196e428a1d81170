"""Drop-in for the reference package ``depthlib`` (its stereo path): same names, same parameter dict,
same error texts, the per-frame arithmetic on the GPU. ``MonocularDepthEstimator`` (a HuggingFace
model wrapper) is a different path and is not part of this engine."""
from .StereoDepthEstimator import StereoDepthEstimator
from .StereoDepthEstimatorVideo import StereoDepthEstimatorVideo
from .stereo_core import StereoCore
from .threaded_stereo import ThreadedStereoCapture
from .visualizations import (visualize_stereo_pair, visualize_disparity, visualize_depth,
                             visualize_disparity_and_depth)

__all__ = [
    'StereoDepthEstimator', 'StereoDepthEstimatorVideo', 'StereoCore', 'ThreadedStereoCapture',
    'visualize_stereo_pair', 'visualize_disparity', 'visualize_depth', 'visualize_disparity_and_depth',
]
