"""depthestimation_b200 — B200-native (sm_100a) engine for depthlib's stereo disparity hot path.

``engine.StereoEngine`` is the thin Python face of the C ABI (include/b200stereo.h);
``depthlib`` (sub-package) mirrors the reference's StereoCore / StereoDepthEstimator API on top.
"""
__version__ = "0.1.0"
