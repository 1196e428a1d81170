"""Display helpers of the reference (depthlib/visualizations.py) need matplotlib / OpenCV HighGUI.
Neither exists on a headless B200 box ("unavailable on this box"): the names are kept so
that callers import cleanly, and each raises a clear error when actually asked to draw."""


def _unavailable(name):
    def fn(*_a, **_k):
        raise RuntimeError(f"{name}: display is unavailable on this box (no matplotlib / HighGUI); "
                           "the depth maps are returned as numpy arrays instead")
    fn.__name__ = name
    return fn


visualize_stereo_pair = _unavailable("visualize_stereo_pair")
visualize_disparity = _unavailable("visualize_disparity")
visualize_depth = _unavailable("visualize_depth")
visualize_disparity_and_depth = _unavailable("visualize_disparity_and_depth")
visualize_depth_live = _unavailable("visualize_depth_live")
visualize_depth_live_gray = _unavailable("visualize_depth_live_gray")
