"""Loading stereo pairs and live streams — same functions as depthlib/input.py of the reference.
File / codec I/O is host work (cv2.imread / cv2.VideoCapture) and outside the GPU hot path."""
from __future__ import annotations

from typing import Iterable, Tuple, Union

import numpy as np


def _cv2():
    import cv2
    return cv2


def _area_downscale(img: np.ndarray, factor: float, device: int = 0) -> np.ndarray:
    """cv2.resize(img, (int(W * f), int(H * f)), INTER_AREA) of input.py:39-43 / threaded_stereo.py:41-46, on the GPU
    (b200s_resize_area_u8, bit-exact), through the calling thread's engine on ``device``. The reference only ever
    shrinks here (0 < factor <= 1)."""
    if factor == 1.0:
        return img
    from ..engine import thread_engine
    return thread_engine(device).resize_area(img, (int(img.shape[1] * factor), int(img.shape[0] * factor)))


def load_stereo_pair(left_image_path, right_image_path, downscale_factor=1.0, device: int = 0):
    """Two image files -> two RGB arrays, optionally INTER_AREA-downscaled (reference: input.py:9-45)."""
    cv2 = _cv2()
    left, right = cv2.imread(left_image_path), cv2.imread(right_image_path)
    if left is None or right is None:
        raise FileNotFoundError("One or both image paths are invalid.")
    left, right = cv2.cvtColor(left, cv2.COLOR_BGR2RGB), cv2.cvtColor(right, cv2.COLOR_BGR2RGB)
    return _area_downscale(left, downscale_factor, device), _area_downscale(right, downscale_factor, device)


def open_capture(source: Union[int, str]):
    cap = _cv2().VideoCapture(source)
    if not cap.isOpened():
        raise RuntimeError(f"Unable to open video source: {source}")
    return cap


def _read_frame(cap, downscale_factor: float, device: int = 0) -> np.ndarray:
    ok, frame = cap.read()
    if not ok or frame is None:
        raise RuntimeError("Failed to read frame from video source")
    return _area_downscale(frame, downscale_factor, device)


def stereo_stream(left_source, right_source, downscale_factor: float = 1.0, device: int = 0) -> Iterable[Tuple[np.ndarray, np.ndarray]]:
    """Yield (left, right) BGR frames until either stream fails (then RuntimeError, as upstream: input.py:58-61)."""
    if downscale_factor <= 0 or downscale_factor > 1.0:
        raise ValueError("downscale_factor must be between 0 and 1.")
    cap_L, cap_R = open_capture(left_source), open_capture(right_source)
    try:
        while True:
            yield _read_frame(cap_L, downscale_factor, device), _read_frame(cap_R, downscale_factor, device)
    finally:
        cap_L.release()
        cap_R.release()
