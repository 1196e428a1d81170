"""Background frame capture — same class as depthlib/threaded_stereo.py of the reference
(daemon producer thread, bounded queue, drop-oldest policy for live cameras, 1 s read timeout)."""
from __future__ import annotations

import queue
import threading
from typing import Optional, Tuple

import numpy as np

from .input import _area_downscale, open_capture


class ThreadedStereoCapture:
    def __init__(self, left_source, right_source, downscale_factor=1.0, buffer_size=2, drop_frames=True, device=0):
        self.left_source, self.right_source = left_source, right_source
        self.device = device            # GPU that runs the INTER_AREA downscale of this capture thread
        self.downscale_factor = downscale_factor
        self.buffer_size = buffer_size
        self.drop_frames = drop_frames
        self._frame_queue: queue.Queue = queue.Queue(maxsize=buffer_size)
        self._stop_event = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._cap_L = None
        self._cap_R = None

    def start(self):
        self._cap_L, self._cap_R = open_capture(self.left_source), open_capture(self.right_source)
        self._stop_event.clear()
        self._thread = threading.Thread(target=self._capture_loop, daemon=True)
        self._thread.start()

    def _read_frame(self, cap) -> Optional[np.ndarray]:
        ok, frame = cap.read()
        if not ok or frame is None:
            return None
        return _area_downscale(frame, self.downscale_factor, self.device)

    def _capture_loop(self):
        while not self._stop_event.is_set():
            if self._cap_L is None or self._cap_R is None:
                break
            left, right = self._read_frame(self._cap_L), self._read_frame(self._cap_R)
            if left is None or right is None:
                break
            if self.drop_frames and self._frame_queue.full():
                try:
                    self._frame_queue.get_nowait()   # live camera: keep latency low, drop the oldest pair
                except queue.Empty:
                    pass
            self._frame_queue.put((left, right))      # video file: block, never skip a frame
        self._stop_event.set()

    def read(self) -> Optional[Tuple[np.ndarray, np.ndarray]]:
        if self._stop_event.is_set() and self._frame_queue.empty():
            return None
        try:
            return self._frame_queue.get(timeout=1.0)
        except queue.Empty:
            return None

    def stop(self):
        self._stop_event.set()
        if self._thread is not None:
            self._thread.join(timeout=2.0)
        for cap in (self._cap_L, self._cap_R):
            if cap is not None:
                cap.release()
