"""StereoDepthEstimatorVideo — same façade as depthlib/StereoDepthEstimatorVideo.py of the reference:
a generator that yields one depth map per stereo frame, with threaded capture, ``target_fps`` pacing and
``drop_frames``. Differences, both forced by the platform and documented in DESIGN.md:

* the HighGUI window / waitKey calls of the reference (:81-82, :111, :138) are left out — display is
  "unavailable on this box"; ``visualize_live=True`` raises when the first frame would be drawn
* ``devices=[...]`` (new, optional) shards chunks of frames across several B200s through
  ``sharding.ThreadedSharder`` and still yields in frame order. Each GPU worker feeds the frames AS CAPTURED
  (3-channel, full resolution) to ``b200s_pipeline_frames_submit``: INTER_AREA downscale, size fix-up, gray conversion,
  remap, matcher, post-filters and depth all run on that worker's device, up to three chunks in flight per GPU, page-
  locked chunk buffers, and only the depth plane (what this generator yields, :103,:130) comes back over PCIe.
"""
from __future__ import annotations

import threading
import time
import warnings
from concurrent.futures import ThreadPoolExecutor
from typing import Optional, Sequence

import numpy as np

from ..sharding import ThreadedSharder
from .input import _area_downscale, stereo_stream
from .stereo_core import StereoCore
from .threaded_stereo import ThreadedStereoCapture
from .visualizations import visualize_depth_live, visualize_depth_live_gray


def _is_frame_sequence(src) -> bool:
    """Neither a path nor a camera index but an iterable of frames."""
    import os
    return not isinstance(src, (str, bytes, int, os.PathLike)) and hasattr(src, "__iter__")


class _FrameWorker:
    """One GPU of the sharded generator: engine + maps + a ring of page-locked chunk buffers (sharding.ThreadedSharder's
    submit / wait protocol). Two host threads per GPU: a feeder (copies the chunk's frames into page-locked memory and
    enqueues upload + kernels + download) and the sharder's worker thread (waits for a chunk, copies the depth maps into
    an array the caller owns); the frame copies themselves fan out over a small pool (numpy releases the GIL)."""
    max_inflight = 3

    def __init__(self, device, matcher_kw, maps, pp, dp, use_maps, size, area, cap=1, copy_threads=8):
        from ..engine import StereoEngine
        self.eng = StereoEngine(device, **matcher_kw)
        if maps is not None:
            self.eng.set_maps(maps['map1_L'], maps['map2_L'], maps['map1_R'], maps['map2_R'])
        self.pp, self.dp, self.use_maps, self.size, self.area = pp, dp, use_maps, size, area
        self.want_depth = bool(dp.enabled)
        self._free = []          # page-locked (L, R, out) triples of the current chunk geometry
        self._geom = None
        self._sized = None
        self._cap = max(1, int(cap))   # frames per buffer: the full chunk, also for the shorter first and last chunks
        self._lock = threading.Lock()
        self._pool = ThreadPoolExecutor(max(1, int(copy_threads)))
        self._feeder = ThreadPoolExecutor(1)

    def close(self):
        self._feeder.shutdown(wait=True)
        self._pool.shutdown(wait=True)
        self._free = []
        self.eng.close()

    def _buffers(self, B, shape):
        from ..engine import pinned_empty
        with self._lock:
            if self._geom is None or B > self._geom[0] or tuple(shape) != self._geom[1]:
                self._free, self._geom = [], (max(B, self._cap), tuple(shape))   # buffers in flight stay alive through their tokens
            geom = self._geom
            if self._free:
                return self._free.pop(), geom
        aW, aH = self.area if self.area is not None else (shape[1], shape[0])
        W, H = self.size if self.size is not None else (aW, aH)
        Wc = W - self.eng.params.numDisparities
        cap = geom[0]
        return (pinned_empty((cap,) + tuple(shape), np.uint8), pinned_empty((cap,) + tuple(shape), np.uint8),
                pinned_empty((cap, H, Wc), np.float32)), geom

    def _feed(self, Ls, Rs):
        B, shape = len(Ls), Ls[0].shape
        if any(f.shape != shape or f.dtype != np.uint8 for f in Ls) or any(f.shape != shape or f.dtype != np.uint8 for f in Rs):
            raise ValueError("sharded video path: every frame of a stream must be uint8 and of one size")
        bufs, geom = self._buffers(B, shape)
        bL, bR, out = bufs
        if self._sized != geom:
            # size the engine's device scratch for a FULL chunk once: the first chunks of a stream are shorter (ramp),
            # and growing the scratch later frees and reallocates gigabytes in the middle of the stream
            self._sized = geom
            # ... and page-lock the other chunk buffers a stream will have in flight now, not in the middle of a later run
            spare = [self._buffers(B, shape)[0] for _ in range(self.max_inflight - 1)]
            with self._lock:
                if geom == self._geom:
                    self._free.extend(spare)
            if B < geom[0]:
                bL[B:] = 0; bR[B:] = 0
                self.eng.pipeline_batch_wait(self.eng.pipeline_frames_submit(
                    bL, bR, self.pp, self.dp, size=self.size, area=self.area, use_maps=self.use_maps,
                    want_disp=not self.want_depth, want_depth=self.want_depth,
                    out_disp=None if self.want_depth else out, out_depth=out if self.want_depth else None))

        def put(k):
            np.copyto(bL[k], Ls[k]); np.copyto(bR[k], Rs[k])
        list(self._pool.map(put, range(B)))
        want_depth = self.want_depth
        tk = self.eng.pipeline_frames_submit(
            bL[:B], bR[:B], self.pp, self.dp, size=self.size, area=self.area, use_maps=self.use_maps,
            want_disp=not want_depth, want_depth=want_depth,
            out_disp=None if want_depth else out[:B], out_depth=out[:B] if want_depth else None)
        return tk, bufs, B, geom

    def submit(self, Ls, Rs):
        return self._feeder.submit(self._feed, Ls, Rs)          # chunks are fed in submission order

    def wait(self, token):
        tk, bufs, B, geom = token.result()
        self.eng.pipeline_batch_wait(tk)
        src = bufs[2]
        res = np.empty((B,) + src.shape[1:], np.float32)      # the caller owns what the generator yields

        def get(k):
            np.copyto(res[k], src[k])
        list(self._pool.map(get, range(B)))
        with self._lock:
            if geom == self._geom:
                self._free.append(bufs)
        return (None, res) if self.want_depth else (res, None)


class StereoDepthEstimatorVideo:
    def __init__(self, left_source=None, right_source=None, downscale_factor=1.0, visualize_live=False,
                 saving_path=None, fast_mode=False, use_threading=True, target_fps=30, drop_frames=False,
                 visualize_gray=False, devices: Optional[Sequence[int]] = None, chunk: Optional[int] = None,
                 max_chunk: int = 256) -> None:
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
        self.chunk = chunk              # frames per GPU call; None: the engine's preferred batch for the frame size,
        self.max_chunk = max_chunk      # bounded by max_chunk (the latency knob: a frame waits for its chunk to fill)
        self.core = StereoCore(downscale_factor=downscale_factor, fast_mode=fast_mode,
                               device=self.devices[0] if self.devices else 0)

    def configure_sgbm(self, **kwargs):
        self.core.configure_sgbm(**kwargs)

    def close_workers(self):
        """Give back what the sharded generator keeps between calls (per GPU: an engine handle with its device scratch and
        the page-locked chunk buffers). Called when the configuration changes; safe to call at any time between runs."""
        for ws in getattr(self, "_worker_pool", {}).values():
            for w in ws:
                w.close()
        self._worker_key, self._worker_pool = None, {}

    # -- frame sources ----------------------------------------------------------------------------
    def _frames(self, downscale_factor=None):
        f = self.downscale_factor if downscale_factor is None else downscale_factor
        if _is_frame_sequence(self.left_source) and _is_frame_sequence(self.right_source):
            # in-memory sources (new): two iterables of frames, e.g. lists of HxWx3 arrays from another decoder
            for left, right in zip(self.left_source, self.right_source):
                yield _area_downscale(left, f, self.core.device), _area_downscale(right, f, self.core.device)
            return
        if self.use_threading:
            cap = ThreadedStereoCapture(self.left_source, self.right_source, downscale_factor=f,
                                        drop_frames=self.drop_frames, device=self.core.device)
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
            yield from stereo_stream(self.left_source, self.right_source, downscale_factor=f, device=self.core.device)

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
        # the batched multi-GPU path runs the fused pipeline only; anything it cannot express (Telea hole filling on the
        # host, a replaced compute_disparity / _process_pair) goes frame by frame through StereoCore like the reference
        if self.devices and self.core._can_fuse(None, None) and (len(self.devices) > 1 or self.chunk is None or self.chunk > 1):
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
        core = self.core
        p = core.get_sgbm_params()
        use_maps = core._calibration()
        maps = None
        if use_maps:
            if core._rect_cache.engine is None:
                core._rect_cache.engine = core.sgbm.engine
            maps = core._rect_cache.get_maps(p['cam_matrix_L'], p['cam_matrix_R'], p['baseline'], p['image_width'],
                                             p['image_height'], p.get('dist_coeff_L'), p.get('dist_coeff_R'),
                                             p.get('rotation'), p.get('translation'), 1.0)
        pp, dp, kw = core._post_params(), core._depth_params(), core._matcher_kwargs()
        # frames are taken as captured (downscale_factor = 1 in the capture); the INTER_AREA downscale of
        # input.py:39-43 / threaded_stereo.py:41-46 runs on the worker's device inside the pipeline call
        frames = iter(self._frames(downscale_factor=1.0))
        first = next(frames, None)
        if first is None:
            return
        sh, sw = first[0].shape[:2]
        f = self.downscale_factor
        area = (int(sw * f), int(sh * f)) if f != 1.0 else None
        aw, ah = area if area is not None else (sw, sh)
        size = (int(p['image_width']), int(p['image_height'])) if use_maps else None
        if size is not None and size != (aw, ah):
            warnings.warn("Input image size %sx%s does not match calibration %sx%s; resizing for rectification." %
                          (aw, ah, size[0], size[1]), RuntimeWarning, stacklevel=2)
        W, H = size if size is not None else (aw, ah)
        n = None
        try:
            n = min(len(self.left_source), len(self.right_source))
        except TypeError:
            pass
        # workers (engine handle with its device scratch, uploaded maps, page-locked chunk buffers) survive between
        # calls of this generator as long as the configuration is the same. The chunk CAPACITY (what their buffers and
        # scratch are sized for) is part of what survives: the engine's preferred batch depends on the free device memory,
        # which the workers' own scratch changes, so it is asked for once per configuration and not on every call
        key = (tuple(sorted((k, str(v)) for k, v in kw.items())), id(maps), size, area, bytes(pp), bytes(dp),
               self.chunk, int(self.max_chunk), tuple(self.devices), (sh, sw))
        if getattr(self, "_worker_key", None) != key:
            self.close_workers()
            cap = self.chunk
            if cap is None:
                cap = max(1, min(core.sgbm.engine.preferred_batch(H, W), int(self.max_chunk)))
            self._worker_key, self._worker_pool, self._worker_cap = key, {}, int(cap)
        cap = chunk = self._worker_cap
        if self.chunk is None and n:
            # a source of known length (a list of frames, a short clip): equal chunks, the same number for every GPU
            # (chunk i goes to GPU i mod G) and at least two each, so that its copies overlap its kernels
            G = len(self.devices)
            nchunks = max(2 * G, -(-n // cap))
            nchunks = -(-nchunks // G) * G
            chunk = max(1, -(-n // nchunks))
        # shorter first chunks start the GPUs before a full chunk has been captured; frames that are all there already
        # (several GPUs, known length) are split evenly instead
        ramp = n is None or len(self.devices) == 1
        pool, used, lock = self._worker_pool, [], threading.Lock()

        def make_worker(dev):
            with lock:
                w = pool.get(dev, []).pop() if pool.get(dev) else None
            if w is None:
                w = _FrameWorker(dev, kw, maps, pp, dp, use_maps, size, area, cap=cap)
            with lock:
                used.append((dev, w))
            return w

        def pairs():
            yield first
            yield from frames

        t0 = time.time()
        try:
            for _disp, depth_m in ThreadedSharder(self.devices, make_worker, chunk, max_inflight=4, ramp=ramp).run(pairs()):
                yield depth_m
                self._show(depth_m)
                t0 = self._pace(t0)
        finally:
            with lock:
                for dev, w in used:
                    pool.setdefault(dev, []).append(w)
