#!/usr/bin/env python
"""Host-side probe of the in-process multi-GPU video driver (StereoDepthEstimatorVideo(devices=[...])):

(1) what first-touch page faults of freshly allocated result arrays cost on this host, alone and from 8 threads at once;
(2) the 240-frame 720p job of tools/multigpu_jobs.py on one GPU and, on a multi-GPU node, that job and the 64-pair 4K job on
    all GPUs, five runs each, with the wall time per phase of a worker (chunk buffers, copies into / out of page-locked
    memory, the submit call, the wait for the GPU), the count of page-locked allocations and workers built per run, and a
    per-chunk timeline (feed / wait, worker, frames, start and end in ms) of the last run.

    python tools/host_probe.py > gpurun_out/host_probe.txt"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def fault_probe():
    src = np.ones((15, 720, 1152), np.float32)
    out = {}
    for threads in (1, 8):
        def fresh(_):
            a = np.empty_like(src); np.copyto(a, src); return a.shape
        keep = [np.empty_like(src) for _ in range(threads)]
        for k in keep:
            k[:] = 0
        def reuse(i):
            np.copyto(keep[i % threads], src); return 0
        for name, fn in (("fresh", fresh), ("reused", reuse)):
            with ThreadPoolExecutor(threads) as ex:
                list(ex.map(fn, range(threads)))
                t0 = time.perf_counter()
                list(ex.map(fn, range(16)))
                dt = time.perf_counter() - t0
            out[f"{name}_{threads}thr_GBps"] = round(16 * src.nbytes / dt / 1e9, 2)
    return out


ACC = {}
EVENTS = []


def _instrument():
    """Accumulate wall time per phase of a worker: whole _feed (copy in + submit), the submit call alone, whole wait (GPU wait +
    copy out), the GPU wait alone."""
    import threading
    import depthestimation_b200.depthlib  # noqa: F401  (the package re-exports the class under the module's name)
    mod = sys.modules["depthestimation_b200.depthlib.StereoDepthEstimatorVideo"]
    from depthestimation_b200.engine import StereoEngine
    lock = threading.Lock()

    def wrap(cls, name, key):
        fn = getattr(cls, name)
        def timed(*a, **k):
            t0 = time.perf_counter()
            try:
                return fn(*a, **k)
            finally:
                with lock:
                    ACC[key] = ACC.get(key, 0.0) + time.perf_counter() - t0
        setattr(cls, name, timed)
    from concurrent.futures import ThreadPoolExecutor as TPE
    pmap = TPE.map
    def timed_map(self, fn, *it, **k):
        key = "copy_out_map" if "worker" in threading.current_thread().name else "copy_in_map"
        t0 = time.perf_counter()
        try:
            return list(pmap(self, fn, *it, **k))
        finally:
            with lock:
                ACC[key] = ACC.get(key, 0.0) + time.perf_counter() - t0
    TPE.map = timed_map
    import depthestimation_b200.engine as engmod
    pe = engmod.pinned_empty
    def counted(shape, dtype):
        with lock:
            ACC["pinned_allocs"] = ACC.get("pinned_allocs", 0) + 1
            ACC["pinned_GB"] = ACC.get("pinned_GB", 0.0) + float(np.prod(shape)) * np.dtype(dtype).itemsize / 1e9
        return pe(shape, dtype)
    engmod.pinned_empty = counted
    init = mod._FrameWorker.__init__
    def counted_init(self, *a, **k):
        with lock:
            ACC["workers_created"] = ACC.get("workers_created", 0) + 1
        return init(self, *a, **k)
    mod._FrameWorker.__init__ = counted_init
    feed0, wait0 = mod._FrameWorker._feed, mod._FrameWorker.wait
    def ev_feed(self, Ls, Rs):
        t0 = time.perf_counter()
        r = feed0(self, Ls, Rs)
        with lock:
            EVENTS.append(("feed", id(self), len(Ls), t0, time.perf_counter()))
        return r
    def ev_wait(self, tok):
        t0 = time.perf_counter()
        r = wait0(self, tok)
        with lock:
            EVENTS.append(("wait", id(self), r[1].shape[0] if r[1] is not None else r[0].shape[0], t0, time.perf_counter()))
        return r
    mod._FrameWorker._feed, mod._FrameWorker.wait = ev_feed, ev_wait
    wrap(mod._FrameWorker, "_buffers", "buffers")
    wrap(mod._FrameWorker, "_feed", "feed_total")
    wrap(mod._FrameWorker, "wait", "wait_total")
    wrap(StereoEngine, "pipeline_frames_submit", "submit_call")
    wrap(StereoEngine, "pipeline_batch_wait", "gpu_wait")


def job(nworkers, H=720, W=1280, D=128, n=240, cn=3, devices=None):
    from depthestimation_b200.depthlib import StereoDepthEstimatorVideo
    from depthestimation_b200.synth import make_calibration, make_pair
    d = []
    for k in range(2):
        l, r, _ = make_pair(H, W, D, seed=1200 + k)
        if cn == 3:
            l, r = np.ascontiguousarray(np.stack([l] * 3, -1)), np.ascontiguousarray(np.stack([r] * 3, -1))
        d.append((l, r))
    lefts, rights = [d[i % 2][0] for i in range(n)], [d[i % 2][1] for i in range(n)]
    cfg = dict(make_calibration(H, W)); cfg.update(num_disp=D, max_depth=100.0)
    vid = StereoDepthEstimatorVideo(lefts, rights, target_fps=0, devices=devices or [0] * nworkers)
    vid.configure_sgbm(**cfg)
    ts = []
    phases = []
    for _ in range(5):
        ACC.clear()
        EVENTS.clear()
        t0 = time.perf_counter()
        c = sum(1 for _ in vid.estimate_depth())
        ts.append(time.perf_counter() - t0)
        phases.append({k: round(v, 3) for k, v in sorted(ACC.items())})
        ids = {}
        timeline = [(k, ids.setdefault(w, len(ids)), b, round(1e3 * (a - t0), 1), round(1e3 * (e - t0), 1)) for k, w, b, a, e in sorted(EVENTS, key=lambda x: x[3])]
    assert c == n
    vid.close_workers()
    return {"shape": f"{W}x{H}", "frames": n, "workers": nworkers, "devices": devices or [0] * nworkers,
            "seconds": [round(t, 4) for t in ts], "phases_last": phases[-1], "timeline_last_ms": timeline}


if __name__ == "__main__":
    print(json.dumps({"cpus": os.cpu_count(), **fault_probe()}), flush=True)
    import torch
    _instrument()
    nd = torch.cuda.device_count()
    print(json.dumps(job(1)), flush=True)
    if nd > 1:
        print(json.dumps(job(nd, devices=list(range(nd)))), flush=True)
        print(json.dumps(job(nd, 2160, 3840, 256, 64, 1, devices=list(range(nd)))), flush=True)
