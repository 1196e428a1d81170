#!/usr/bin/env python
"""Copy rates of the 4K job's host side in isolation: pageable frames -> page-locked chunk buffer (8 copy threads), page-locked
depth planes -> a fresh / a reused result array."""
import json
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from depthestimation_b200.engine import pinned_empty, StereoEngine  # noqa: E402

eng = StereoEngine(0)
B, H, W, Wc = 13, 2160, 3840, 3584
frames = [np.full((H, W), k, np.uint8) for k in range(B)]
bL = pinned_empty((B, H, W), np.uint8)
out = pinned_empty((B, H, Wc), np.float32)
out[:] = 1.0
pool = ThreadPoolExecutor(8)
res = {}
for rep in range(3):
    t0 = time.perf_counter()
    list(pool.map(lambda k: np.copyto(bL[k], frames[k]), range(B)))
    res[f"in_pool8_GBps_{rep}"] = round(bL.nbytes / (time.perf_counter() - t0) / 1e9, 2)
t0 = time.perf_counter()
for k in range(B):
    np.copyto(bL[k], frames[k])
res["in_1thr_GBps"] = round(bL.nbytes / (time.perf_counter() - t0) / 1e9, 2)
for rep in range(3):
    t0 = time.perf_counter()
    r = np.empty((B, H, Wc), np.float32)
    list(pool.map(lambda k: np.copyto(r[k], out[k]), range(B)))
    res[f"out_fresh_pool8_GBps_{rep}"] = round(out.nbytes / (time.perf_counter() - t0) / 1e9, 2)
    del r
r = np.empty((B, H, Wc), np.float32)
r[:] = 0
for rep in range(2):
    t0 = time.perf_counter()
    list(pool.map(lambda k: np.copyto(r[k], out[k]), range(B)))
    res[f"out_reused_pool8_GBps_{rep}"] = round(out.nbytes / (time.perf_counter() - t0) / 1e9, 2)
t0 = time.perf_counter()
r2 = np.empty((B, H, Wc), np.float32)
r2[:] = 0
res["fresh_touch_1thr_GBps"] = round(out.nbytes / (time.perf_counter() - t0) / 1e9, 2)
print(json.dumps(res))
