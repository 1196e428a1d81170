#!/usr/bin/env python
"""Pairs/s and per-call latency against the number of pairs per call (KITTI shape, `sgbm` and `sgbm_3way`), device-resident
and end to end from page-locked host buffers: one JSON line per (mode, B).

    python tools/batch_curve.py [--batches 1,2,4,8,16,32,75,150] [--modes sgbm,sgbm_3way]

device: b200s_pipeline_batch_dev (rectify -> match -> post -> depth), calls issued back to back on one stream, CUDA events;
latency_p50_ms: one call, synchronised, median of 15; e2e: b200s_pipeline_frames_submit / _wait, up to four calls in
flight, float32 disparity + depth back (wall clock)."""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
H, W, D = 375, 1242, 128


def run(mode, B, torch):
    from depthestimation_b200.depthlib.rectify import RectificationCache
    from depthestimation_b200.engine import StereoEngine, make_depth_params, make_post_params, pinned_empty
    from depthestimation_b200.synth import make_calibration, make_pair
    eng = StereoEngine(0, num_disp=D, sgbm_mode=mode)
    cal = make_calibration(H, W)
    maps = RectificationCache(engine=eng).get_maps(cal["cam_matrix_L"], cal["cam_matrix_R"], cal["baseline"], W, H,
                                                   cal["dist_coeff_L"], cal["dist_coeff_R"], None, None, 1.0)
    eng.set_maps(maps["map1_L"], maps["map2_L"], maps["map1_R"], maps["map2_R"])
    pp, dp = make_post_params(False, 100, 1.0, 2.5, True), make_depth_params(cal["focal_length"], cal["baseline"], 0.0, eps=0)
    pairs = [make_pair(H, W, D, seed=1000 + k)[:2] for k in range(min(4, B))]
    L = np.stack([pairs[i % len(pairs)][0] for i in range(B)])
    R = np.stack([pairs[i % len(pairs)][1] for i in range(B)])
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    Wc = W - D
    disp = torch.empty((B, H, Wc), dtype=torch.float32, device="cuda")
    depth = torch.empty_like(disp)

    def call():
        eng.pipeline_batch_dev(dL.data_ptr(), dR.data_ptr(), B, H, W, pp, dp, disp.data_ptr(), depth.data_ptr(), use_maps=True)

    for _ in range(3):
        call()
    torch.cuda.synchronize()
    lat = []
    for _ in range(15):
        t0 = time.perf_counter()
        call()
        stream.synchronize()
        lat.append(1e3 * (time.perf_counter() - t0))
    n = max(4, min(200, 600 // B))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(n):
        call()
    e1.record(stream)
    stream.synchronize()
    dev_fps = B * n / (e0.elapsed_time(e1) * 1e-3)
    # end to end
    sets = []
    for _ in range(4):
        nL, nR = pinned_empty(L.shape, np.uint8), pinned_empty(R.shape, np.uint8)
        nL[...] = L; nR[...] = R
        sets.append((nL, nR, pinned_empty((B, H, Wc), np.float32), pinned_empty((B, H, Wc), np.float32)))

    def e2e(m):
        fly = []
        for i in range(m):
            if len(fly) == 4:
                eng.pipeline_batch_wait(fly.pop(0))
            s = sets[i % 4]
            fly.append(eng.pipeline_frames_submit(s[0], s[1], pp, dp, use_maps=True, out_disp=s[2], out_depth=s[3]))
        for tk in fly:
            eng.pipeline_batch_wait(tk)
    e2e(6)
    t0 = time.perf_counter()
    e2e(n)
    e2e_fps = B * n / (time.perf_counter() - t0)
    ok = bool(np.array_equal(sets[0][2], disp.cpu().numpy()))
    eng.close()
    return {"mode": mode, "pairs_per_call": B, "device_pairs_per_s": round(dev_fps, 1), "latency_p50_ms": round(float(np.median(lat)), 3),
            "ms_per_pair_in_call": round(float(np.median(lat)) / B, 4), "e2e_pairs_per_s": round(e2e_fps, 1), "calls": n,
            "e2e_equals_device": ok, "workload": f"kitti_{W}x{H}_D{D}_{mode}_rectify+match+post+depth"}


def main():
    import torch
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1,2,4,8,16,32,75,150")
    ap.add_argument("--modes", default="sgbm,sgbm_3way")
    a = ap.parse_args()
    for mode in a.modes.split(","):
        for B in [int(b) for b in a.batches.split(",")]:
            print(json.dumps(run(mode, B, torch)), flush=True)


if __name__ == "__main__":
    main()
