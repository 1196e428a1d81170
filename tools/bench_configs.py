#!/usr/bin/env python
"""Device-resident throughput of every BASELINE.json configuration (bench.py measures configs[1] only; the
others are parity cases there). One JSON line per configuration:

    python tools/bench_configs.py [--only kitti,middlebury,...] [--steps K] [--warmup W]

Per configuration: synthetic pairs of the named shape (SURVEY.md §8(d) generator), the matcher mode the config
names, the whole StereoCore.estimate_depth path (rectify through cached maps -> match -> crop -> postprocess ->
depth) through b200s_pipeline_batch_dev, timed with CUDA events on the launching stream, plus the library's own
per-stage CUDA-event times and the algorithmic-bytes roofline of the matcher (10 B/cell for sgbm / sgbm_3way,
16 B/cell for hh / hh4)."""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

# name: (H, W, D, mode, frames in the job the config names (0 = a single pair; run with --batch 1 for its latency))
CONFIGS = {
    "example": (994, 1482, 140, "sgbm_3way", 0),     # configs[0]: example_stereo.py after its 0.5 downscale
    "kitti": (375, 1242, 128, "sgbm", 0),            # configs[1]
    "middlebury": (1080, 1920, 256, "hh", 0),        # configs[2]: 8-path + LR check + speckle filter
    "video720": (720, 1280, 128, "sgbm_3way", 240),  # configs[3]
    "uhd4k": (2160, 3840, 256, "sgbm_3way", 64),     # configs[4]
    "kitti_3way": (375, 1242, 128, "sgbm_3way", 0),  # the reference's default mode on the KITTI shape
    "kitti_hh": (375, 1242, 128, "hh", 0),
    "hd1080_sgbm": (1080, 1920, 128, "sgbm", 0),     # a wide frame at D = 128: two column slots per thread in the vertical cluster kernel
}
BYTES_PER_CELL = {"sgbm": 10, "sgbm_3way": 10, "hh": 16, "hh4": 16}


def run(name, steps, warmup, batch_override=0):
    import torch
    from depthestimation_b200.engine import StereoEngine, make_post_params, make_depth_params
    from depthestimation_b200.synth import make_pair, make_calibration
    from depthestimation_b200.depthlib.rectify import RectificationCache

    H, W, D, mode, nframes = CONFIGS[name]
    kw = dict(min_disp=0, num_disp=D, block_size=5, disp12_max_diff=1, prefilter_cap=31, uniqueness_ratio=10,
              speckle_window_size=50, speckle_range=2, sgbm_mode=mode)
    eng = StereoEngine(0, **kw)
    g = eng.geometry(W)
    V = H * g["W1"] * D
    B = batch_override or eng.preferred_batch(H, W)
    free, _ = torch.cuda.mem_get_info()
    per_frame = 2 * H * g["W1"] * g["Dp"] * 2 + H * W * 64
    B = max(1, min(B, int(0.6 * free / per_frame)))
    if nframes:
        B = min(B, nframes)
    pairs = [make_pair(H, W, D, seed=1000 + k)[:2] for k in range(min(4, B))]
    L = np.stack([pairs[i % len(pairs)][0] for i in range(B)])
    R = np.stack([pairs[i % len(pairs)][1] for i in range(B)])
    cal = make_calibration(H, W)
    maps = RectificationCache(engine=eng).get_maps(cal["cam_matrix_L"], cal["cam_matrix_R"], cal["baseline"], W, H,
                                                   cal["dist_coeff_L"], cal["dist_coeff_R"], None, None, 1.0)
    eng.set_maps(maps["map1_L"], maps["map2_L"], maps["map1_R"], maps["map2_R"])
    pp = make_post_params(False, 100, 1.0, 2.5, True)
    dp = make_depth_params(cal["focal_length"], cal["baseline"], 0.0, eps=0, max_depth=None)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    Wc = W - D
    with torch.cuda.stream(stream):
        dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
        disp = torch.empty((B, H, Wc), dtype=torch.float32, device="cuda")
        depth = torch.empty((B, H, Wc), dtype=torch.float32, device="cuda")

        def step():
            eng.pipeline_batch_dev(dL.data_ptr(), dR.data_ptr(), B, H, W, pp, dp, disp.data_ptr(), depth.data_ptr(),
                                   use_maps=True)
        for _ in range(warmup):
            step()
        stream.synchronize()
        eng.enable_timing(True)
        eng.stage_times(reset=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        stream.synchronize()
    ms = e0.elapsed_time(e1) / steps
    st = {k: v[0] / steps for k, v in eng.stage_times().items()}
    ln = {k: v[1] // steps for k, v in eng.stage_times().items()}
    eng.enable_timing(False)
    fps = B / ms * 1e3
    matcher_ms = sum(st[k] for k in ("prefilter", "cost", "paths", "wta", "median", "speckle"))
    peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("hbm_gbs", 6539.5)
    ach = BYTES_PER_CELL[mode] * V * B / (matcher_ms * 1e-3) / 1e9
    line = {"config": name, "shape": f"{W}x{H}", "D": D, "mode": mode, "pairs_per_step": B, "ms_per_step": ms,
            "frames_per_s": fps, "mde_per_s": fps * H * W * D / 1e6, "us_per_frame": 1e3 * ms / B,
            "stage_us_per_frame": {k: round(1e3 * v / B, 2) for k, v in st.items()}, "stage_launches": ln,
            "matcher_roofline": {"bytes_per_cell": BYTES_PER_CELL[mode], "achieved_gbs": ach, "peak_gbs": peak,
                                 "frac": ach / peak}}
    if nframes > 1:
        line["job_seconds_1gpu"] = nframes / fps
    print(json.dumps(line), flush=True)
    eng.close()
    del dL, dR, disp, depth
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0)
    a = ap.parse_args()
    names = [n for n in a.only.split(",") if n] or list(CONFIGS)
    for n in names:
        run(n, a.steps, a.warmup, a.batch)


if __name__ == "__main__":
    main()
