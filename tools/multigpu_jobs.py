#!/usr/bin/env python
"""The two multi-GPU jobs BASELINE.json names, end to end through the reference-shaped API, on 1, 2, 4 and 8 GPUs of one
node in ONE process (frames sharded chunk-per-GPU by sharding.ThreadedSharder, host gather in frame order, no collective):

  video720  configs[3]: 240 stereo frames 1280x720 (3-channel, as a capture delivers them), D = 128, default sgbm_3way,
            calibrated: rectify -> match -> post -> depth, one depth map per frame yielded in order
  uhd4k     configs[4]: a batch of 64 pairs 3840x2160, D = 256, sgbm_3way, the same full pipeline

    python tools/multigpu_jobs.py [--gpus 1,2,4,8] [--only video720,uhd4k]

One JSON line per (job, N): wall-clock seconds of the whole job from host frames to host depth maps (strong scaling: the
job is fixed), frames/s, and the speed-up over N = 1. Three warm-up runs per configuration (allocations), then the median of five."""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

JOBS = {"video720": (720, 1280, 128, 240, 3, None), "uhd4k": (2160, 3840, 256, 64, 1, None)}


def main():
    import torch
    from depthestimation_b200.depthlib import StereoDepthEstimatorVideo
    from depthestimation_b200.synth import make_calibration, make_pair
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2,4,8")
    ap.add_argument("--only", default="")
    ap.add_argument("--isolate", type=int, default=0, help="1: every (job, N) in a process of its own (no engines, page-locked "
                    "buffers or host threads left over from the previous configuration; costs a start-up per configuration: "
                    "minutes of an 8-GPU box)")
    a = ap.parse_args()
    ndev = torch.cuda.device_count()
    if a.isolate:
        import subprocess
        for name in [n for n in a.only.split(",") if n] or list(JOBS):
            base = None
            for n in [int(g) for g in a.gpus.split(",")]:
                if n > ndev:
                    continue
                out = subprocess.run([sys.executable, __file__, "--isolate", "0", "--only", name, "--gpus", str(n)],
                                     capture_output=True, text=True).stdout.strip().splitlines()
                if not out:
                    continue
                d = json.loads(out[-1])
                base = base or d["seconds"]
                d["speedup_vs_1gpu"] = round(base / d["seconds"], 2)
                d.pop("first_frame_equals_1gpu", None)
                print(json.dumps(d), flush=True)
        return
    for name in [n for n in a.only.split(",") if n] or list(JOBS):
        H, W, D, nframes, cn, chunk = JOBS[name]
        distinct = []
        for k in range(2):
            l, r, _ = make_pair(H, W, D, seed=1200 + k)
            if cn == 3:
                l, r = np.ascontiguousarray(np.stack([l] * 3, -1)), np.ascontiguousarray(np.stack([r] * 3, -1))
            distinct.append((l, r))
        lefts = [distinct[i % 2][0] for i in range(nframes)]
        rights = [distinct[i % 2][1] for i in range(nframes)]
        cfg = dict(make_calibration(H, W))
        cfg.update(num_disp=D, max_depth=100.0)
        base = None
        for n in [int(g) for g in a.gpus.split(",")]:
            if n > ndev:
                continue
            vid = StereoDepthEstimatorVideo(lefts, rights, target_fps=0, devices=list(range(n)), chunk=chunk)
            vid.configure_sgbm(**cfg)

            def run():
                cnt, first = 0, None
                for depth_m in vid.estimate_depth():
                    if first is None:
                        first = depth_m.copy()
                    cnt += 1
                return cnt, first
            for _ in range(3):       # which worker gets which chunks differs from run to run: let every worker see a full chunk
                run()
            times = []
            for _ in range(5):
                t0 = time.perf_counter()
                cnt, first = run()
                times.append(time.perf_counter() - t0)
            assert cnt == nframes
            t = float(np.median(times))
            if base is None:
                base, ref = t, first
            same = bool(np.array_equal(first, ref))
            vid.close_workers()     # give the device scratch and page-locked buffers back
            print(json.dumps({"job": name, "shape": f"{W}x{H}", "D": D, "mode": "sgbm_3way", "frames": nframes, "channels": cn,
                              "n_gpus": n, "seconds": round(t, 4), "frames_per_s": round(nframes / t, 1),
                              "speedup_vs_1gpu": round(base / t, 2), "first_frame_equals_1gpu": same,
                              "api": "StereoDepthEstimatorVideo(devices=[0..N-1]) generator, in-memory frame sources, depth plane only"}),
                  flush=True)
            del vid
            import gc
            gc.collect()


if __name__ == "__main__":
    main()
