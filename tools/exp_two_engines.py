#!/usr/bin/env python
"""Experiment: N independent engines (own stream + scratch each) running the device-resident pipeline concurrently
on one GPU, against one engine. Does giving the GPU two kernel queues fill the gaps (tails, small kernels, the
ALU-bound cost kernel beside the DRAM-bound row kernel)?   python tools/exp_two_engines.py [--engines 2] [--batch 75]"""
import argparse, sys, threading, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench as B


def main():
    import torch
    from depthestimation_b200.engine import StereoEngine, make_post_params, make_depth_params
    ap = argparse.ArgumentParser()
    ap.add_argument("--engines", type=int, default=2)
    ap.add_argument("--batch", type=int, default=75)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--stagger-ms", type=float, default=0.0, help="start engine i that many ms after engine i-1")
    a = ap.parse_args()
    H, W, D = B.H, B.W, B.D
    L, R = B.make_inputs(a.batch)
    maps, cal = B.make_maps()
    pp = make_post_params(False, 100, 1.0, 2.5, True)
    dp = make_depth_params(cal["focal_length"], cal["baseline"], 0.0, eps=0, max_depth=None)
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    engs = []
    for i in range(a.engines):
        e = StereoEngine(0, **B.SGBM_KW)
        e.set_maps(maps["map1_L"], maps["map2_L"], maps["map1_R"], maps["map2_R"])
        s = torch.cuda.Stream()
        e.set_stream(s.cuda_stream)
        disp = torch.empty((a.batch, H, W - D), dtype=torch.float32, device="cuda")
        depth = torch.empty_like(disp)
        engs.append((e, s, disp, depth))

    def run(e, s, disp, depth, n, delay=0.0):
        time.sleep(delay)
        for _ in range(n):
            e.pipeline_batch_dev(dL.data_ptr(), dR.data_ptr(), a.batch, H, W, pp, dp, disp.data_ptr(), depth.data_ptr(), use_maps=True)

    for k in (1, a.engines):
        for e, s, disp, depth in engs[:k]:
            run(e, s, disp, depth, 3)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        th = [threading.Thread(target=run, args=(*engs[i], a.steps, i * a.stagger_ms * 1e-3)) for i in range(k)]
        [t.start() for t in th]
        [t.join() for t in th]
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"{k} engine(s): {k * a.steps * a.batch / dt:.0f} pairs/s ({dt * 1e3 / a.steps:.2f} ms per round of {k} x {a.batch} pairs)")
    same = all(torch.equal(engs[0][2], x[2]) for x in engs[1:])
    print("results identical across engines:", same)


if __name__ == "__main__":
    main()
