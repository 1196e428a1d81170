#!/usr/bin/env python
"""A few tiny pairs through every product kernel (all modes, exact and padded disparity counts, the pipeline tail),
checked against the oracle. Meant as the workload for `compute-sanitizer --tool memcheck|racecheck` where that tool is
available (it is closed on the shared B200 pool); on its own it is a 2-second end-to-end sanity run."""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import oracle
from depthestimation_b200.engine import StereoEngine, make_post_params, make_depth_params
from depthestimation_b200.synth import make_pair

e = StereoEngine(0)
bad = 0
for mode, H, W, D in [("sgbm", 24, 200, 128), ("sgbm_3way", 20, 230, 140), ("hh", 18, 120, 64), ("hh4", 16, 90, 44), ("sgbm_3way", 14, 100, 16)]:
    L, R, _ = make_pair(H, W, D, seed=5 + D)
    kw = dict(min_disp=0, num_disp=D, block_size=5, sgbm_mode=mode)
    e.configure(**kw)
    got = e.compute(L, R)
    ref = oracle.sgbm_compute(L, R, oracle.make_params(**kw))
    n = int((got != ref).sum())
    bad += n
    print(mode, H, W, D, "mismatching px:", n)
L, R, _ = make_pair(24, 200, 64, seed=3)
e.configure(min_disp=0, num_disp=64, block_size=5, sgbm_mode="sgbm")
disp, depth, _, _ = e.pipeline(L, R, make_post_params(False, 100, 1.0, 2.5, True), make_depth_params(160.0, 0.54, 0.0, eps=0))
print("pipeline ok", disp.shape, float(np.nanmax(disp)))
e.close()
sys.exit(1 if bad else 0)
