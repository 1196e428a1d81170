"""Shared helpers for the parity tests (golden-record decoding)."""
import numpy as np


def golden_sgbm_kwargs(g):
    mind, D, bs, d12, cap, uniq, sws, sr = (int(v) for v in g["params"])
    return dict(min_disp=mind, num_disp=D, block_size=bs, disp12_max_diff=d12, prefilter_cap=cap,
                uniqueness_ratio=uniq, speckle_window_size=sws, speckle_range=sr, sgbm_mode=str(g["mode"]))


def golden_max_depth(g):
    v = float(g["max_depth"])
    return None if np.isnan(v) else v
