"""Pins the C restatement (oracle/*.c) bit-exact against the live cv2 binary, called exactly the
way depthlib calls it (oracle/cv2_reference.py). Skipped where cv2 is not importable."""
import numpy as np
import pytest

import oracle
from oracle import cv2_reference as cr
from depthestimation_b200.synth import make_pair, make_adversarial, make_calibration

pytestmark = pytest.mark.skipif(not cr.HAVE_CV2, reason="cv2 not importable")

MODES = ["sgbm", "hh", "sgbm_3way", "hh4"]


def _check(L, R, **kw):
    ref = cr.compute_disparity16(L, R, **kw)
    got = oracle.sgbm_compute(L, R, oracle.make_params(**kw))
    bad = int((ref != got).sum())
    assert bad == 0, f"{bad} px differ for {kw}"


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("D,bs", [(64, 5), (48, 3), (32, 9), (16, 1), (40, 11), (44, 5), (62, 7)])
def test_modes_blocks(mode, D, bs):
    L, R, _ = make_pair(64, 180, D, seed=1000 + D + bs)
    _check(L, R, min_disp=0, num_disp=D, block_size=bs, sgbm_mode=mode)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("mind", [-24, -8, 3, 8])
def test_min_disp(mode, mind):
    L, R, _ = make_pair(56, 170, 32, seed=77)
    _check(L, R, min_disp=mind, num_disp=32, block_size=5, sgbm_mode=mode)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("uniq,d12,cap,sws", [(0, 1, 31, 0), (30, -1, 5, 50), (75, 3, 63, 20), (10, 0, 20, 100)])
def test_filters(mode, uniq, d12, cap, sws):
    L, R, _ = make_pair(60, 160, 40, seed=5)
    _check(L, R, min_disp=0, num_disp=40, block_size=5, sgbm_mode=mode, uniqueness_ratio=uniq,
           disp12_max_diff=d12, prefilter_cap=cap, speckle_window_size=sws, speckle_range=1)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("kind", ["noise", "zeros", "quantized", "periodic"])
def test_adversarial_inputs(mode, kind):
    L, R = make_adversarial(kind, 48, 150)
    _check(L, R, min_disp=0, num_disp=32, block_size=5, sgbm_mode=mode)


@pytest.mark.parametrize("mode", ["sgbm", "sgbm_3way"])
def test_kitti_shape_defaults(mode):
    L, R, _ = make_pair(375, 1242, 128, seed=1000)
    _check(L, R, min_disp=0, num_disp=128, block_size=5, sgbm_mode=mode)


def test_remap_gray_median_speckle_post_depth():
    import cv2
    rng = np.random.default_rng(3)
    H, W = 90, 210
    cal = make_calibration(H, W)
    maps = cr.build_maps(cal, alpha=1.0)
    img = rng.integers(0, 256, (H, W), dtype=np.uint8)
    assert np.array_equal(oracle.remap_u8(img, maps["map1_L"], maps["map2_L"]),
                          cv2.remap(img, maps["map1_L"], maps["map2_L"], cv2.INTER_LINEAR))
    col = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    assert np.array_equal(oracle.bgr2gray(col), cv2.cvtColor(col, cv2.COLOR_BGR2GRAY))
    L, R, _ = make_pair(H, W, 48, seed=9)
    d16 = cr.compute_disparity16(L, R, num_disp=48, sgbm_mode="sgbm", speckle_window_size=0)
    assert np.array_equal(oracle.median3_s16(d16), cv2.medianBlur(d16, 3))
    sp = d16.copy()
    cv2.filterSpeckles(sp, -16, 40, 16)
    assert np.array_equal(oracle.filter_speckles(d16, -16, 40, 16), sp)
    d = (d16.astype(np.float32) / 16.0)[:, 48:]
    assert np.array_equal(oracle.median3_f32(d), cv2.medianBlur(d, 3))
    assert np.array_equal(oracle.postprocess(d, 100, 1.0, 2.5), cr.postprocess_disparity(d, 100, 1.0, 2.5))
    post = oracle.postprocess(d, 100, 1.0, 2.5)
    for md in (None, 30.0):
        a = oracle.disp_to_depth(post, 0.8 * W, 0.54, 1.25, eps=0, max_depth=md)
        b = cr.disparity_to_depth(post, 0.8 * W, 0.54, 1.25, eps=0, max_depth=md)
        assert np.array_equal(a, b)


@pytest.mark.parametrize("H,W,f,cn", [(60, 90, 0.7, 3), (61, 97, 0.5, 3), (50, 83, 0.33, 1), (64, 100, 0.9, 3),
                                       (200, 302, 0.5, 3), (201, 303, 1 / 3, 3), (200, 304, 0.25, 1),
                                       (497, 741, 0.5, 3), (97, 131, 0.77, 1), (64, 64, 1.0, 3), (33, 40, 0.1, 3)])
def test_resize_area_matches_cv2(H, W, f, cn):
    """INTER_AREA downscale as depthlib/input.py:39-43 calls it (size = int(dim * factor)): 2x2 fast path,
    integer-factor fast path, and the weighted-table path, uint8 gray and 3-channel."""
    import cv2
    rng = np.random.default_rng(H * W + cn)
    img = rng.integers(0, 256, (H, W, cn) if cn > 1 else (H, W), dtype=np.uint8)
    dsize = (int(W * f), int(H * f))
    ref = cv2.resize(img, dsize, interpolation=cv2.INTER_AREA)
    assert np.array_equal(oracle.resize_area(img, dsize), ref)


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 7, 9, 11])
def test_fill_nearest_matches_cv2(k):
    """fill_holes(method='nearest') exactly as depthlib/postprocess.py:106-116 spells it with cv2 (ellipse element, k dilations
    under the hole mask), default mask (disparity <= 0) and an explicit one."""
    import cv2
    rng = np.random.default_rng(40 + k)
    H, W = 57, 83
    d = (rng.random((H, W)) * 60).astype(np.float32)
    d[rng.random((H, W)) < 0.35] = 0.0
    d[10:30, 20:50] = 0.0                      # a hole wider than the element
    d[0:3, :] = 0.0                            # holes on the border
    for mask in (None, rng.random((H, W)) < 0.3):
        m = (d <= 0) if mask is None else mask
        kernel = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (k, k))
        ref = d.copy()
        for _ in range(k):
            ref = np.where(m, cv2.dilate(ref, kernel), ref)
        assert np.array_equal(oracle.fill_nearest(d, mask, k), ref)


def test_resize_linear_matches_cv2():
    """INTER_LINEAR resize of uint8 images as rectify.py:92-105 calls it (the size fix-up): random up / down scales, the
    exact 2x case (which cv2 may route to INTER_AREA), identity, 1- and 3-channel."""
    import cv2
    rng = np.random.default_rng(11)
    cases = [(100, 200, 50, 100), (99, 200, 49, 100), (90, 120, 30, 40), (40, 50, 160, 200), (7, 9, 70, 90),
             (64, 64, 64, 64), (5, 5, 1, 1), (6, 8, 3, 4), (375, 1242, 370, 1226), (370, 1226, 375, 1242)]
    for _ in range(40):
        cases.append((int(rng.integers(5, 200)), int(rng.integers(5, 300)), int(rng.integers(3, 260)), int(rng.integers(3, 400))))
    for i, (sH, sW, dH, dW) in enumerate(cases):
        cn = 3 if i % 2 else 1
        img = rng.integers(0, 256, (sH, sW, cn) if cn == 3 else (sH, sW), dtype=np.uint8)
        ref = cv2.resize(img, (dW, dH), interpolation=cv2.INTER_LINEAR)
        got = oracle.resize_linear(img, (dW, dH))
        assert np.array_equal(got, ref), f"{(sH, sW)} -> {(dH, dW)} x{cn}: {(got != ref).sum()} of {ref.size} differ"


def test_maps_oracle_matches_cv2():
    """oracle/maps.py (numpy restatement of cv2.initUndistortRectifyMap, depthlib/rectify.py:68-73) against the live
    binary: five calibrations (synthetic, rotated / translated rig, 8-, 12- and 14-coefficient models), alpha 0 and 1.
    cv2's dispatched SIMD path accumulates the column coordinate several columns at a time (eight on an AVX-512 host); the
    restatement evaluates every pixel directly. Over 157 M map values of the builder's survey 9 differed, each by one float32 ulp:
    the bar here is at most one value per map pair, by at most one ulp, and none on the two standard cases."""
    import cv2
    from oracle import maps as om
    for name, KL, dL, KR, dR, size, R, T in om.calibration_cases():
        for alpha in (0.0, 1.0):
            rL, rR, pL, pR, *_ = cv2.stereoRectify(KL, dL, KR, dR, size, R, T, flags=cv2.CALIB_ZERO_DISPARITY, alpha=alpha)
            for K, d, r, p in ((KL, dL, rL, pL), (KR, dR, rR, pR)):
                want = cv2.initUndistortRectifyMap(K, d, r, p, size, cv2.CV_32FC1)
                got = om.init_undistort_rectify_map(K, d, r, p, size)
                bad = sum(int((g != w).sum()) for g, w in zip(got, want))
                assert bad <= (0 if name in ("synthetic", "rotated") else 1), (name, alpha, bad)
                for g, w in zip(got, want):
                    assert np.all(np.abs(g.astype(np.float64) - w) <= np.spacing(np.abs(w)).astype(np.float64))
