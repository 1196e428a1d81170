"""GPU parity: the CUDA matcher (through the C ABI) against the C oracle and the golden vectors.
Bit-exact is the bar for int16 disparity (BASELINE.json north_star)."""
import numpy as np
import pytest

import oracle
from conftest import golden_cases
from helpers import golden_sgbm_kwargs
from depthestimation_b200.synth import make_pair, make_adversarial

pytestmark = pytest.mark.gpu

MODES = ["sgbm", "hh", "sgbm_3way", "hh4"]


@pytest.fixture(scope="module")
def engine():
    from depthestimation_b200.engine import StereoEngine
    e = StereoEngine(0)
    yield e
    e.close()


def _check(engine, L, R, stages=True, **kw):
    """Product flow (fused row kernel) against the oracle's disparity; then the debug flow (generic
    kernels, full S materialised) against the oracle's C / S / raw stages."""
    engine.configure(**kw)
    engine.set_debug_flow(False)
    got = engine.compute(L, R)
    ref = oracle.sgbm_compute(L, R, oracle.make_params(**kw), stages=stages)
    H, W = L.shape
    d = ref["disp"] if stages else ref
    assert np.array_equal(got, d), f"disp: {(got != d).sum()} px differ {kw}"
    if stages and ref["raw"] is not None:
        raw = engine.debug_stage("raw", H, W)
        assert np.array_equal(raw, ref["raw"]), f"raw(fused): {(raw != ref['raw']).sum()} px differ {kw}"
    engine.set_debug_flow(True)
    got = engine.compute(L, R)
    engine.set_debug_flow(False)
    if stages and ref["C"] is not None:
        C = engine.debug_stage("C", H, W)
        assert np.array_equal(C, ref["C"]), f"C: {(C != ref['C']).sum()} cells differ {kw}"
        S = engine.debug_stage("S", H, W)
        assert np.array_equal(S, ref["S"]), f"S: {(S != ref['S']).sum()} cells differ {kw}"
        raw = engine.debug_stage("raw", H, W)
        assert np.array_equal(raw, ref["raw"]), f"raw: {(raw != ref['raw']).sum()} px differ {kw}"
    d = ref["disp"] if stages else ref
    assert np.array_equal(got, d), f"disp: {(got != d).sum()} px differ {kw}"


@pytest.mark.parametrize("name", golden_cases())
def test_golden_disp16(engine, name, golden):
    g = golden(name)
    kw = golden_sgbm_kwargs(g)
    engine.configure(**kw)
    got = engine.compute(g["left_rect"], g["right_rect"])
    assert np.array_equal(got, g["disp16"]), f"{(got != g['disp16']).sum()} px differ"


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("D,bs", [(64, 5), (48, 3), (32, 9), (16, 1), (40, 11), (44, 5), (62, 7), (128, 5), (140, 5)])
def test_modes_blocks(engine, mode, D, bs):
    L, R, _ = make_pair(64, D + 120, D, seed=1000 + D + bs)
    _check(engine, L, R, min_disp=0, num_disp=D, block_size=bs, sgbm_mode=mode)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("mind", [-24, -8, 3, 8])
def test_min_disp(engine, mode, mind):
    L, R, _ = make_pair(56, 170, 32, seed=77)
    _check(engine, L, R, min_disp=mind, num_disp=32, block_size=5, sgbm_mode=mode)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("uniq,d12,cap,sws", [(0, 1, 31, 0), (30, -1, 5, 50), (75, 3, 63, 20), (10, 0, 20, 100)])
def test_filters(engine, mode, uniq, d12, cap, sws):
    L, R, _ = make_pair(60, 160, 40, seed=5)
    _check(engine, L, R, min_disp=0, num_disp=40, block_size=5, sgbm_mode=mode, uniqueness_ratio=uniq,
           disp12_max_diff=d12, prefilter_cap=cap, speckle_window_size=sws, speckle_range=1)


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("kind", ["noise", "zeros", "quantized", "periodic"])
def test_adversarial_inputs(engine, mode, kind):
    L, R = make_adversarial(kind, 48, 150)
    _check(engine, L, R, min_disp=0, num_disp=32, block_size=5, sgbm_mode=mode)


@pytest.mark.parametrize("mode", MODES)
def test_kitti_shape(engine, mode):
    L, R, _ = make_pair(375, 1242, 128, seed=1000)
    _check(engine, L, R, stages=False, min_disp=0, num_disp=128, block_size=5, sgbm_mode=mode)


def test_ragged_and_tiny(engine):
    # widths/heights that are not multiples of any tile size, image barely wider than D, 1-row image
    for (H, W, D) in [(37, 131, 24), (5, 70, 64), (1, 90, 16), (9, 41, 40), (33, 300, 8)]:
        L, R, _ = make_pair(H, W, D, seed=H * W)
        for mode in MODES:
            _check(engine, L, R, min_disp=0, num_disp=D, block_size=3, sgbm_mode=mode)


def test_no_full_range_column(engine):
    L, R, _ = make_pair(20, 50, 16, seed=3)
    engine.configure(num_disp=64, sgbm_mode="sgbm")
    got = engine.compute(L, R)
    assert np.all(got == -16)


def test_strided_input(engine):
    L, R, _ = make_pair(40, 200, 32, seed=11)
    big = np.zeros((40, 512), np.uint8)
    big2 = np.zeros((40, 512), np.uint8)
    big[:, 7:207] = L
    big2[:, 7:207] = R
    engine.configure(num_disp=32, sgbm_mode="sgbm")
    assert np.array_equal(engine.compute(big[:, 7:207], big2[:, 7:207]), engine.compute(L, R))


def test_invalid_params(engine):
    with pytest.raises(ValueError):
        engine.configure(num_disp=0)
    with pytest.raises(ValueError):
        engine.configure(block_size=4)
    with pytest.raises(ValueError):
        engine.configure(block_size=13)


@pytest.mark.parametrize("mode", MODES)
def test_max_disparity_range(engine, mode):
    """numDisparities = 256 (the engine's limit, BASELINE configs #3 / #5) against the oracle."""
    L, R, _ = make_pair(72, 640, 256, seed=77)
    _check(engine, L, R, stages=False, min_disp=0, num_disp=256, block_size=5, sgbm_mode=mode)
    # D between the lane layouts, incl. round_up(D, 8) % 16 == 8 (a 16-disparity lane must not straddle the pitch)
    for D, seed in [(200, 78), (136, 79), (72, 80), (66, 81)]:
        L, R, _ = make_pair(40, 420, D, seed=seed)
        _check(engine, L, R, min_disp=0, num_disp=D, block_size=3, sgbm_mode=mode)


def test_baseline_config_shapes_vs_oracle(engine):
    """BASELINE.json configs at their own sizes where the CPU oracle finishes in seconds:
    #1 example_stereo after its 0.5 downscale (994x1482, D=140, sgbm_3way) and #4 video frames (720x1280, D=128, sgbm_3way)."""
    for (H, W, D, seed) in [(994, 1482, 140, 1001), (720, 1280, 128, 1004)]:
        L, R, _ = make_pair(H, W, D, seed=seed)
        _check(engine, L, R, stages=False, min_disp=0, num_disp=D, block_size=5, sgbm_mode="sgbm_3way")


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("D", [70, 89, 100, 140, 150])
def test_lane_groups_follow_d(engine, mode, D):
    """Disparity counts whose lane count is not a power of two (5, 6, 7, 9, 10 lanes of 16 disparities: a warp then
    carries 6, 5, 4, 3, 3 lines instead of rounding up to 8 or 16 lanes): D = 140 is what example_stereo.py produces, 89
    what the video path's re-scaling quirk yields. All four modes, every kernel family (cluster
    kernel, generic line scan, row kernel), fused and debug flow with stage taps, against the oracle."""
    L, R, _ = make_pair(40, D + 90, D, seed=4000 + D)
    _check(engine, L, R, min_disp=0, num_disp=D, block_size=5, sgbm_mode=mode)
    from depthestimation_b200 import _lib
    engine.configure(min_disp=0, num_disp=D, block_size=5, sgbm_mode=mode)
    engine.compute(L, R)
    if mode in ("sgbm", "hh"):
        assert engine.last_flow & _lib.FLOW_VERT_DOWN, "the cluster kernel must take these shapes, not the fallback"


def _flows_agree(engine, H, W, D, mode, seed):
    """Size-independent property for shapes the scalar oracle cannot do in seconds: the product flow (cluster kernel +
    fused row kernel, S never materialised) and the debug flow (one generic kernel per direction + a selection kernel
    on the full S, each diffed against the oracle stage by stage at small sizes) are two implementations of the same
    spec and must give the same int16 image; plus the invariants of A.6/A.8 on the result."""
    L, R, _ = make_pair(H, W, D, seed=seed)
    engine.configure(min_disp=0, num_disp=D, block_size=5, sgbm_mode=mode)
    engine.set_debug_flow(False)
    a = engine.compute(L, R)
    again = engine.compute(L, R)
    engine.set_debug_flow(True)
    b = engine.compute(L, R)
    engine.set_debug_flow(False)
    assert np.array_equal(a, again), "not deterministic"
    assert np.array_equal(a, b), f"{(a != b).sum()} px differ between the fused and the generic flow"
    assert np.all(a[:, :D] == -16), "columns without a full disparity range must be invalid"
    valid = a[a != -16]
    assert valid.size > 0.5 * H * (W - D)
    assert valid.min() >= 0 and valid.max() <= (D - 1) * 16


def test_full_size_1080p_hh_d256(engine, monkeypatch):
    """BASELINE config #3, one pair. A single pair on this shape's clamped-record cluster plan goes to the generic line
    kernel (vert.cu: under a third of the SMs busy); with the rule switched off the cluster kernel itself runs: both
    must agree with the debug flow."""
    from depthestimation_b200 import _lib
    _flows_agree(engine, 1080, 1920, 256, "hh", 1003)
    engine.compute(*make_pair(1080, 1920, 256, seed=1003)[:2])
    assert engine.last_flow & _lib.FLOW_PATH_VERTICAL and not engine.last_flow & _lib.FLOW_VERT_DOWN
    monkeypatch.setenv("B200S_VERT_SMALL_GENERIC", "0")
    _flows_agree(engine, 1080, 1920, 256, "hh", 1003)
    engine.compute(*make_pair(1080, 1920, 256, seed=1003)[:2])
    assert engine.last_flow & _lib.FLOW_VERT_DOWN and engine.last_flow & _lib.FLOW_VERT_UP


def test_full_size_4k_d256(engine):
    _flows_agree(engine, 2160, 3840, 256, "sgbm", 1005)        # BASELINE config #5 (one pair of the 64)


def _full_size_vs_cv2(engine, H, W, D, mode, seed, **extra):
    """The two largest BASELINE configurations against the REAL oracle at full size: cv2.StereoSGBM.compute on the host
    (cv2 is on the GPU box; seconds per frame), then the whole per-frame pipeline (crop, speckles, outliers, median,
    depth) against the same chain of cv2 calls. Volumes here are 0.9 - 4 GB per frame: pitch, padding and 32-bit offset
    mistakes show up at this size and nowhere else."""
    pytest.importorskip("cv2")
    from oracle import cv2_reference as cr
    from depthestimation_b200.engine import make_depth_params, make_post_params
    L, R, _ = make_pair(H, W, D, seed=seed)
    kw = dict(min_disp=0, num_disp=D, block_size=5, sgbm_mode=mode, **extra)
    engine.configure(**kw)
    engine.set_debug_flow(False)
    got = engine.compute(L, R)
    want = cr.compute_disparity16(L, R, **kw)
    assert np.array_equal(got, want), f"{(got != want).sum()} of {got.size} px differ from cv2 ({mode} {W}x{H} D={D})"
    pp = make_post_params(False, 100, 1.0, 2.5, True)
    dp = make_depth_params(0.8 * W, 0.54, 0.0, eps=0, max_depth=100.0)
    disp, depth, _, _ = engine.pipeline(L, R, pp, dp)
    wd, wz = cr.Cv2StereoPipeline(kw, None, 1.0, False, 0.8 * W, 0.54, 0.0, 100.0)(L, R)
    assert np.array_equal(disp, wd), f"{(disp != wd).sum()} px of the post-processed disparity differ"
    fin = np.isfinite(wz)
    assert np.array_equal(np.isfinite(depth), fin) and np.allclose(depth[fin], wz[fin], rtol=1e-4, atol=0)


@pytest.mark.parametrize("small_generic", ["1", "0"])
def test_full_size_1080p_hh_d256_vs_cv2(engine, monkeypatch, small_generic):
    """BASELINE config #3: 1920x1080, D = 256, 8-path (hh) + LR check + speckle filter; through the generic line kernel
    (what a single pair of this shape takes) and through the clamped-record cluster kernel (what a batch takes)."""
    monkeypatch.setenv("B200S_VERT_SMALL_GENERIC", small_generic)
    _full_size_vs_cv2(engine, 1080, 1920, 256, "hh", 1003, disp12_max_diff=1, speckle_window_size=100, speckle_range=2)
    from depthestimation_b200 import _lib
    assert bool(engine.last_flow & _lib.FLOW_VERT_DOWN) == (small_generic == "0")


@pytest.mark.parametrize("mode", ["sgbm_3way", "sgbm"])
def test_full_size_4k_d256_vs_cv2(engine, mode):
    """BASELINE config #5 (one pair of the 64): 3840x2160, D = 256, in the mode tools/bench_configs.py measures
    (sgbm_3way, vertical path through the generic kernel: the cluster kernel's 31-bit offset guard) and in sgbm."""
    _full_size_vs_cv2(engine, 2160, 3840, 256, mode, 1005)


_CLAMP_SCRIPT = r"""
import sys, numpy as np
sys.path.insert(0, {root!r})
import oracle
from depthestimation_b200.engine import StereoEngine
from depthestimation_b200.synth import make_pair
e = StereoEngine(0)
bad = 0
for mode in ("sgbm", "hh"):
    for (H, W, D, seed) in [(36, 360, 256, 91), (33, 353, 250, 92)]:
        L, R, _ = make_pair(H, W, D, seed=seed)
        kw = dict(min_disp=0, num_disp=D, block_size=3, sgbm_mode=mode)
        e.configure(**kw)
        got = e.compute(L, R)
        ref = oracle.sgbm_compute(L, R, oracle.make_params(**kw))
        bad += int((got != ref).sum())
        if not e.last_flow & 1 or e.last_flow & 4:        # B200S_FLOW_VERT_DOWN / B200S_FLOW_PATH_VERTICAL
            bad += 1000000
e.close()
print("MISMATCH", bad)
"""


def test_vertical_kernel_clamped_records():
    """k_vert<.., CL = true> (strip state sized NC + 1 records, surplus groups redirected to the last column) against
    the oracle. The planner only picks it for shapes like 1920x1080 at D = 256; forcing one strip per frame
    (B200S_VERT_CS=1, read once per process, hence the subprocess) selects it at a size the oracle does in seconds:
    W1 = 104 columns, 4 slots of 32 groups = 129 records of 2080 B do not fit, 105 do."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = str(Path(__file__).resolve().parents[1])
    # (B200S_VERT_SMALL_GENERIC=0: a single small pair on a clamped plan would otherwise go to the generic line kernel)
    env = dict(os.environ, B200S_VERT_CS="1", B200S_VERBOSE="1", B200S_VERT_SMALL_GENERIC="0")
    r = subprocess.run([sys.executable, "-c", _CLAMP_SCRIPT.format(root=root)], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "clamped records" in r.stderr, "the clamped instantiation was not selected:\n" + r.stderr[-1000:]
    assert "MISMATCH 0" in r.stdout, r.stdout[-500:]
