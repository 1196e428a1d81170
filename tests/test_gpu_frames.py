"""GPU parity for the round-2 entry points: optional outputs (depth only / int16 disparity), the frame pipeline with
input preparation on the device (INTER_AREA -> INTER_LINEAR fix-up -> gray -> remap), batches at the bench's own size,
several handles in several threads / on several devices, and the video generator on top of them.
The checker is the live cv2 binary driven as depthlib drives it (oracle/cv2_reference.py) and the C oracle."""
import threading
from unittest.mock import patch

import numpy as np
import pytest

import oracle
from depthestimation_b200.synth import make_calibration, make_pair

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    from depthestimation_b200.engine import StereoEngine
    e = StereoEngine(0)
    yield e
    e.close()


def _pp_dp(W, fast=False):
    from depthestimation_b200.engine import make_depth_params, make_post_params
    return make_post_params(fast, 100, 1.0, 2.5, True), make_depth_params(0.8 * W, 0.54, 0.0, eps=0, max_depth=80.0)


def test_optional_outputs_depth_only_and_int16(engine):
    """disp == NULL skips the disparity download (the video generator's case), depth == NULL the depth download, and the
    int16 plane is exactly disparity x 16."""
    H, W, D, B = 60, 200, 32, 5
    engine.configure(num_disp=D, sgbm_mode="sgbm")
    pairs = [make_pair(H, W, D, seed=3000 + k)[:2] for k in range(B)]
    Ls, Rs = np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    pp, dp = _pp_dp(W)
    disp, depth = engine.pipeline_batch(Ls, Rs, pp, dp)
    none, depth_only = engine.pipeline_batch(Ls, Rs, pp, dp, want_disp=False)
    assert none is None and np.array_equal(depth_only, depth)
    d2, d16, z2 = engine.pipeline_frames(Ls, Rs, pp, dp, want_disp=True, want_disp16=True, want_depth=False)
    assert z2 is None and np.array_equal(d2, disp)
    assert np.array_equal(d16.astype(np.float32) / np.float32(16), disp)
    _, d16b, z3 = engine.pipeline_frames(Ls, Rs, pp, dp, want_disp=False, want_disp16=True, want_depth=True)
    assert np.array_equal(d16b, d16) and np.array_equal(z3, depth)
    # single-pair entry point: disp may be NULL too
    from depthestimation_b200 import _lib
    import ctypes as C
    z = np.empty_like(depth[0])
    rc = engine._lib.b200s_pipeline(engine._h, Ls[0].ctypes.data, Rs[0].ctypes.data, H, W, W, 0, C.byref(pp), C.byref(dp),
                                    None, z.ctypes.data, None, None)
    assert rc == 0 and np.array_equal(z, depth[0])
    rc = engine._lib.b200s_pipeline(engine._h, Ls[0].ctypes.data, Rs[0].ctypes.data, H, W, W, 0, C.byref(pp), C.byref(dp),
                                    None, None, None, None)
    assert rc == _lib.E_INVALID


@pytest.mark.parametrize("case", ["bgr_maps", "bgr_area_maps", "bgr_area_linear_maps", "bgr_nomaps", "gray_area_frac"])
def test_pipeline_frames_input_prep_matches_cv2(engine, case):
    """Frames as captured -> [INTER_AREA] -> [INTER_LINEAR fix-up] -> gray -> [remap] -> pipeline, on the device, against
    the same chain made of cv2 calls in the reference's order (input.py:39-43, rectify.py:92-119,185-186)."""
    cv2 = pytest.importorskip("cv2")
    from oracle import cv2_reference as cr
    H, W, D, B = 72, 208, 32, 3
    rng = np.random.default_rng(5)
    cfg = dict(bgr_maps=(1.0, 3, (W, H), True), bgr_area_maps=(2.0, 3, (W, H), True),
               bgr_area_linear_maps=(2.0, 3, (W + 10, H - 6), True), bgr_nomaps=(1.0, 3, (W, H), False),
               gray_area_frac=(1.6, 1, (W, H), False))[case]
    up, cn, size, use_maps = cfg
    frames_L, frames_R = [], []
    for k in range(B):
        l, r, _ = make_pair(H, W, D, seed=3100 + k)
        sh, sw = int(round(H * up)), int(round(W * up))
        big = [cv2.resize(x, (sw, sh), interpolation=cv2.INTER_CUBIC) for x in (l, r)]
        if cn == 3:   # three different channels, so the channel order matters
            big = [np.stack([b, np.roll(b, 1, 1), 255 - b], -1).astype(np.uint8) for b in big]
        frames_L.append(big[0]); frames_R.append(big[1])
    Ls, Rs = np.ascontiguousarray(np.stack(frames_L)), np.ascontiguousarray(np.stack(frames_R))
    sh, sw = Ls.shape[1:3]
    area = (W, H) if up != 1.0 else None
    calib = make_calibration(size[1], size[0])
    engine.configure(num_disp=D, sgbm_mode="sgbm_3way")
    maps = None
    if use_maps:
        maps = cr.build_maps(calib)
        engine.set_maps(maps["map1_L"], maps["map2_L"], maps["map1_R"], maps["map2_R"])
    pp, dp = _pp_dp(size[0])
    disp, d16, depth = engine.pipeline_frames(Ls, Rs, pp, dp, size=size if use_maps else None, area=area,
                                              use_maps=use_maps, want_disp=True, want_disp16=True, want_depth=True)
    ref = cr.Cv2StereoPipeline(dict(num_disp=D, sgbm_mode="sgbm_3way"), None, 1.0, False, 0.8 * size[0], 0.54, 0.0, 80.0)
    for k in range(B):
        l, r = Ls[k], Rs[k]
        if area is not None:
            l, r = (cv2.resize(x, area, interpolation=cv2.INTER_AREA) for x in (l, r))
        if use_maps and (l.shape[1], l.shape[0]) != size:
            l, r = (cv2.resize(x, size, interpolation=cv2.INTER_LINEAR) for x in (l, r))
        if use_maps:
            l, r = cr.rectify_pair(l, r, maps)
        want_disp, want_depth = ref(l, r)
        assert np.array_equal(disp[k], want_disp), f"frame {k}: {(disp[k] != want_disp).sum()} px differ"
        assert np.array_equal(d16[k], np.rint(want_disp * 16).astype(np.int16))
        fin = np.isfinite(want_depth)
        assert np.array_equal(np.isfinite(depth[k]), fin) and np.allclose(depth[k][fin], want_depth[fin], rtol=1e-4, atol=0)


@pytest.mark.parametrize("mode", ["sgbm", "hh", "sgbm_3way", "hh4"])
def test_bench_sized_batch_equals_single_pairs(engine, mode):
    """The launch configuration a large batch takes (row kernel with full CTAs, 128-row cost segments, many co-resident
    clusters) differs from the single-pair one: every frame of a 24-pair KITTI-shape batch must equal the single-pair
    result, and the first and last must equal cv2."""
    cv2 = pytest.importorskip("cv2")
    import torch
    from oracle import cv2_reference as cr
    H, W, D, B = 375, 1242, 128, 24
    engine.configure(num_disp=D, sgbm_mode=mode)
    distinct = [make_pair(H, W, D, seed=1000 + k)[:2] for k in range(3)]
    singles = [engine.compute(l, r) for l, r in distinct]
    dL = torch.from_numpy(np.stack([distinct[k % 3][0] for k in range(B)])).cuda()
    dR = torch.from_numpy(np.stack([distinct[k % 3][1] for k in range(B)])).cuda()
    out = torch.empty((B, H, W), dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    engine.compute_batch_dev(dL.data_ptr(), dR.data_ptr(), B, H, W, out.data_ptr())
    engine.sync()
    got = out.cpu().numpy()
    for k in range(B):
        assert np.array_equal(got[k], singles[k % 3]), f"frame {k} ({mode}): {(got[k] != singles[k % 3]).sum()} px differ"
    for k in (0, B - 1):
        want = cr.compute_disparity16(*distinct[k % 3], num_disp=D, sgbm_mode=mode)
        assert np.array_equal(got[k], want), f"frame {k} vs cv2 ({mode}): {(got[k] != want).sum()} px differ"


def _run_in_threads(jobs):
    errs, outs = [], [None] * len(jobs)

    def run(i, fn):
        try:
            outs[i] = fn()
        except BaseException as exc:  # noqa: BLE001
            errs.append(exc)
    ts = [threading.Thread(target=run, args=(i, fn)) for i, fn in enumerate(jobs)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]
    return outs


def _thread_job(device, H, W, D, mode, seeds, reps=3):
    from depthestimation_b200 import _lib
    from depthestimation_b200.engine import StereoEngine

    def job():
        eng = StereoEngine(device, num_disp=D, sgbm_mode=mode)
        res = []
        for _ in range(reps):
            for s in seeds:
                l, r, _ = make_pair(H, W, D, seed=s)
                res.append(eng.compute(l, r))
                assert eng.last_flow & _lib.FLOW_ROWS_FUSED
                assert eng.last_flow & _lib.FLOW_VERT_DOWN, "the cluster kernel must run, not the generic fallback"
                assert not eng.last_flow & _lib.FLOW_PATH_VERTICAL
        eng.close()
        return res
    return job


def test_two_handles_two_threads_different_shapes():
    """Two engines in two host threads with different (W1, Dp): the cluster kernel's cached plan is per device and
    guarded, so neither thread may launch with the other's geometry."""
    shapes = [(96, 320, 64, "sgbm", [11, 12]), (80, 420, 128, "sgbm", [13, 14])]
    outs = _run_in_threads([_thread_job(0, *s) for s in shapes])
    for (H, W, D, mode, seeds), res in zip(shapes, outs):
        for i, got in enumerate(res):
            l, r, _ = make_pair(H, W, D, seed=seeds[i % len(seeds)])
            want = oracle.sgbm_compute(l, r, oracle.make_params(num_disp=D, sgbm_mode=mode), stages=False)
            assert np.array_equal(got, want)


def test_two_devices_in_one_process():
    """In-process multi-GPU (what ThreadedSharder / StereoDepthEstimatorVideo(devices=[0, 1]) do): device 1 must run the
    fused cluster kernel too (function attributes are per device) and give the same bits."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    shapes = [(96, 320, 64, "sgbm", [21, 22]), (96, 320, 64, "hh", [21, 22])]
    outs = _run_in_threads([_thread_job(dev, *shapes[dev % 2]) for dev in (0, 1)])
    for (H, W, D, mode, seeds), res in zip(shapes, outs):
        for i, got in enumerate(res):
            l, r, _ = make_pair(H, W, D, seed=seeds[i % len(seeds)])
            want = oracle.sgbm_compute(l, r, oracle.make_params(num_disp=D, sgbm_mode=mode), stages=False)
            assert np.array_equal(got, want)


class _FakeCapture:
    frames = []

    def __init__(self, *a, downscale_factor=1.0, **k):
        self.i, self.f = 0, downscale_factor

    def start(self):
        pass

    def read(self):
        if self.i >= len(self.frames):
            return None
        self.i += 1
        l, r = self.frames[self.i - 1]
        if self.f != 1.0:   # what threaded_stereo.py:41-46 does in the capture thread
            import cv2
            l, r = (cv2.resize(x, (int(x.shape[1] * self.f), int(x.shape[0] * self.f)), interpolation=cv2.INTER_AREA)
                    for x in (l, r))
        return l, r

    def stop(self):
        pass


@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 1]])
def test_video_generator_bgr_downscale_calibrated(devices):
    """BGR frames at capture resolution through StereoDepthEstimatorVideo(devices=...) with downscale 0.5 and a
    calibration: the sharded path (INTER_AREA, gray, remap, matcher, post, depth on each worker's device, depth plane
    only back) yields, in order, exactly what the frame-by-frame StereoCore path yields."""
    cv2 = pytest.importorskip("cv2")
    import torch
    if max(devices) >= torch.cuda.device_count():
        pytest.skip("needs two GPUs")
    from depthestimation_b200.depthlib import StereoCore, StereoDepthEstimatorVideo
    H, W, D = 64, 192, 64          # after the 0.5 downscale; num_disp is scaled by 0.5 twice (reference quirk): 64 -> 32 -> 16
    frames = []
    for k in range(9):
        l, r, _ = make_pair(H, W, 16, seed=3300 + k)
        big = [cv2.resize(x, (2 * W, 2 * H), interpolation=cv2.INTER_LINEAR) for x in (l, r)]
        frames.append(tuple(np.stack([b, 255 - b, np.roll(b, 2, 0)], -1).astype(np.uint8) for b in big))
    _FakeCapture.frames = frames
    calib = make_calibration(H, W)
    cfg = dict(calib)
    cfg.update(num_disp=D, sgbm_mode="sgbm", focal_length=0.8 * W * 2, max_depth=60.0)
    with patch("depthestimation_b200.depthlib.StereoDepthEstimatorVideo.ThreadedStereoCapture", _FakeCapture):
        ref = StereoDepthEstimatorVideo("l.mp4", "r.mp4", downscale_factor=0.5, target_fps=0)      # frame by frame
        ref.configure_sgbm(**cfg)
        expect = list(ref.estimate_depth())
        vid = StereoDepthEstimatorVideo("l.mp4", "r.mp4", downscale_factor=0.5, target_fps=0, devices=devices, chunk=4)
        vid.configure_sgbm(**cfg)
        got = list(vid.estimate_depth())
        auto = StereoDepthEstimatorVideo("l.mp4", "r.mp4", downscale_factor=0.5, target_fps=0, devices=devices)
        auto.configure_sgbm(**cfg)
        got_auto = list(auto.estimate_depth())
    assert len(expect) == len(frames) == len(got) == len(got_auto)
    assert ref.core.sgbm_params["num_disp"] == 16
    for a, b, c in zip(got, expect, got_auto):
        assert a.shape == b.shape and np.array_equal(a, b) and np.array_equal(c, b)
    # and the frame-by-frame path itself is the reference chain in cv2 calls
    from oracle import cv2_reference as cr
    maps = cr.build_maps(calib)
    l, r = (cv2.resize(x, (W, H), interpolation=cv2.INTER_AREA) for x in frames[0])
    l, r = cr.rectify_pair(l, r, maps)
    pipe = cr.Cv2StereoPipeline(dict(num_disp=16, sgbm_mode="sgbm"), None, 0.5, False, 0.8 * W * 2 * 0.25, 0.54, 0.0, 60.0)
    _, want = pipe(l, r)
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(expect[0]), fin) and np.allclose(expect[0][fin], want[fin], rtol=1e-4, atol=0)


def test_video_generator_hole_filling_not_sharded():
    """hole_filling=True needs the host Telea step between the GPU stages: the generator must take the frame-by-frame
    path (same output as StereoCore.estimate_depth), not silently drop the stage."""
    pytest.importorskip("cv2")
    from depthestimation_b200.depthlib import StereoCore, StereoDepthEstimatorVideo
    H, W, D = 48, 160, 32
    frames = [tuple(np.stack([x] * 3, -1) for x in make_pair(H, W, D, seed=3400 + k)[:2]) for k in range(3)]
    _FakeCapture.frames = frames
    cfg = dict(num_disp=D, sgbm_mode="sgbm", focal_length=0.8 * W, baseline=0.54, hole_filling=True)
    core = StereoCore()
    core.configure_sgbm(**cfg)
    expect = [core.estimate_depth(l, r)[1] for l, r in frames]
    with patch("depthestimation_b200.depthlib.StereoDepthEstimatorVideo.ThreadedStereoCapture", _FakeCapture):
        vid = StereoDepthEstimatorVideo("l.mp4", "r.mp4", target_fps=0, devices=[0, 0], chunk=2)
        vid.configure_sgbm(**cfg)
        got = list(vid.estimate_depth())
    assert len(got) == 3
    for a, b in zip(got, expect):
        assert np.array_equal(a, b)
    core2 = StereoCore()
    core2.configure_sgbm(**dict(cfg, hole_filling=False))
    assert not np.array_equal(core2.estimate_depth(*frames[0])[1], expect[0]), "the case must exercise hole filling"


def test_no_silent_host_fallbacks():
    from depthestimation_b200.depthlib.postprocess import fill_holes
    from depthestimation_b200.depthlib.rectify import _ensure_image_size
    d = np.zeros((8, 8), np.float32)
    with pytest.raises(ValueError, match="kernel_size"):
        fill_holes(d, method="nearest", kernel_size=33)
    with pytest.raises(ValueError, match="float32"):
        fill_holes(d.astype(np.float64), method="nearest", kernel_size=3)
    with pytest.warns(RuntimeWarning):
        with pytest.raises(ValueError, match="uint8"):
            _ensure_image_size(np.zeros((8, 8), np.float32), (16, 16))


def test_build_maps_matches_oracle_and_cv2(engine):
    """b200s_build_maps (cv2.initUndistortRectifyMap on the GPU, double precision, no contraction): bit-exact against the
    numpy restatement for every calibration model, and against cv2 itself up to the documented one-ulp flips of its
    SIMD column accumulation (at most one value per map pair; none on the synthetic and the rotated / translated rig)."""
    cv2 = pytest.importorskip("cv2")
    from oracle import maps as om
    cases = om.calibration_cases()
    cal = make_calibration(375, 1242)
    cases.append(("kitti", cal["cam_matrix_L"], cal["dist_coeff_L"], cal["cam_matrix_R"], cal["dist_coeff_R"], (1242, 375),
                  np.eye(3), np.array([-0.54, 0.0, 0.0])))
    for name, KL, dL, KR, dR, size, R, T in cases:
        for alpha in (0.0, 1.0):
            rL, rR, pL, pR, *_ = cv2.stereoRectify(KL, dL, KR, dR, size, R, T, flags=cv2.CALIB_ZERO_DISPARITY, alpha=alpha)
            for K, d, r, p in ((KL, dL, rL, pL), (KR, dR, rR, pR)):
                got = engine.build_maps(K, d, r, p, size)
                ref = om.init_undistort_rectify_map(K, d, r, p, size)
                want = cv2.initUndistortRectifyMap(K, d, r, p, size, cv2.CV_32FC1)
                assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]), name
                bad = sum(int((g != w).sum()) for g, w in zip(got, want))
                assert bad <= (0 if name in ("synthetic", "rotated", "kitti") else 1), (name, alpha, bad)
    # P = None (new camera matrix = K), R = None, no distortion: the identity map
    K = cases[0][1]
    m1, m2 = engine.build_maps(K, None, None, None, (320, 200))
    w1, w2 = cv2.initUndistortRectifyMap(K, None, None, K, (320, 200), cv2.CV_32FC1)
    assert np.array_equal(m1, w1) and np.array_equal(m2, w2)


def test_rectify_images_uncached_rebuilds_maps_on_gpu(engine):
    """rectify_images without a cache (depthlib/rectify.py:209-235) rebuilds the maps on every call: stereoRectify on the
    host, initUndistortRectifyMap and both remaps on the GPU; same rectified images as the cv2 chain."""
    cv2 = pytest.importorskip("cv2")
    from oracle import cv2_reference as cr
    from depthestimation_b200.depthlib.rectify import rectify_images
    H, W = 120, 300
    L, R, _ = make_pair(H, W, 32, seed=77)
    cal = make_calibration(H, W)
    gl, gr = rectify_images(L, R, cal["cam_matrix_L"], cal["cam_matrix_R"], cal["baseline"], W, H, cal["dist_coeff_L"],
                            cal["dist_coeff_R"], alpha=1.0, engine=engine)
    wl, wr = cr.rectify_pair(L, R, cr.build_maps(cal, alpha=1.0))
    assert np.array_equal(gl, wl) and np.array_equal(gr, wr)


def test_video_generator_in_memory_sources():
    """Two lists of frames as sources (no capture object): same depth maps, in order, as the frame-by-frame core; the
    chunk size is derived from the known length (at least two chunks per GPU)."""
    from depthestimation_b200.depthlib import StereoCore, StereoDepthEstimatorVideo
    H, W, D = 48, 176, 32
    pairs = [make_pair(H, W, D, seed=3500 + k)[:2] for k in range(11)]
    lefts = [np.stack([l] * 3, -1) for l, _ in pairs]
    rights = [np.stack([r] * 3, -1) for _, r in pairs]
    cfg = dict(num_disp=D, sgbm_mode="hh4", focal_length=0.8 * W, baseline=0.54)
    core = StereoCore()
    core.configure_sgbm(**cfg)
    expect = [core.estimate_depth(l, r)[1] for l, r in zip(lefts, rights)]
    vid = StereoDepthEstimatorVideo(lefts, rights, target_fps=0, devices=[0])
    vid.configure_sgbm(**cfg)
    for _ in range(2):                       # second pass reuses the cached worker (engine, page-locked buffers)
        got = list(vid.estimate_depth())
        assert len(got) == len(expect)
        for a, b in zip(got, expect):
            assert np.array_equal(a, b)
    assert len(vid._worker_pool[0]) == 1
    # the chunk length is fixed per configuration: the engine's preferred batch follows the free device memory (which the
    # worker's own scratch changes), and a different answer on a later call must not throw the worker away
    w0, chunk0 = vid._worker_pool[0][0], vid._worker_cap
    with patch.object(type(vid.core.sgbm.engine), "preferred_batch", lambda self, H, W: 1):
        assert len(list(vid.estimate_depth())) == len(expect)
    assert vid._worker_pool[0][0] is w0 and vid._worker_cap == chunk0
    vid.left_source, vid.right_source = lefts[:2], rights[:2]      # nor does a source of another length
    assert len(list(vid.estimate_depth())) == 2 and vid._worker_pool[0][0] is w0
    vid.left_source, vid.right_source = lefts, rights
    vid.configure_sgbm(num_disp=D // 2)     # a new configuration does: the old worker's engine and buffers are given back
    assert len(list(vid.estimate_depth())) == len(expect)
    assert vid._worker_pool[0][0] is not w0 and not w0.eng._h
    vid.close_workers()
    assert vid._worker_pool == {}


def test_pipeline_frames_rejects_bad_requests(engine):
    """Argument errors of the frame pipeline come back as B200S_E_INVALID -> ValueError with the reason, before any work."""
    from depthestimation_b200.engine import make_depth_params, make_post_params
    engine.configure(num_disp=32, sgbm_mode="sgbm")
    pp, dp = make_post_params(False, 100, 1.0, 2.5, True), make_depth_params(100.0, 0.5)
    L = np.zeros((2, 40, 96), np.uint8)
    with pytest.raises(ValueError, match="only shrinks"):
        engine.pipeline_frames(L, L, pp, dp, area=(200, 80))                 # INTER_AREA upscaling
    with pytest.raises(ValueError, match="no output requested"):
        engine.pipeline_frames(L, L, pp, dp, want_disp=False, want_disp16=False, want_depth=False)
    with pytest.raises(ValueError, match="no output requested"):
        engine.pipeline_frames(L, L, pp, make_depth_params(None, None), want_disp=False, want_depth=True)   # depth disabled
    with pytest.raises(ValueError, match="num_disp"):
        engine.pipeline_frames(np.zeros((1, 40, 24), np.uint8), np.zeros((1, 40, 24), np.uint8), pp, dp)
    with pytest.raises(ValueError):
        engine.pipeline_frames(np.zeros((1, 40, 96, 2), np.uint8), np.zeros((1, 40, 96, 2), np.uint8), pp, dp)   # 2 channels
    with pytest.raises(ValueError, match="does not match the rectification maps"):
        engine.set_maps(*[np.zeros((30, 64), np.float32)] * 4)
        engine.pipeline_frames(L, L, pp, dp, use_maps=True)
    # and the handle still works afterwards
    d, _, z = engine.pipeline_frames(L, L, pp, dp)
    assert d.shape == (2, 40, 64) and z.shape == d.shape
