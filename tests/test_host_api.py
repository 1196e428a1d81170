"""CPU-only tests of the host side: the depthlib-compatible API surface (mirrors what the reference's
tests/test_config.py, test_inputs.py, test_rectification_cache.py pin), the C-ABI export list, and the
frame-sharding logic incl. a world_size-2 gloo gather. No compute call is made (no GPU here)."""
import ctypes
import re
from pathlib import Path
from unittest.mock import patch

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_abi_library_exports_every_declared_symbol():
    hdr = (ROOT / "include" / "b200stereo.h").read_text()
    declared = set(re.findall(r"\b(b200s_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    from depthestimation_b200 import _lib
    lib = _lib.load()
    assert lib.b200s_abi_version() == 1
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)


def test_abi_struct_layouts_match_header():
    from depthestimation_b200 import _lib
    assert ctypes.sizeof(_lib.B200sParams) == 11 * 4
    assert ctypes.sizeof(_lib.B200sPostParams) == 5 * 4
    assert ctypes.sizeof(_lib.B200sDepthParams) == 6 * 4


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from depthestimation_b200.engine import StereoEngine, EngineError
    with pytest.raises(EngineError):
        StereoEngine(0)


def test_configure_sgbm_scaling_and_validation():
    from depthestimation_b200.depthlib import StereoCore
    core = StereoCore(downscale_factor=0.5)
    assert core.sgbm_params["num_disp"] == 128 and core.sgbm_params["sgbm_mode"] == "sgbm_3way"
    assert len(core.sgbm_params) == 22
    core.configure_sgbm(num_disp=64, block_size=7, focal_length=1000.0, doffs=10.0)
    p = core.get_sgbm_params()
    assert p["num_disp"] == 32 and p["block_size"] == 7 and p["focal_length"] == 500.0 and p["doffs"] == 5.0
    p["num_disp"] = 1
    assert core.sgbm_params["num_disp"] == 32           # get_sgbm_params returns a copy
    core.configure_sgbm(**core.get_sgbm_params())       # the video façade's re-apply: scales again (quirk C.3)
    assert core.sgbm_params["num_disp"] == 16
    with pytest.raises(ValueError, match=r"Invalid parameter 'bogus'. Valid parameters: \["):
        core.configure_sgbm(bogus=1)
    assert core.sgbm.kw["num_disp"] == 16 and core.sgbm.kw["block_size"] == 7


def test_estimator_facades_validate():
    from depthestimation_b200.depthlib import StereoDepthEstimator, StereoDepthEstimatorVideo
    for bad in (0, -1, 1.5):
        with pytest.raises(ValueError, match="downscale_factor must be between 0 and 1."):
            StereoDepthEstimator(downscale_factor=bad)
    est = StereoDepthEstimator()
    with pytest.raises(ValueError, match="Left and right sources must be provided"):
        est.estimate_depth()
    with pytest.raises(ValueError, match="Disparity map not computed"):
        est.visualize_results()
    vid = StereoDepthEstimatorVideo()
    with pytest.raises(ValueError, match="Both left_source and right_source must be provided"):
        next(vid.estimate_depth())
    from depthestimation_b200.depthlib import StereoCore
    with pytest.raises(ValueError, match="Left and right sources must be set before estimating depth."):
        StereoCore().estimate_depth(None, np.zeros((4, 4), np.uint8))


def test_load_stereo_pair_invalid_paths():
    from depthestimation_b200.depthlib.input import load_stereo_pair
    with patch("cv2.imread", return_value=None):
        with pytest.raises(FileNotFoundError, match="One or both image paths are invalid"):
            load_stereo_pair("a.png", "b.png")


class _HostMapEngine:
    """Stands in for the GPU map kernel in the CPU suite (the cache logic is what is tested here): the oracle's
    restatement of initUndistortRectifyMap."""
    def build_maps(self, K, dist, R, P, size):
        from oracle import maps as om
        return om.init_undistort_rectify_map(K, dist, R, P, size)


def test_rectification_cache_identity():
    pytest.importorskip("cv2")
    from depthestimation_b200.depthlib.rectify import RectificationCache
    K = np.array([[500.0, 0, 320], [0, 500.0, 240], [0, 0, 1]])
    cache = RectificationCache(engine=_HostMapEngine())
    m1 = cache.get_maps(K, K, 0.1, 640, 480)
    m2 = cache.get_maps(K, K, 0.1, 640, 480)
    assert m1 is m2 and set(m1) == {"map1_L", "map2_L", "map1_R", "map2_R"}
    assert m1["map1_L"].shape == (480, 640) and m1["map1_L"].dtype == np.float32
    m3 = cache.get_maps(K, K, 0.2, 640, 480)
    assert m3 is not m1
    cache.clear()
    assert cache._maps is None


def test_threaded_capture_drop_and_timeout():
    from depthestimation_b200.depthlib.threaded_stereo import ThreadedStereoCapture

    class FakeCap:
        def __init__(self, n): self.n, self.i = n, 0
        def isOpened(self): return True
        def read(self):
            if self.i >= self.n: return False, None
            self.i += 1
            return True, np.full((4, 6, 3), self.i, np.uint8)
        def release(self): pass

    with patch("depthestimation_b200.depthlib.threaded_stereo.open_capture", side_effect=lambda s: FakeCap(5)):
        cap = ThreadedStereoCapture("l", "r", drop_frames=False)
        cap.start()
        got = []
        while True:
            pair = cap.read()
            if pair is None:
                break
            got.append(int(pair[0][0, 0, 0]))
        cap.stop()
    assert got == [1, 2, 3, 4, 5]       # video-file mode never skips a frame


# ---- frame sharding -------------------------------------------------------------------------------
def test_frames_for_rank_partition():
    from depthestimation_b200.sharding import frames_for_rank
    for n in (0, 1, 7, 240):
        for world in (1, 2, 4, 8):
            parts = [frames_for_rank(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        frames_for_rank(4, 2, 2)


def test_threaded_sharder_order_and_failover():
    from depthestimation_b200.sharding import ThreadedSharder
    frames = [(np.full((3, 4), i, np.uint8), np.full((3, 4), 100 + i, np.uint8)) for i in range(23)]
    used = []

    def make_worker(dev):
        def fn(L, R):
            if dev == 1 and len(used) > 2:
                raise RuntimeError("injected GPU failure")
            used.append(dev)
            return L.astype(np.float32) * 2, R.astype(np.float32)
        return fn

    out = list(ThreadedSharder([0, 1, 2], make_worker, chunk=4).run(iter(frames)))
    assert len(out) == 23
    for i, (disp, depth) in enumerate(out):
        assert disp.shape == (3, 4) and float(disp[0, 0]) == 2 * i and float(depth[0, 0]) == 100 + i

    def all_fail(dev):
        def fn(L, R):
            raise RuntimeError("boom")
        return fn
    with pytest.raises(RuntimeError, match="all GPU workers failed"):
        list(ThreadedSharder([0, 1], all_fail, chunk=2).run(iter(frames)))


def test_threaded_sharder_async_workers_ramp_and_failover():
    """The submit / wait worker protocol the video generator uses (several chunks in flight per GPU, chunks handed over as
    lists of frames, ramped-up first chunks, depth-only results): frame order, chunk sizes, in-flight bound, and a worker
    that dies with chunks in flight hands all of them to the survivors."""
    import threading
    import time as _time
    from depthestimation_b200.sharding import ThreadedSharder
    frames = [(np.full((2, 3), i, np.uint8), np.full((2, 3), 200 - i, np.uint8)) for i in range(61)]
    lock = threading.Lock()
    sizes, peak, per_dev = [], {}, {}

    class Worker:
        max_inflight = 3

        def __init__(self, dev, fail_after=None):
            self.dev, self.fly, self.n, self.fail_after = dev, 0, 0, fail_after

        def submit(self, Ls, Rs):
            assert isinstance(Ls, list) and len(Ls) == len(Rs)
            with lock:
                sizes.append(len(Ls))
                per_dev.setdefault(self.dev, []).append(len(Ls))
                self.fly += 1
                self.n += 1
                peak[self.dev] = max(peak.get(self.dev, 0), self.fly)
            if self.fail_after is not None and self.n > self.fail_after:
                raise RuntimeError("injected failure with chunks in flight")
            return np.stack(Rs).astype(np.float32)

        def wait(self, token):
            _time.sleep(0.002)
            with lock:
                self.fly -= 1
            return None, token            # depth only

    out = list(ThreadedSharder([0, 1], lambda d: Worker(d), chunk=8, ramp=True).run(iter(frames)))
    assert [float(depth[0, 0]) for _, depth in out] == [200.0 - i for i in range(61)] and all(d is None for d, _ in out)
    assert sorted(sizes)[:5] == [1, 2, 2, 4, 4] and sum(sizes) == 61 and max(sizes) == 8      # quarter, half, then full chunks
    # chunk i goes to worker i mod G: each of the two sees a quarter, a half, then full chunks, and the same number of chunks
    assert per_dev[0][:3] == [2, 4, 8] and per_dev[1][:3] == [2, 4, 8] and abs(len(per_dev[0]) - len(per_dev[1])) <= 1
    assert max(peak.values()) <= 3
    sizes.clear()
    out = list(ThreadedSharder([0, 1, 2], lambda d: Worker(d, fail_after=2 if d == 1 else None), chunk=4).run(iter(frames)))
    assert [float(depth[0, 0]) for _, depth in out] == [200.0 - i for i in range(61)]
    with pytest.raises(RuntimeError, match="all GPU workers failed"):
        list(ThreadedSharder([0], lambda d: Worker(d, fail_after=0), chunk=4).run(iter(frames)))


def _gloo_worker(rank, world, port, n, q):
    import os
    import torch.distributed as dist
    from depthestimation_b200.sharding import frames_for_rank, gather_in_order
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = {i: np.full((2, 3), i, np.float32) * (1 + 0) for i in frames_for_rank(n, rank, world)}
    res = gather_in_order(mine, n)
    if rank == 0:
        q.put([float(a[0, 0]) for a in res])
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_gather_in_frame_order():
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n = 11
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [float(i) for i in range(n)]


def test_bench_reference_arm_runs_without_a_gpu():
    """`bench.py --impl reference` (the driver's reference arm) is the reference's own CPU path: it must not touch the
    product library (it runs here, where no GPU exists) and its one JSON line carries the contract's keys."""
    import json
    import subprocess
    import sys
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["metric"] == "stereo_pairs_per_second" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["gpu_launches"] == 0 and line["config"]["workload"].startswith("kitti_1242x375_D128")


def test_threaded_sharder_consumer_stops_early():
    """A consumer that abandons the generator: the workers finish what they have in flight, queued chunks are dropped, the
    threads end."""
    import threading
    from depthestimation_b200.sharding import ThreadedSharder
    frames = [(np.full((2, 3), i % 200, np.uint8), np.full((2, 3), i % 200, np.uint8)) for i in range(400)]
    calls = []

    def make_worker(dev):
        def fn(L, R):
            calls.append(L.shape[0])
            return L.astype(np.float32), R.astype(np.float32)
        return fn
    before = threading.active_count()
    gen = ThreadedSharder([0, 1], make_worker, chunk=4, max_inflight=4).run(iter(frames))
    first = [next(gen) for _ in range(3)]
    gen.close()
    assert [float(d[0, 0]) for d, _ in first] == [0.0, 1.0, 2.0]
    assert sum(calls) <= 4 * 4 * 2 + 8 and threading.active_count() <= before      # nothing beyond the queued window ran
