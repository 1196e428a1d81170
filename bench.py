#!/usr/bin/env python
"""bench.py — throughput of depthlib's stereo hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl b200|reference]

Workload (BASELINE.json configs[1]): KITTI-shape synthetic stereo pairs 1242x375, 128 disparities,
5-path `sgbm`, and per pair the whole StereoCore.estimate_depth path (depthlib/stereo_core.py:274-293):
rectify (remap through cached maps) -> match -> crop -> postprocess -> depth.
A "step" = one batch of B pairs through that path. `value` = pairs/s with inputs resident in HBM;
`e2e` = the same through the host-buffer C-ABI call (pinned host in, pinned host out, copies timed).
One process per GPU under torchrun; pairs shard frame-per-GPU, no collective on the data path.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

H, W, D = 375, 1242, 128
MODE = "sgbm"
SGBM_KW = dict(min_disp=0, num_disp=D, block_size=5, disp12_max_diff=1, prefilter_cap=31, uniqueness_ratio=10,
               speckle_window_size=50, speckle_range=2, sgbm_mode=MODE)
WORKLOAD = f"kitti_{W}x{H}_D{D}_{MODE}_5path_rectify+match+post+depth"
METRIC = "stereo_pairs_per_second"
UNIT = "frames/s"
N_DISTINCT = 4  # distinct synthetic pairs, tiled to the batch


def mde(fps: float) -> float:
    return fps * H * W * D / 1e6


def make_inputs(batch: int):
    from depthestimation_b200.synth import make_pair
    pairs = [make_pair(H, W, D, seed=1000 + k)[:2] for k in range(min(N_DISTINCT, batch))]
    L = np.stack([pairs[i % len(pairs)][0] for i in range(batch)])
    R = np.stack([pairs[i % len(pairs)][1] for i in range(batch)])
    return L, R


def make_maps():
    """Rectification maps for the synthetic calibration, built by the PRODUCT package exactly as StereoCore does per
    calibration (depthestimation_b200.depthlib.rectify.RectificationCache, alpha = 1.0 as stereo_core.py:150)."""
    from depthestimation_b200.depthlib.rectify import RectificationCache
    from depthestimation_b200.synth import make_calibration
    cal = make_calibration(H, W)
    maps = RectificationCache().get_maps(cal["cam_matrix_L"], cal["cam_matrix_R"], cal["baseline"], W, H,
                                         cal["dist_coeff_L"], cal["dist_coeff_R"], None, None, 1.0)
    return maps, cal


def make_reference_maps():
    """The reference arm's maps: the reference's own cv2 calls (rectify.py:63-73), else the oracle's restatement of them.
    Nothing of the product library is on this arm's path, so it also runs on a box without a GPU."""
    from depthestimation_b200.synth import make_calibration
    from oracle import cv2_reference as cr
    cal = make_calibration(H, W)
    if cr.HAVE_CV2:
        return cr.build_maps(cal, alpha=1.0), cal
    # no cv2 (and so no stereoRectify) on this host: identity maps time the same remap work
    xs, ys = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32))
    return dict(map1_L=xs, map2_L=ys, map1_R=xs.copy(), map2_R=ys.copy()), cal


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (profiling guide recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 8:
                continue
            try:
                sm.append(float(p[1])); mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's own per-frame path (cv2 driven exactly as depthlib does), frame-parallel
# over all host cores; falls back to the C oracle port only when cv2 is not importable.
# ---------------------------------------------------------------------------------------------
class CpuReference:
    def __init__(self, maps, cal):
        from oracle import cv2_reference as cr
        self.cores = os.cpu_count() or 1
        self.have_cv2 = cr.HAVE_CV2
        self.maps, self.cal = maps, cal
        if self.have_cv2:
            import cv2
            cv2.setNumThreads(1)  # sgbm is single-threaded inside OpenCV; parallelism is across frames
            self.kind = "reference"
            self.desc = f"cv2 {cr.CV2_VERSION} driven as depthlib/stereo_core.py:125-293 does"
            self.pipes = []
            for _ in range(self.cores):
                p = cr.Cv2StereoPipeline(SGBM_KW, calib=None, downscale=1.0, fast_mode=False,
                                         focal_length=cal["focal_length"], baseline=cal["baseline"])
                p.maps = maps
                self.pipes.append(p)
        else:
            import oracle
            self.kind = "port"
            self.desc = "C oracle port (cv2 not importable)"
            self.oracle = oracle
            self.params = oracle.make_params(**SGBM_KW)

    def _one(self, wid, L, R):
        if self.have_cv2:
            return self.pipes[wid](L, R)
        o = self.oracle
        l = o.remap_u8(L, self.maps["map1_L"], self.maps["map2_L"])
        r = o.remap_u8(R, self.maps["map1_R"], self.maps["map2_R"])
        d = (o.sgbm_compute(l, r, self.params).astype(np.float32) / 16.0)[:, D:]
        d = o.postprocess(d, 100, 1.0, 2.5)
        return d, o.disp_to_depth(d, self.cal["focal_length"], self.cal["baseline"], 0.0, eps=0)

    def as_shipped(self, L, R, nframes: int) -> dict:
        """BASELINE.md section 3's other CPU figure: ONE stream, cv2's default thread pool (what a user of the reference
        gets from StereoDepthEstimator.estimate_depth in a loop), with the per-stage split of SURVEY section 6."""
        if not self.have_cv2:
            return {}
        import cv2
        cv2.setNumThreads(-1)
        timers: dict = {}
        self.pipes[0](L[0], R[0])
        t0 = time.perf_counter()
        for i in range(nframes):
            self.pipes[0](L[i % len(L)], R[i % len(R)], timers)
        t = time.perf_counter() - t0
        cv2.setNumThreads(1)
        return {"as_shipped_fps": nframes / t, "as_shipped_sample": f"{nframes} frames, one stream, cv2.setNumThreads(-1)",
                "stage_ms_per_frame": {k: 1e3 * v / nframes for k, v in timers.items()}}

    def run_frames(self, L, R, nframes: int) -> float:
        """Process nframes (cycled from the batch) on all cores; returns seconds."""
        from concurrent.futures import ThreadPoolExecutor
        idx = list(range(nframes))
        t0 = time.perf_counter()
        with ThreadPoolExecutor(self.cores) as ex:
            def work(w):
                for i in idx[w::self.cores]:
                    self._one(w, L[i % len(L)], R[i % len(R)])
            list(ex.map(work, range(self.cores)))
        return time.perf_counter() - t0


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    L, R = make_inputs(N_DISTINCT)
    maps, cal = make_reference_maps()
    ref = CpuReference(maps, cal)
    per_step = ref.cores  # one frame per host thread and step: a bounded sample of the workload
    for _ in range(min(args.warmup, 1)):
        ref.run_frames(L, R, per_step)
    t = 0.0
    for _ in range(args.steps):
        t += ref.run_frames(L, R, per_step)
    fps = args.steps * per_step / t
    line = {
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16/u8 (+f32 post)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": per_step, "distinct_pairs": N_DISTINCT},
        "mde_per_s": mde(fps),
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": ref.cores, "kind": ref.kind,
                         "sample": f"{args.steps} steps x {per_step} frames, frame-parallel, {ref.desc}"},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def algorithmic_bytes(eng, batch):
    """u16 C and S volumes, V = H*W1*D cells per frame; canonical dataflow 10 B/cell:
    k_cost writes C (2V); k_vert (the three top-down directions fused in one launch) reads C and writes S (4V);
    k_rows (both horizontal directions + winner selection + LR check in one launch) reads C and S (4V).
    Each is ONE launch per step over the whole batch."""
    g = eng.geometry(W)
    V = H * g["W1"] * D
    per_kernel = {"cost": ("k_cost<5,64,1> (prefiltered rows -> C)", 2.0 * V * batch),
                  "paths": ("k_vert<8,8,0,512,0> (3 vertical-family directions, one cluster per frame)", 4.0 * V * batch),
                  "wta": ("k_rows<8,8,0,0,1> (2 horizontal directions + per-chunk winner selection + LR check)", 4.0 * V * batch)}
    return V, per_kernel, 10.0 * V * batch


def ncu_traffic(stage, batch):
    """DRAM bytes per launch of the stage's kernel from the committed `ncu --set full` capture (profiles/), scaled
    from the captured batch to this run's batch (traffic is proportional to the number of frames); None if absent."""
    try:
        t = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text())[stage]
        return float(t["dram_bytes_per_launch"]) * batch / float(t["batch"])
    except Exception:
        return None


def parity_check(L, R, maps, cal, d_disp, d_depth, h_disp16, frames):
    """Frames `frames` of the timed batch: device results against the reference chain (checker = oracle/, as in tests)."""
    from oracle import cv2_reference as cr
    disp, depth = d_disp.cpu().numpy(), d_depth.cpu().numpy()
    mism = 0
    if cr.HAVE_CV2:
        pipe = cr.Cv2StereoPipeline(SGBM_KW, calib=None, downscale=1.0, fast_mode=False,
                                    focal_length=cal["focal_length"], baseline=cal["baseline"])
        pipe.maps = maps
        checker = f"cv2 {cr.CV2_VERSION} as depthlib drives it"
        ref = lambda l, r: pipe(l, r)
    else:
        import oracle
        prm = oracle.make_params(**SGBM_KW)
        checker = "C oracle"

        def ref(l, r):
            l = oracle.remap_u8(l, maps["map1_L"], maps["map2_L"]); r = oracle.remap_u8(r, maps["map1_R"], maps["map2_R"])
            d = oracle.postprocess((oracle.sgbm_compute(l, r, prm).astype(np.float32) / 16.0)[:, D:], 100, 1.0, 2.5)
            return d, oracle.disp_to_depth(d, cal["focal_length"], cal["baseline"], 0.0, eps=0)
    for k in frames:
        wd, wz = ref(L[k], R[k])
        fin = np.isfinite(wz)
        ok = (np.array_equal(disp[k], wd) and np.array_equal(h_disp16[k], np.rint(wd * 16).astype(np.int16)) and
              np.array_equal(np.isfinite(depth[k]), fin) and np.allclose(depth[k][fin], wz[fin], rtol=1e-4, atol=0))
        mism += 0 if ok else 1
    return {"frames": len(frames), "mismatch": mism, "checker": checker,
            "what": "float32 disparity and int16 x16 disparity bit-exact, depth within 1e-4 relative"}


VH, VW, VD, VFRAMES = 720, 1280, 128, 240


def video_api_leg(local, rank, world, barrier, dist, torch):
    """BASELINE config #4 through the reference-shaped API: VFRAMES synthetic 1280x720 3-channel frames per rank (frame i
    of the global stream goes to rank i mod world: weak scaling, like the rest of this line) through
    StereoDepthEstimatorVideo(devices=[local]); the generator yields one depth map per frame, in order."""
    from depthestimation_b200.depthlib import StereoDepthEstimatorVideo
    from depthestimation_b200.synth import make_calibration, make_pair
    distinct = []
    for k in range(3):
        l, r, _ = make_pair(VH, VW, VD, seed=1100 + k)
        distinct.append((np.ascontiguousarray(np.stack([l] * 3, -1)), np.ascontiguousarray(np.stack([r] * 3, -1))))
    lefts = [distinct[i % 3][0] for i in range(VFRAMES)]
    rights = [distinct[i % 3][1] for i in range(VFRAMES)]
    cal = make_calibration(VH, VW)
    cfg = dict(cal)
    cfg.update(num_disp=VD, max_depth=100.0)

    vid = StereoDepthEstimatorVideo(lefts, rights, target_fps=0, devices=[local])
    vid.configure_sgbm(**cfg)

    def run():
        n, sink = 0, 0.0
        for depth_m in vid.estimate_depth():
            n += 1
            sink += float(depth_m[0, 0])
        return n

    run()                                         # warm-up: allocations, maps, page-locked buffers
    run()                                         # (a fresh box's first pass also pays for page faults and clock ramps)
    reps = []
    for _ in range(3):                            # the job lasts ~0.12 s: median of three, each the max over ranks
        barrier()
        t0 = time.perf_counter()
        n = run()
        te = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        assert n == VFRAMES
        reps.append(float(te.item()))
    te = torch.tensor([sorted(reps)[1]], dtype=torch.float64)
    # the 240-frame job of config #4 lasts ~0.1 s (three chunks): filling the first chunk and draining the last are
    # a fifth of it. The same generator on a stream four times as long, for the steady-state rate:
    vid.left_source, vid.right_source = lefts * 4, rights * 4
    barrier()
    t0 = time.perf_counter()
    n4 = run()
    t4 = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t4, op=dist.ReduceOp.MAX)
    assert n4 == 4 * VFRAMES
    # the same frames through the asynchronous C ABI directly (page-locked chunk buffers already filled, depth plane
    # only, three chunks in flight): what the Python generator above is measured against
    from depthestimation_b200.engine import pinned_empty
    w = vid._worker_pool[local][0]
    eng = w.eng
    Bc = min(eng.preferred_batch(VH, VW), VFRAMES)
    bufs = []
    for _ in range(3):
        bL, bR = pinned_empty((Bc, VH, VW, 3), np.uint8), pinned_empty((Bc, VH, VW, 3), np.uint8)
        for k in range(Bc):
            bL[k] = lefts[k]; bR[k] = rights[k]
        bufs.append((bL, bR, pinned_empty((Bc, VH, VW - VD), np.float32)))

    def cabi(nchunks):
        fly = []
        for i in range(nchunks):
            if len(fly) == 3:
                eng.pipeline_batch_wait(fly.pop(0))
            bL, bR, out = bufs[i % 3]
            fly.append(eng.pipeline_frames_submit(bL, bR, w.pp, w.dp, size=w.size, area=w.area, use_maps=w.use_maps,
                                                  want_disp=False, want_depth=True, out_depth=out))
        for tk in fly:
            eng.pipeline_batch_wait(tk)

    cabi(3)
    barrier()
    nch = 8
    t0 = time.perf_counter()
    cabi(nch)
    tc = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(tc, op=dist.ReduceOp.MAX)
    return {"value": world * VFRAMES / float(te.item()), "unit": UNIT, "frames_per_rank": VFRAMES,
            "job_seconds_3_runs": [round(r, 4) for r in reps],
            "value_960_frames": world * n4 / float(t4.item()),
            "cabi_async_same_shape_fps": world * nch * Bc / float(tc.item()), "cabi_chunk": Bc,
            "workload": f"video_{VW}x{VH}_D{VD}_sgbm_3way_bgr_frames_rectify+match+post+depth",
            "h2d_bytes_per_frame": 2 * VH * VW * 3, "d2h_bytes_per_frame": VH * (VW - VD) * 4,
            "api": "depthlib.StereoDepthEstimatorVideo(devices=[gpu]).estimate_depth() generator, in-memory frame sources"}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from depthestimation_b200.engine import StereoEngine, make_post_params, make_depth_params

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the engine has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    eng = StereoEngine(local, **SGBM_KW)
    B = args.batch if args.batch > 0 else eng.preferred_batch(H, W)
    L, R = make_inputs(B)
    maps, cal = make_maps()
    eng.set_maps(maps["map1_L"], maps["map2_L"], maps["map1_R"], maps["map2_R"])
    pp = make_post_params(False, 100, 1.0, 2.5, True)            # stereo_core.py:175-184 at downscale 1.0
    dp = make_depth_params(cal["focal_length"], cal["baseline"], 0.0, eps=0, max_depth=None)
    # the engine launches on a dedicated torch stream, so torch.cuda.Event on it times our kernels
    # (NULL / the legacy default stream would make the handle fall back to its own private stream)
    stream = torch.cuda.Stream()
    assert stream.cuda_stream != 0
    eng.set_stream(stream.cuda_stream)

    Wc = W - D
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    disp = torch.empty((B, H, Wc), dtype=torch.float32, device="cuda")
    depth = torch.empty_like(disp)

    def step_dev():
        eng.pipeline_batch_dev(dL.data_ptr(), dR.data_ptr(), B, H, W, pp, dp, disp.data_ptr(), depth.data_ptr(),
                               use_maps=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_dev()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = eng.launch_count
    # per-stage CUDA events (recorded by the library on the launching stream) run INSIDE the timed region: the
    # roofline's launch durations are those of the timed steps themselves (cost: one host sync per step)
    eng.enable_timing(True)
    eng.stage_times(reset=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_dev()
    e1.record(stream)
    barrier()
    stream.synchronize()
    stages = eng.stage_times(reset=True)
    eng.enable_timing(False)
    nst = args.steps
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
    launches = eng.launch_count - l0
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    fps = world * B * args.steps / (total_ms / 1e3)

    V, per_kernel, matcher_bytes = algorithmic_bytes(eng, B)
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    kern = {}
    for st, (name, nbytes) in per_kernel.items():
        k_ms, k_n = stages[st]
        launch_ms = k_ms / max(k_n, 1)
        ach = nbytes / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else 0.0
        kern[st] = {"bound": "hbm", "kernel": name, "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": ach / peak if peak else None, "traffic": ncu_traffic(st, B), "launch_ms": launch_ms,
                    "launches_per_step": k_n / max(nst, 1), "algorithmic_bytes_per_launch": nbytes,
                    "peak_source": peak_src}
    dominant = max(kern, key=lambda k: stages[k][0])
    matcher_ms = sum(stages[k][0] for k in ("prefilter", "cost", "paths", "wta", "median", "speckle")) / nst
    step_ms_sum = sum(v[0] for v in stages.values()) / nst

    # ---- end to end through the host-buffer C-ABI calls (pinned host in / out) -------------------
    # b200s_pipeline_batch_submit / _wait, up to four chunks in flight as a frame loop uses them: every step copies
    # its own inputs host->device and its own results device->host inside the timed region; the copies of step k+1 /
    # k-1 overlap the kernels of step k (the extra slots keep the uploads of k+1, k+2 from waiting for the download of
    # k-1 when the host's PCIe throughput is shared by eight ranks). Four buffer sets rotate.
    # Three output selections: float disparity + float depth (what StereoCore.estimate_depth returns: the headline
    # `e2e`), depth only (what the reference's video generator yields, StereoDepthEstimatorVideo.py:103,130), and
    # int16 x16 disparity + depth (what sgbm.compute itself returns, stereo_core.py:231, 2 B/px).
    from depthestimation_b200.engine import pinned_empty
    DEPTH = 4
    sets = []
    for _ in range(DEPTH):
        nL, nR = pinned_empty(L.shape, np.uint8), pinned_empty(R.shape, np.uint8)
        nL[...] = L; nR[...] = R
        sets.append(dict(L=nL, R=nR, disp=pinned_empty((B, H, Wc), np.float32), depth=pinned_empty((B, H, Wc), np.float32),
                         disp16=pinned_empty((B, H, Wc), np.int16)))

    def e2e_leg(want_disp, want_disp16, want_depth, n):
        def submit(i):
            st = sets[i % DEPTH]
            return eng.pipeline_frames_submit(st["L"], st["R"], pp, dp, use_maps=True, want_disp=want_disp,
                                              want_disp16=want_disp16, want_depth=want_depth, out_disp=st["disp"],
                                              out_disp16=st["disp16"], out_depth=st["depth"])
        probe = "depth" if want_depth else "disp"

        def loop(n):
            sink = 0.0
            inflight = []                                   # (ticket, step)
            for i in range(n):
                if len(inflight) == DEPTH:                  # the slot / buffer set of step i is that of step i - DEPTH
                    tk, j = inflight.pop(0)
                    eng.pipeline_batch_wait(tk)
                    sink += float(sets[j % DEPTH][probe][0, 0, 0])  # device->host result is consumed
                inflight.append((submit(i), i))
            for tk, j in inflight:
                eng.pipeline_batch_wait(tk)
                sink += float(sets[j % DEPTH][probe][0, 0, 0])
            return sink

        loop(max(DEPTH, min(args.warmup, 3)))
        barrier()
        t0 = time.perf_counter()
        loop(n)
        torch.cuda.synchronize()
        mine = time.perf_counter() - t0
        te = torch.tensor([mine], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        d2h = B * H * Wc * (4 * int(want_disp) + 2 * int(want_disp16) + 4 * int(want_depth))
        rates = torch.tensor([d2h * n / mine / 1e9], device="cuda", dtype=torch.float64)
        allr = [torch.zeros_like(rates) for _ in range(world)]
        if world > 1:
            dist.all_gather(allr, rates)
        else:
            allr = [rates]
        return {"value": world * B * n / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": int(2 * B * H * W),
                "d2h_bytes_per_step": int(d2h), "steps": n,
                "d2h_gbs_per_rank": [round(float(r.item()), 2) for r in allr]}

    ke = max(2, min(args.steps, 10))
    e2e = e2e_leg(True, False, True, ke)
    same = bool(all(np.array_equal(st["disp"], sets[0]["disp"]) for st in sets) and
                np.array_equal(sets[0]["disp"], disp.cpu().numpy()) and np.array_equal(sets[0]["depth"], depth.cpu().numpy()))
    e2e_depth = e2e_leg(False, False, True, ke)
    same = same and bool(np.array_equal(sets[0]["depth"], depth.cpu().numpy()))
    e2e_d16 = e2e_leg(False, True, True, ke)
    same = same and bool(np.array_equal(sets[0]["disp16"].astype(np.float32) / np.float32(16), disp.cpu().numpy()))
    e2e_fps = e2e["value"]
    # the blocking one-call form, for reference
    t0 = time.perf_counter()
    for i in range(ke):
        eng.pipeline_batch(sets[0]["L"], sets[0]["R"], pp, dp, use_maps=True, out_disp=sets[0]["disp"], out_depth=sets[0]["depth"])
    e2e_sync_fps = B * ke / (time.perf_counter() - t0)

    # ---- the reference-shaped API: StereoDepthEstimatorVideo on BASELINE config #4 (1280x720 BGR frames, D = 128,
    # default sgbm_3way, calibrated: rectify -> match -> post -> depth, depth maps yielded in frame order) ------------
    video = None
    if not args.no_video:
        video = video_api_leg(local, rank, world, barrier, dist if world > 1 else None, torch)

    # ---- parity of the TIMED workload (rank 0, outside every timed region): the first and the last frame of the batch
    # this run just timed, against cv2 driven as depthlib drives it (C oracle if cv2 is missing) ------------------------
    parity = None
    if rank == 0 and not args.no_parity:
        parity = parity_check(L, R, maps, cal, disp, depth, sets[0]["disp16"], [0, B - 1])

    # ---- CPU baseline beside it (rank 0, N=1 only): bounded sample of the same workload ----------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            ref = CpuReference(maps, cal)
            ref.run_frames(L, R, ref.cores)               # warm-up
            n = 2 * ref.cores
            t = ref.run_frames(L, R, n)
            while t < 8.0 and n < 64 * ref.cores:          # grow the sample to roughly 10-20 s of work
                n *= 2
                t = ref.run_frames(L, R, n)
            cpu = {"value": n / t, "unit": UNIT, "cores": ref.cores, "kind": ref.kind,
                   "sample": f"{n} frames of the same workload, frame-parallel on {ref.cores} threads, {ref.desc}"}
            cpu.update(ref.as_shipped(L, R, 8))
        except Exception as exc:  # the baseline is a report, not a dependency of the GPU number
            cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {exc}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u16x2 packed int (matcher), f32 (post/depth)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_per_step_per_gpu": B, "distinct_pairs": min(N_DISTINCT, B),
                       "cells_per_frame": V, "parallelism": f"frame-sharded x{world}, no collective",
                       "l2": "inputs larger than L2: C+S volumes are %.1f GB per step (L2 = 126 MB)" %
                             (2 * 2 * V * B / 1e9)},
            "mde_per_s": mde(fps),
            "e2e": dict(e2e, mde_per_s=mde(e2e_fps), outputs="float32 disparity + float32 depth",
                        api="b200s_pipeline_frames_submit / b200s_pipeline_batch_wait, up to 4 chunks in flight (page-locked host buffers)",
                        blocking_call_fps_rank0=e2e_sync_fps, matches_device_path=same),
            "e2e_depth_only": dict(e2e_depth, outputs="float32 depth (what the reference's video generator yields)"),
            "e2e_disp16_depth": dict(e2e_d16, outputs="int16 x16 disparity + float32 depth"),
            "e2e_video_api": video,
            "parity_checked": parity,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": kern[dominant],
            "roofline_other_kernels": {k: v for k, v in kern.items() if k != dominant},
            "roofline_matcher": {"bytes_per_cell": 10, "achieved": matcher_bytes / (matcher_ms * 1e-3) / 1e9,
                                 "frac": matcher_bytes / (matcher_ms * 1e-3) / 1e9 / peak, "ms_per_step": matcher_ms,
                                 "nominal_8TBs_frac": matcher_bytes / (matcher_ms * 1e-3) / 1e9 / 8000.0},
            "stage_ms_per_step": {k: v[0] / nst for k, v in stages.items()},
            "stage_ms_sum": step_ms_sum,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    eng.close()
    if rank == 0 and parity is not None and parity["mismatch"]:
        print("bench.py: the timed workload does NOT match the oracle", file=sys.stderr)
        return 3
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0,
                    help="stereo pairs per step and GPU (0: the engine's preferred batch for this shape)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-video", action="store_true", help="skip the StereoDepthEstimatorVideo (config #4) leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the cv2 check of the timed batch")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
