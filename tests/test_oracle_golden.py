"""The C oracle against the golden vectors recorded from the reference's own StereoCore
(tools/make_golden.py). CPU-only; runs here and on the GPU box."""
import numpy as np
import pytest

import oracle
from conftest import golden_cases
from helpers import golden_sgbm_kwargs, golden_max_depth


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_disp16_matches_reference(name, golden):
    g = golden(name)
    kw = golden_sgbm_kwargs(g)
    got = oracle.sgbm_compute(g["left_rect"], g["right_rect"], oracle.make_params(**kw))
    assert got.dtype == np.int16 and got.shape == g["disp16"].shape
    assert np.array_equal(got, g["disp16"]), f"{(got != g['disp16']).sum()} pixels differ"


@pytest.mark.parametrize("name", [n for n in golden_cases() if n.startswith("calib")])
def test_oracle_remap_matches_reference(name, golden):
    g = golden(name)
    assert np.array_equal(oracle.remap_u8(g["left"], g["map1_L"], g["map2_L"]), g["left_rect"])
    assert np.array_equal(oracle.remap_u8(g["right"], g["map1_R"], g["map2_R"]), g["right_rect"])


@pytest.mark.parametrize("name", golden_cases())
def test_oracle_post_and_depth_match_reference(name, golden):
    g = golden(name)
    kw = golden_sgbm_kwargs(g)
    d = g["disp16"].astype(np.float32) / 16.0
    d = d[:, kw["num_disp"]:]
    if int(g["fast"]):
        post = oracle.median3_f32(d)
    else:
        post = oracle.postprocess(d, max_speckle_size=int(100 * float(g["downscale"])), max_diff=1.0,
                                  outlier_threshold=2.5)
    assert np.array_equal(post, g["disp"])
    depth = oracle.disp_to_depth(post, float(g["focal_length"]), float(g["baseline"]), float(g["doffs"]),
                                 eps=kw["min_disp"], max_depth=golden_max_depth(g))
    ref = g["depth"]
    assert np.array_equal(np.isinf(depth), np.isinf(ref))
    fin = np.isfinite(ref)
    assert np.allclose(depth[fin], ref[fin], rtol=1e-6, atol=0)


@pytest.mark.parametrize("name", golden_cases())
def test_cv2_reference_pipeline_matches_reference(name, golden):
    """The assembled cv2 pipeline that `bench.py --impl reference` and its cpu_baseline leg time
    (oracle/cv2_reference.py::Cv2StereoPipeline) against what the reference's own StereoCore.estimate_depth produced for
    the same inputs (tests/golden, recorded by tools/make_golden.py): rectified pair -> post-processed disparity bit for
    bit, depth within 1e-6 relative. Pins the reference arm itself, not just its pieces."""
    from oracle import cv2_reference as cr
    if not cr.HAVE_CV2:
        pytest.skip("cv2 not importable")
    g = golden(name)
    kw = golden_sgbm_kwargs(g)
    pipe = cr.Cv2StereoPipeline(kw, calib=None, downscale=float(g["downscale"]), fast_mode=bool(int(g["fast"])),
                                focal_length=float(g["focal_length"]), baseline=float(g["baseline"]),
                                doffs=float(g["doffs"]), max_depth=golden_max_depth(g))
    if "map1_L" in g:      # calibrated cases: the maps the reference built, raw frames in
        pipe.maps = {k: g[k] for k in ("map1_L", "map2_L", "map1_R", "map2_R")}
        disp, depth = pipe(g["left"], g["right"])
    else:
        disp, depth = pipe(g["left_rect"], g["right_rect"])
    assert np.array_equal(disp, g["disp"]), f"{(disp != g['disp']).sum()} px differ"
    ref = g["depth"]
    assert np.array_equal(np.isinf(depth), np.isinf(ref))
    fin = np.isfinite(ref)
    assert np.allclose(depth[fin], ref[fin], rtol=1e-6, atol=0)
