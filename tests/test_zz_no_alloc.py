"""Allocation-free speckle filter (b200s_filter_speckles_noalloc, include/b200stereo.h) against the existing entry point.

Collected last on purpose (file name). The new entry point keeps its labels / sizes
scratch in a static __device__ array; it is called through the C ABI on DEVICE buffers (torch is used for memory only)
and compared, bit for bit (int16 work: the tolerance of the other speckle tests is zero), with
``b200s_postprocess(B200S_POST_SPECKLE_ONLY)`` == postprocess.filter_speckles (depthlib/postprocess.py:6-35) on the same
image, on two calls in a row with different images (the second call sees the first call's scratch contents).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _speckled(rng, H, W):
    """Piecewise-constant int16 'disparity' (x16) with small islands, zeros and isolated noise."""
    img = np.full((H, W), 40 * 16, np.int16)
    img[H // 4: H // 2, W // 3: 2 * W // 3] = 90 * 16
    for _ in range(300):
        y, x = int(rng.integers(0, H - 6)), int(rng.integers(0, W - 6))
        h, w = int(rng.integers(1, 7)), int(rng.integers(1, 7))
        img[y:y + h, x:x + w] = int(rng.integers(0, 128)) * 16
    noise = rng.random((H, W)) < 0.02
    img[noise] = rng.integers(-16, 2048, int(noise.sum())).astype(np.int16)
    img[rng.random((H, W)) < 0.01] = 0
    img += rng.integers(0, 3, (H, W)).astype(np.int16)
    return img


def _existing(eng, img16, max_size, max_diff16):
    from depthestimation_b200._lib import B200sPostParams
    pp = B200sPostParams(2, int(max_size), int(max_diff16), 2.5, 0)   # B200S_POST_SPECKLE_ONLY
    out = eng.postprocess(img16.astype(np.float32) / np.float32(16.0), pp)
    return np.rint(out * np.float32(16.0)).astype(np.int16)


def test_noalloc_speckle_matches_existing_entry_point_twice():
    import torch
    from depthestimation_b200 import _lib
    from depthestimation_b200.engine import StereoEngine
    lib = _lib.load()
    eng = StereoEngine(0)
    torch.cuda.set_device(0)
    rng = np.random.default_rng(77)
    err = C.create_string_buffer(256)
    stream = torch.cuda.Stream()
    # two calls in a row, different shapes and contents; the second runs in place on a non-default stream
    cases = [(375, 1114, 100, 16, False), (271, 333, 37, 24, True)]
    # the device buffers exist before the calls; the calls themselves allocate nothing
    bufs = []
    for H, W, _, _, inplace in cases:
        img = _speckled(rng, H, W)
        d_src = torch.from_numpy(img).cuda()
        d_dst = d_src if inplace else torch.empty_like(d_src)
        bufs.append((img, d_src, d_dst))
    torch.cuda.synchronize()
    for (H, W, max_size, max_diff, inplace), (img, d_src, d_dst) in zip(cases, bufs):
        st = stream.cuda_stream if inplace else 0
        rc = lib.b200s_filter_speckles_noalloc(d_src.data_ptr(), d_dst.data_ptr(), H, W, 0, max_size, max_diff, st,
                                               err, len(err))
        assert rc == 0, err.value
        torch.cuda.synchronize()
        got = d_dst.cpu().numpy()
        want = _existing(eng, img, max_size, max_diff)
        assert (want != img).any(), "the case must actually remove speckles"
        np.testing.assert_array_equal(got, want)
    eng.close()


def test_noalloc_speckle_rejects_oversize():
    import torch
    from depthestimation_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.int16, device="cuda")
    err = C.create_string_buffer(256)
    rc = lib.b200s_filter_speckles_noalloc(d.data_ptr(), d.data_ptr(), 2048, 1025, 0, 10, 16, 0, err, len(err))
    assert rc == _lib.E_INVALID
    assert b"B200S_NOALLOC_MAX_PIXELS" in err.value
    rc = lib.b200s_filter_speckles_noalloc(None, d.data_ptr(), 4, 4, 0, 10, 16, 0, err, len(err))
    assert rc == _lib.E_INVALID and b"null" in err.value
