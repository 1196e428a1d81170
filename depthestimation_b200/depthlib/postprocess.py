"""Post-processing of disparity maps — drop-in for depthlib/postprocess.py of the reference.

``filter_speckles``, ``detect_outliers`` and ``postprocess_disparity`` run on the GPU through the C ABI
(``b200s_postprocess``; reference: postprocess.py:6-70, 120-171). ``fill_holes`` (off by default,
``hole_filling=True`` in the parameter dict): the 'nearest' branch runs on the GPU (``b200s_fill_nearest``); 'inpaint'
is cv2.inpaint(TELEA), a sequential fast-marching method that SURVEY.md §8f.2 keeps on host cv2 — applied between
the GPU outlier mask and the GPU median. Nothing else in this module touches the host.
"""
from __future__ import annotations

import numpy as np

from ..engine import make_post_params
from .rectify import _engine


def filter_speckles(disparity, max_speckle_size=100, max_diff=1, engine=None):
    """Remove connected regions of at most ``max_speckle_size`` pixels (postprocess.py:6-35):
    (d*16) -> int16 -> filterSpeckles(newVal=0, maxDiff=int(max_diff*16)) -> /16."""
    pp = make_post_params(False, max_speckle_size, max_diff, 0.0, False)
    pp.fastMode = 2  # speckle stage only
    return _engine(engine).postprocess(np.asarray(disparity, dtype=np.float32), pp)


def detect_outliers(disparity, threshold=3.0, kernel_size=5, engine=None):
    """Boolean mask of pixels further than ``threshold`` local standard deviations from the 5x5 local
    mean, valid (>0) pixels only (postprocess.py:37-70). Only the reference's kernel_size=5 is on the GPU."""
    if kernel_size != 5:
        raise ValueError("detect_outliers: the GPU path implements the reference's 5x5 window only")
    d = np.asarray(disparity, dtype=np.float32)
    pp = make_post_params(False, 0, 0.0, threshold, True)
    pp.fastMode = 3  # outlier stage only: outliers come back as 0
    masked = _engine(engine).postprocess(d, pp)
    return (masked != d) & (d > 0)


_FILL_NEAREST_MAX_K = 31


def fill_holes(disparity, mask=None, method='inpaint', kernel_size=5, engine=None):
    """Hole filling (postprocess.py:72-118). method='nearest' (iterated dilation under the hole mask) runs on the GPU
    (``b200s_fill_nearest``, bit-exact) for 2-D float32 maps and element sizes 1..31; anything else raises (no silent
    host fallback). method='inpaint' is Telea fast marching, a sequential method that SURVEY 8(f2) keeps on host cv2:
    that is the one documented host step of this module. Other methods return an unchanged copy, as upstream."""
    if mask is None:
        mask = (disparity <= 0)
    if method == 'nearest':
        d = np.asarray(disparity)
        if d.dtype != np.float32 or d.ndim != 2:
            raise ValueError("fill_holes(method='nearest') runs on the GPU for 2-D float32 disparity maps only "
                             f"(got {d.dtype}, {d.ndim}-D); there is no CPU fallback")
        if not 1 <= int(kernel_size) <= _FILL_NEAREST_MAX_K:
            raise ValueError(f"fill_holes(method='nearest'): kernel_size must be in [1, {_FILL_NEAREST_MAX_K}] "
                             f"(got {kernel_size}); there is no CPU fallback")
        return _engine(engine).fill_nearest(d, mask, int(kernel_size))
    filled = disparity.copy()
    if method == 'inpaint':
        import cv2
        hole_mask = mask.astype(np.uint8) * 255
        filled = cv2.inpaint(filled.astype(np.float32), hole_mask, kernel_size, cv2.INPAINT_TELEA)
    return filled


def postprocess_disparity(disparity, **kwargs):
    """speckles -> outlier masking -> [hole filling] -> 3x3 median (postprocess.py:120-171).
    Keyword defaults are the reference's *code* defaults: max_speckle_size 50, max_diff 1,
    outlier_threshold 3.0, apply_outlier_removal True, apply_hole_filling True, fill_kernel 3."""
    eng = _engine(kwargs.get('engine'))
    d = np.asarray(disparity, dtype=np.float32)
    speckle = kwargs.get('max_speckle_size', 50)
    max_diff = kwargs.get('max_diff', 1)
    thr = kwargs.get('outlier_threshold', 3.0)
    outl = kwargs.get('apply_outlier_removal', True)
    if kwargs.get('outlier_kernel', 5) != 5:
        raise ValueError("postprocess_disparity: the GPU path implements the reference's 5x5 outlier window only")
    if not kwargs.get('apply_hole_filling', True):
        return eng.postprocess(d, make_post_params(False, speckle, max_diff, thr, outl))   # one C-ABI call
    # hole filling requested: GPU stages around the host inpaint
    pp = make_post_params(False, speckle, max_diff, thr, outl)
    pp.fastMode = 4  # speckle (+ outliers), no median
    result = eng.postprocess(d, pp)
    result = fill_holes(result, method=kwargs.get('fill_method', 'inpaint'), kernel_size=kwargs.get('fill_kernel', 3), engine=eng)
    return eng.postprocess(result.astype(np.float32), make_post_params(True))
