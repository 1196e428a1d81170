"""ctypes binding of libb200stereo.so (include/b200stereo.h). There is no CPU fallback: if the
shared library is missing this raises, and every compute call fails without a CUDA device."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libb200stereo.so"
CSRC = _PKG / "csrc"

MODES = {"sgbm": 0, "hh": 1, "sgbm_3way": 2, "hh4": 3}

OK, E_INVALID, E_CUDA, E_STATE = 0, -1, -2, -3
FLOW_VERT_DOWN, FLOW_VERT_UP, FLOW_PATH_VERTICAL, FLOW_ROWS_FUSED = 1, 2, 4, 8
STAGES = ("remap", "prefilter", "cost", "paths", "wta", "median", "speckle", "post", "depth")


class B200sParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "minDisparity", "numDisparities", "blockSize", "P1", "P2", "disp12MaxDiff", "preFilterCap",
        "uniquenessRatio", "speckleWindowSize", "speckleRange", "mode")]


class B200sPostParams(C.Structure):
    _fields_ = [("fastMode", C.c_int32), ("maxSpeckleSize", C.c_int32), ("maxDiff16", C.c_int32),
                ("outlierThreshold", C.c_float), ("applyOutlierRemoval", C.c_int32)]


class B200sFrameIO(C.Structure):
    _fields_ = [("L", C.c_void_p), ("R", C.c_void_p), ("B", C.c_int32), ("srcH", C.c_int32), ("srcW", C.c_int32),
                ("channels", C.c_int32), ("areaH", C.c_int32), ("areaW", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
                ("use_maps", C.c_int32), ("disp", C.c_void_p), ("disp16", C.c_void_p), ("depth", C.c_void_p)]


class B200sDepthParams(C.Structure):
    _fields_ = [("enabled", C.c_int32), ("fB", C.c_float), ("doffs", C.c_float), ("eps", C.c_float),
                ("maxDepth", C.c_float), ("hasMaxDepth", C.c_int32)]


# every symbol include/b200stereo.h declares: name -> (restype, argtypes)
_H = C.c_void_p
_VP = C.c_void_p
_PP, _DP = C.POINTER(B200sPostParams), C.POINTER(B200sDepthParams)
_IP = C.POINTER(C.c_int)
SYMBOLS = {
    "b200s_abi_version": (C.c_int, []),
    "b200s_create": (C.c_int, [C.c_int, C.POINTER(B200sParams), C.POINTER(_H)]),
    "b200s_configure": (C.c_int, [_H, C.POINTER(B200sParams)]),
    "b200s_destroy": (C.c_int, [_H]),
    "b200s_last_error": (C.c_char_p, [_H]),
    "b200s_set_stream": (C.c_int, [_H, _VP]),
    "b200s_sync": (C.c_int, [_H]),
    "b200s_set_maps": (C.c_int, [_H, _VP, _VP, _VP, _VP, C.c_int, C.c_int]),
    "b200s_bgr2gray": (C.c_int, [_H, _VP, C.c_int, C.c_int, C.c_size_t, _VP]),
    "b200s_resize_area_u8": (C.c_int, [_H, _VP, C.c_int, C.c_int, C.c_int, C.c_size_t, _VP, C.c_int, C.c_int]),
    "b200s_resize_linear_u8": (C.c_int, [_H, _VP, C.c_int, C.c_int, C.c_int, C.c_size_t, _VP, C.c_int, C.c_int]),
    "b200s_remap_u8": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_size_t, _VP, _VP]),
    "b200s_sgbm_compute": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_size_t, _VP]),
    "b200s_postprocess": (C.c_int, [_H, _VP, C.c_int, C.c_int, _PP]),
    "b200s_fill_nearest": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_int]),
    "b200s_disp_to_depth": (C.c_int, [_H, _VP, C.c_size_t, _DP, _VP]),
    "b200s_pipeline": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_size_t, C.c_int, _PP, _DP,
                                 _VP, _VP, _VP, _VP]),
    "b200s_sgbm_compute_batch_dev": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_int, _VP]),
    "b200s_pipeline_batch_dev": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_int, C.c_int, _PP, _DP, _VP, _VP]),
    "b200s_pipeline_batch": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_int, C.c_int, _PP, _DP, _VP, _VP]),
    "b200s_pipeline_batch_submit": (C.c_int, [_H, _VP, _VP, C.c_int, C.c_int, C.c_int, C.c_int, _PP, _DP, _VP, _VP, _IP]),
    "b200s_pipeline_batch_wait": (C.c_int, [_H, C.c_int]),
    "b200s_pipeline_frames_submit": (C.c_int, [_H, C.POINTER(B200sFrameIO), _PP, _DP, _IP]),
    "b200s_pipeline_frames": (C.c_int, [_H, C.POINTER(B200sFrameIO), _PP, _DP]),
    "b200s_last_flow": (C.c_int, [_H]),
    "b200s_build_maps": (C.c_int, [_H, _VP, _VP, C.c_int, _VP, _VP, C.c_int, C.c_int, _VP, _VP]),
    "b200s_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    "b200s_host_free": (C.c_int, [C.c_void_p]),
    "b200s_filter_speckles_noalloc": (C.c_int, [_VP, _VP, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _VP,
                                                C.c_char_p, C.c_size_t]),
    "b200s_stage_times": (C.c_int, [_H, C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.c_int]),
    "b200s_reset_stage_times": (C.c_int, [_H]),
    "b200s_geometry": (C.c_int, [_H, C.c_int, _IP, _IP, _IP, _IP]),
    "b200s_debug_stage": (C.c_int, [_H, C.c_int, C.c_int, _VP]),
    "b200s_preferred_batch": (C.c_int, [_H, C.c_int, C.c_int, _IP]),
    "b200s_set_debug_flow": (C.c_int, [_H, C.c_int]),
    "b200s_launch_count": (C.c_longlong, [_H]),
    "b200s_enable_timing": (C.c_int, [_H, C.c_int]),
    "b200s_last_ms": (C.c_float, [_H]),
}

_lib = None


def build(verbose: bool = False) -> Path:
    """Compile the CUDA sources for sm_100a into depthestimation_b200/libb200stereo.so (make)."""
    r = subprocess.run(["make", "-C", str(CSRC), "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libb200stereo.so failed")
    return LIB_PATH


def load() -> C.CDLL:
    """Load the engine. Raises (no fallback) if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = Path(os.environ.get("B200S_LIB", LIB_PATH))   # B200S_LIB: a tuning build of the same engine (tools/)
    if not path.exists():
        raise RuntimeError(
            f"{path} not found: build the CUDA engine first (python -c 'import __graft_entry__ as g; "
            f"g.build()' or make -C {CSRC}). There is no CPU fallback.")
    lib = C.CDLL(str(path))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
