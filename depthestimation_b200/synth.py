"""Deterministic synthetic stereo inputs — numpy only, so the same pairs can be
rebuilt on the GPU box (no cv2 / no reference needed).

A pair is a blurred-noise texture seen through three fronto-parallel disparity planes
(background D/8, centre box D/2, inset box D-8), forward-warped into the right view, plus +-3
noise. Pair ``k`` uses ``seed = 1000 + k``. Also provides the adversarial inputs of the parity
suite (independent noise, all-zero, quantised) and the synthetic calibration used to exercise
rectification.
"""
from __future__ import annotations

import numpy as np

__all__ = ["make_pair", "make_adversarial", "make_calibration", "mde_per_frame"]


def _gauss_blur(img: np.ndarray, sigma: float = 1.2) -> np.ndarray:
    r = max(1, int(round(3 * sigma)))
    xs = np.arange(-r, r + 1, dtype=np.float64)
    k = np.exp(-(xs * xs) / (2 * sigma * sigma))
    k /= k.sum()
    f = img.astype(np.float64)
    p = np.pad(f, ((0, 0), (r, r)), mode="reflect")
    f = sum(k[i] * p[:, i:i + f.shape[1]] for i in range(2 * r + 1))
    p = np.pad(f, ((r, r), (0, 0)), mode="reflect")
    f = sum(k[i] * p[i:i + f.shape[0], :] for i in range(2 * r + 1))
    return f


def make_pair(H: int, W: int, D: int, seed: int = 1000, contrast: float = 3.0):
    """Return (left u8 HxW, right u8 HxW, gt_disparity int32 HxW)."""
    rng = np.random.default_rng(seed)
    base = _gauss_blur(rng.integers(0, 256, (H, W + D), dtype=np.uint8))
    base = np.clip((base - 127.5) * contrast + 127.5, 0, 255)
    left = base[:, D:D + W]
    gt = np.full((H, W), max(D // 8, 1), dtype=np.int32)
    gt[H // 4:3 * H // 4, W // 4:3 * W // 4] = max(D // 2, 1)
    gt[3 * H // 8:5 * H // 8, 3 * W // 8:5 * W // 8] = max(D - 8, 1)
    right = np.zeros((H, W), dtype=np.float64)
    # forward warp, far-to-near so nearer surfaces overwrite farther ones
    xs = np.arange(W)
    for d in np.unique(gt):
        m = gt == d
        ys, xl = np.nonzero(m)
        xr = xl - d
        ok = xr >= 0
        right[ys[ok], xr[ok]] = left[ys[ok], xl[ok]]
    del xs
    noise = rng.integers(-3, 4, (H, W))
    right = np.clip(np.rint(right) + noise, 0, 255).astype(np.uint8)
    left = np.clip(np.rint(left), 0, 255).astype(np.uint8)
    return np.ascontiguousarray(left), np.ascontiguousarray(right), gt


def make_adversarial(kind: str, H: int, W: int, seed: int = 7):
    """'noise' (independent uniform L/R), 'zeros', 'quantized' ((img//32)*32), 'periodic'."""
    rng = np.random.default_rng(seed)
    if kind == "noise":
        return (rng.integers(0, 256, (H, W), dtype=np.uint8),
                rng.integers(0, 256, (H, W), dtype=np.uint8))
    if kind == "zeros":
        z = np.zeros((H, W), dtype=np.uint8)
        return z, z.copy()
    if kind == "quantized":
        l, r, _ = make_pair(H, W, max(8, min(64, W // 4)), seed)
        return (l // 32) * 32, (r // 32) * 32
    if kind == "periodic":
        x = np.arange(W)
        row = (((x // 4) % 2) * 200 + 20).astype(np.uint8)
        l = np.tile(row, (H, 1))
        return l, np.roll(l, -3, axis=1).copy()
    raise ValueError(kind)


def make_calibration(H: int, W: int):
    """Synthetic calibration with mild radial distortion."""
    K = np.array([[0.8 * W, 0, W / 2], [0, 0.8 * W, H / 2], [0, 0, 1]], dtype=np.float64)
    return dict(
        cam_matrix_L=K, cam_matrix_R=K.copy(),
        dist_coeff_L=np.array([-0.05, 0.01, 0, 0, 0], dtype=np.float64),
        dist_coeff_R=np.array([-0.04, 0.01, 0, 0, 0], dtype=np.float64),
        baseline=0.54, image_width=W, image_height=H, focal_length=0.8 * W,
    )


def mde_per_frame(H: int, W: int, D: int) -> float:
    """Nominal million disparity evaluations per frame."""
    return H * W * D / 1e6
