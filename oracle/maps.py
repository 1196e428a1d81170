"""oracle/maps.py — TEST INFRASTRUCTURE ONLY (imported by tests/ only; the product path is csrc/imgops.cu).

numpy double-precision restatement of ``cv2.initUndistortRectifyMap(K, dist, R, P, size, cv2.CV_32FC1)`` as the reference
calls it at depthlib/rectify.py:68-73 and :229-234. The arithmetic lives in the third-party opencv wheel (no source on this
box); the published algorithm is restated here and pinned bit-exact against the live cv2 binary in
tests/test_oracle_vs_cv2.py::test_maps_oracle_matches_cv2.

    iR = (P[:, :3] R)^-1 ; (x, y, w) = iR (j, i, 1) ; x /= w ; y /= w ; r2 = x^2 + y^2
    kr = (1 + ((k3 r2 + k2) r2 + k1) r2) / (1 + ((k6 r2 + k5) r2 + k4) r2)
    xd = x kr + p1 2xy + p2 (r2 + 2 x^2) + s1 r2 + s2 r2^2 ;  yd likewise with (p1, p2) swapped and (s3, s4)
    (xd, yd, 1) through the tilt projection, map1 = fx xd' + u0, map2 = fy yd' + v0, rounded to float32
"""
import numpy as np


def _inv3(a):
    a = a.ravel()
    d = a[0] * (a[4] * a[8] - a[5] * a[7]) - a[1] * (a[3] * a[8] - a[5] * a[6]) + a[2] * (a[3] * a[7] - a[4] * a[6])
    r = 1.0 / d
    return np.array([(a[4] * a[8] - a[5] * a[7]) * r, (a[2] * a[7] - a[1] * a[8]) * r, (a[1] * a[5] - a[2] * a[4]) * r,
                     (a[5] * a[6] - a[3] * a[8]) * r, (a[0] * a[8] - a[2] * a[6]) * r, (a[2] * a[3] - a[0] * a[5]) * r,
                     (a[3] * a[7] - a[4] * a[6]) * r, (a[1] * a[6] - a[0] * a[7]) * r, (a[0] * a[4] - a[1] * a[3]) * r])


def tilt_matrix(tx, ty):
    cx, sx, cy, sy = np.cos(tx), np.sin(tx), np.cos(ty), np.sin(ty)
    rx = np.array([[1, 0, 0], [0, cx, sx], [0, -sx, cx]], np.float64)
    ry = np.array([[cy, 0, -sy], [0, 1, 0], [sy, 0, cy]], np.float64)
    rxy = ry @ rx
    pz = np.array([[rxy[2, 2], 0, -rxy[0, 2]], [0, rxy[2, 2], -rxy[1, 2]], [0, 0, 1]], np.float64)
    return pz @ rxy


def init_undistort_rectify_map(K, dist, R, P, size):
    K = np.asarray(K, np.float64)
    d = np.zeros(14)
    if dist is not None:
        dv = np.asarray(dist, np.float64).ravel()
        d[:dv.size] = dv
    k1, k2, p1, p2, k3, k4, k5, k6, s1, s2, s3, s4, tx, ty = d
    Rm = np.eye(3) if R is None else np.asarray(R, np.float64)
    Pm = K if P is None else np.asarray(P, np.float64)[:, :3]
    A = np.zeros((3, 3))
    for r in range(3):
        for c in range(3):
            acc = 0.0
            for k in range(3):
                acc += Pm[r, k] * Rm[k, c]
            A[r, c] = acc
    ir = _inv3(A)
    W, H = int(size[0]), int(size[1])
    i = np.arange(H, dtype=np.float64)[:, None]
    j = np.arange(W, dtype=np.float64)[None, :]
    _x = (i * ir[1] + ir[2]) + j * ir[0]
    _y = (i * ir[4] + ir[5]) + j * ir[3]
    _w = (i * ir[7] + ir[8]) + j * ir[6]
    w = 1.0 / _w
    x, y = _x * w, _y * w
    x2, y2 = x * x, y * y
    r2, _2xy = x2 + y2, 2.0 * x * y
    kr = (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1.0 + ((k6 * r2 + k5) * r2 + k4) * r2)
    xd = x * kr + p1 * _2xy + p2 * (r2 + 2.0 * x2) + s1 * r2 + (s2 * r2) * r2
    yd = y * kr + p1 * (r2 + 2.0 * y2) + p2 * _2xy + s3 * r2 + (s4 * r2) * r2
    T = tilt_matrix(tx, ty) if (tx != 0.0 or ty != 0.0) else np.eye(3)
    t0 = (T[0, 0] * xd + T[0, 1] * yd) + T[0, 2]
    t1 = (T[1, 0] * xd + T[1, 1] * yd) + T[1, 2]
    t2 = (T[2, 0] * xd + T[2, 1] * yd) + T[2, 2]
    inv = np.where(t2 != 0.0, 1.0 / np.where(t2 != 0.0, t2, 1.0), 1.0)
    u = (K[0, 0] * inv) * t0 + K[0, 2]
    v = (K[1, 1] * inv) * t1 + K[1, 2]
    return u.astype(np.float32), v.astype(np.float32)


def calibration_cases():
    """(name, K_L, d_L, K_R, d_R, size, R, T): the synthetic calibration of SURVEY 8(d), a rotated / translated rig with
    different cameras, and the 8-, 12- and 14-coefficient distortion models."""
    out = []
    W, H = 320, 200
    K = np.array([[0.8 * W, 0, W / 2], [0, 0.8 * W, H / 2], [0, 0, 1]], np.float64)
    out.append(("synthetic", K, np.array([-0.05, 0.01, 0, 0, 0.0]), K.copy(), np.array([-0.04, 0.01, 0, 0, 0.0]), (W, H),
                np.eye(3), np.array([-0.54, 0, 0.0])))
    import cv2
    Rr, _ = cv2.Rodrigues(np.array([0.012, -0.02, 0.006]))
    K2 = np.array([[262.0, 0.3, 158.0], [0, 259.5, 97.0], [0, 0, 1]], np.float64)
    out.append(("rotated", K, np.array([-0.11, 0.03, 0.001, -0.002, 0.004]), K2, np.array([-0.09, 0.02, -0.001, 0.001, 0.0]),
                (W, H), Rr, np.array([-0.31, 0.004, -0.006])))
    out.append(("rational8", K, np.array([0.2, -0.1, 0.001, 0.0005, 0.02, 0.25, -0.08, 0.01]), K2,
                np.array([0.1, 0.05, 0.0, 0.001, 0.01, 0.12, 0.04, 0.02]), (W, H), Rr, np.array([-0.2, 0.0, 0.0])))
    out.append(("prism12", K, np.array([-0.05, 0.01, 0.001, 0.0, 0.0, 0, 0, 0, 0.002, -0.001, 0.0015, 0.0007]), K2,
                np.array([-0.04, 0.0, 0.0, 0.001, 0.0, 0.01, 0, 0, -0.001, 0.0004, 0.0, 0.0011]), (W, H), np.eye(3),
                np.array([-0.2, 0.0, 0.0])))
    out.append(("tilt14", K, np.array([-0.05, 0.01, 0, 0, 0.0, 0, 0, 0, 0, 0, 0, 0, 0.01, -0.02]), K2,
                np.array([-0.04, 0.01, 0, 0, 0.0, 0, 0, 0, 0.001, 0, 0, 0, -0.015, 0.005]), (W, H), np.eye(3),
                np.array([-0.2, 0.0, 0.0])))
    return out
