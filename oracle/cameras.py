"""Camera projection (oracle; test infrastructure).

  fisheye  acinoset_misc.py:1663-1679  pt3d_to_2d_fisheye  (Kannala-Brandt, `r + 1e-12`, no skew)
  pinhole  acinoset_misc.py:1682-1696  pt3d_to_2d          (radial r^2,r^4,r^6 with D[0..2]; kinetic dataset)

A camera is the 20-vector  [fx, fy, cx, cy, D0, D1, D2, D3, R(9 row-major), t(3)].
"""
import numpy as np

FISHEYE = 0
PINHOLE = 1


def pack_camera(K, D, R, t):
    K = np.asarray(K, float)
    D = np.asarray(D, float).reshape(-1)
    cam = np.zeros(20)
    cam[0], cam[1], cam[2], cam[3] = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    cam[4:4 + min(4, D.size)] = D[:4]
    cam[8:17] = np.asarray(R, float).reshape(9)
    cam[17:20] = np.asarray(t, float).reshape(3)
    return cam


def project(X, cam, model=FISHEYE):
    """X[...,3] world points, cam[...,20] broadcastable -> (u,v)[...,2]."""
    X = np.asarray(X, float)
    cam = np.asarray(cam, float)
    R = cam[..., 8:17]
    x = X[..., 0] * R[..., 0] + X[..., 1] * R[..., 1] + X[..., 2] * R[..., 2] + cam[..., 17]
    y = X[..., 0] * R[..., 3] + X[..., 1] * R[..., 4] + X[..., 2] * R[..., 5] + cam[..., 18]
    z = X[..., 0] * R[..., 6] + X[..., 1] * R[..., 7] + X[..., 2] * R[..., 8] + cam[..., 19]
    a = x / z
    b = y / z
    r = (a ** 2 + b ** 2) ** 0.5
    if model == FISHEYE:
        th = np.arctan(r)
        th_d = th * (1 + cam[..., 4] * th ** 2 + cam[..., 5] * th ** 4 + cam[..., 6] * th ** 6 + cam[..., 7] * th ** 8)
        xp = a * th_d / (r + 1e-12)
        yp = b * th_d / (r + 1e-12)
    else:
        d = 1 + cam[..., 4] * r ** 2 + cam[..., 5] * r ** 4 + cam[..., 6] * r ** 6
        xp = a * d
        yp = b * d
    u = cam[..., 0] * xp + cam[..., 2]
    v = cam[..., 1] * yp + cam[..., 3]
    return np.stack([u, v], -1)


def project_jacobian(X, cam, model=FISHEYE):
    """d(u,v)/dX  [...,2,3], analytic, consistent with `project` (incl. the +1e-12)."""
    X = np.asarray(X, float)
    cam = np.asarray(cam, float)
    R = cam[..., 8:17].reshape(cam.shape[:-1] + (3, 3))
    Xc = np.einsum("...ij,...j->...i", R, X) + cam[..., 17:20]
    x, y, z = Xc[..., 0], Xc[..., 1], Xc[..., 2]
    a = x / z
    b = y / z
    r2 = a * a + b * b
    r = np.sqrt(r2)
    if model == FISHEYE:
        th = np.arctan(r)
        t2 = th * th
        poly = 1 + cam[..., 4] * t2 + cam[..., 5] * t2 ** 2 + cam[..., 6] * t2 ** 3 + cam[..., 7] * t2 ** 4
        dpoly = 1 + 3 * cam[..., 4] * t2 + 5 * cam[..., 5] * t2 ** 2 + 7 * cam[..., 6] * t2 ** 3 + 9 * cam[..., 7] * t2 ** 4
        th_d = th * poly
        s = th_d / (r + 1e-12)                       # x' = a * s
        # ds/dr = (dth_d/dr * (r+eps) - th_d) / (r+eps)^2 ; dth_d/dr = dpoly / (1+r^2)
        ds_dr = (dpoly / (1 + r2) * (r + 1e-12) - th_d) / (r + 1e-12) ** 2
        with np.errstate(divide="ignore", invalid="ignore"):
            g = np.where(r > 0, ds_dr / r, 0.0)      # ds/da = g*a, ds/db = g*b
    else:
        s = 1 + cam[..., 4] * r2 + cam[..., 5] * r2 ** 2 + cam[..., 6] * r2 ** 3
        g = 2 * cam[..., 4] + 4 * cam[..., 5] * r2 + 6 * cam[..., 6] * r2 ** 2
    dxp_da = s + a * a * g
    dxp_db = a * b * g
    dyp_da = a * b * g
    dyp_db = s + b * b * g
    # d(a,b)/dXc
    da = np.stack([1 / z, np.zeros_like(z), -a / z], -1)
    db = np.stack([np.zeros_like(z), 1 / z, -b / z], -1)
    du_dXc = cam[..., 0, None] * (dxp_da[..., None] * da + dxp_db[..., None] * db)
    dv_dXc = cam[..., 1, None] * (dyp_da[..., None] * da + dyp_db[..., None] * db)
    J = np.stack([du_dXc, dv_dXc], -2)               # [...,2,3] wrt Xc
    return J @ R
