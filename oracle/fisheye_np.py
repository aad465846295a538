"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) — numpy restatement of the
camera math on the AcinoSet TRI + SBA hot path.

Every function cites the reference line it follows. The reference delegates the
arithmetic to OpenCV (un-vendored, unpinned; 4.13.0 in the build image), so the
formulas below restate OpenCV's published fisheye model (Kannala-Brandt, k1..k4),
``cv::fisheye::undistortPoints`` (Newton on theta, <=10 steps, eps 1e-8),
``cv::triangulatePoints`` (4x4 homogeneous DLT, smallest right singular vector)
and ``cv::Rodrigues``. ``tests/test_oracle_golden.py`` pins them against fixtures
produced by the real reference (``oracle/make_golden.py``).
"""
import numpy as np

NONCONVERGED = -1000000.0   # cv::fisheye::undistortPoints marker for failed points


def _intrinsics(k):
    k = np.asarray(k, dtype=np.float64)
    return k[0, 0], k[1, 1], k[0, 2], k[1, 2], k[0, 1] / k[0, 0]


def rodrigues_vec_to_mat(rvec):
    """cv::Rodrigues (vector -> matrix) as used at src/lib/sba.py:29."""
    r = np.asarray(rvec, dtype=np.float64).reshape(3)
    theta = np.linalg.norm(r)
    if theta < np.finfo(np.float64).eps:
        return np.eye(3)
    kx, ky, kz = r / theta
    c, s = np.cos(theta), np.sin(theta)
    c1 = 1.0 - c
    kk = np.array([[kx * kx, kx * ky, kx * kz], [kx * ky, ky * ky, ky * kz], [kx * kz, ky * kz, kz * kz]])
    kx_ = np.array([[0, -kz, ky], [kz, 0, -kx], [-ky, kx, 0]])
    return c * np.eye(3) + c1 * kk + s * kx_


def project_points_fisheye(obj_pts, k, d, r, t):
    """src/lib/calib.py:132-136 -> cv::fisheye::projectPoints. Returns (N, 2) f64."""
    X = np.asarray(obj_pts, dtype=np.float64).reshape(-1, 3)
    fx, fy, cx, cy, alpha = _intrinsics(k)
    d = np.asarray(d, dtype=np.float64).reshape(-1)
    R = np.asarray(r, dtype=np.float64).reshape(3, 3)
    T = np.asarray(t, dtype=np.float64).reshape(3)
    Xc = X @ R.T + T
    a = Xc[:, 0] / Xc[:, 2]
    b = Xc[:, 1] / Xc[:, 2]
    rr = np.sqrt(a * a + b * b)
    th = np.arctan(rr)
    th2 = th * th
    thd = th * (1.0 + th2 * (d[0] + th2 * (d[1] + th2 * (d[2] + th2 * d[3]))))
    with np.errstate(divide='ignore', invalid='ignore'):
        cdist = np.where(rr > 1e-8, thd / rr, 1.0)
    xd = a * cdist
    yd = b * cdist
    return np.stack([fx * (xd + alpha * yd) + cx, fy * yd + cy], axis=1)


def undistort_points_fisheye(pts, k, d):
    """cv::fisheye::undistortPoints(pts, K, D) with default R=I, P=I, criteria (10, 1e-8),
    as called at src/lib/calib.py:123-124. Returns (M, 2) normalised coords (f64)."""
    p = np.asarray(pts, dtype=np.float64).reshape(-1, 2)
    fx, fy, cx, cy, _alpha = _intrinsics(k)
    d = np.asarray(d, dtype=np.float64).reshape(-1)
    pw0 = (p[:, 0] - cx) / fx
    pw1 = (p[:, 1] - cy) / fy
    thd = np.sqrt(pw0 * pw0 + pw1 * pw1)
    thd = np.minimum(np.maximum(-np.pi / 2.0, thd), np.pi / 2.0)
    out = np.empty_like(p)
    for i in range(p.shape[0]):           # scalar Newton, mirrors the per-point loop
        td = thd[i]
        converged = False
        th = td
        scale = 0.0
        if abs(td) > 1e-8:
            for _ in range(10):
                t2 = th * th; t4 = t2 * t2; t6 = t4 * t2; t8 = t4 * t4
                k0 = d[0] * t2; k1 = d[1] * t4; k2 = d[2] * t6; k3 = d[3] * t8
                fix = (th * (1 + k0 + k1 + k2 + k3) - td) / (1 + 3 * k0 + 5 * k1 + 7 * k2 + 9 * k3)
                th = th - fix
                if abs(fix) < 1e-8:
                    converged = True
                    break
            scale = np.tan(th) / td
        else:
            converged = True
        flipped = (td < 0 and th > 0) or (td > 0 and th < 0)
        if converged and not flipped:
            out[i, 0] = pw0[i] * scale
            out[i, 1] = pw1[i] * scale
        else:
            out[i, 0] = NONCONVERGED
            out[i, 1] = NONCONVERGED
    return out


def dlt_two_view(n1, n2, r1, t1, r2, t2):
    """cv::triangulatePoints(P1, P2, n1, n2) + dehomogenise, src/lib/calib.py:125-128.
    P = [R | t] (no K). Rows x*P[2]-P[0], y*P[2]-P[1] per view; X = last right
    singular vector of the 4x4 system."""
    P1 = np.hstack([np.asarray(r1, np.float64).reshape(3, 3), np.asarray(t1, np.float64).reshape(3, 1)])
    P2 = np.hstack([np.asarray(r2, np.float64).reshape(3, 3), np.asarray(t2, np.float64).reshape(3, 1)])
    n1 = np.asarray(n1, np.float64).reshape(-1, 2)
    n2 = np.asarray(n2, np.float64).reshape(-1, 2)
    M = n1.shape[0]
    A = np.empty((M, 4, 4))
    A[:, 0, :] = n1[:, 0:1] * P1[2] - P1[0]
    A[:, 1, :] = n1[:, 1:2] * P1[2] - P1[1]
    A[:, 2, :] = n2[:, 0:1] * P2[2] - P2[0]
    A[:, 3, :] = n2[:, 1:2] * P2[2] - P2[1]
    _, _, vt = np.linalg.svd(A)
    h = vt[:, 3, :]
    return h[:, :3] / h[:, 3:4]


def triangulate_points_fisheye(img_pts_1, img_pts_2, k1, d1, r1, t1, k2, d2, r2, t2):
    """src/lib/calib.py:120-129."""
    n1 = undistort_points_fisheye(img_pts_1, k1, d1)
    n2 = undistort_points_fisheye(img_pts_2, k2, d2)
    return dlt_two_view(n1, n2, r1, t1, r2, t2)


def ring_triangulate_dense(uv, lik, thresh, k_arr, d_arr, r_arr, t_arr):
    """Dense-layout restatement of get_pairwise_3d_points_from_df (src/lib/utils.py:319-349)
    composed with the caller's ``likelihood > dlc_thresh`` filter (src/core/tri.py:41).

    uv (P, C, 2), lik (P, C). Ring pairs (i, (i+1) % C) (utils.py:321); a point exists iff
    at least one ring pair has both views above threshold; xyz = unweighted mean of the
    pair estimates (utils.py:348). Returns xyz (P, 3) (NaN where missing), exists (P,) bool,
    mask (P, C) bool."""
    uv = np.asarray(uv, np.float64)
    P, C = uv.shape[:2]
    mask = np.asarray(lik) > thresh
    d_arr = np.asarray(d_arr, np.float64).reshape(C, 4)
    acc = np.zeros((P, 3))
    cnt = np.zeros(P, dtype=np.int64)
    for a in range(C):
        b = (a + 1) % C
        both = mask[:, a] & mask[:, b]
        idx = np.nonzero(both)[0]
        if idx.size == 0:
            continue
        X = triangulate_points_fisheye(uv[idx, a], uv[idx, b], k_arr[a], d_arr[a], r_arr[a], t_arr[a],
                                       k_arr[b], d_arr[b], r_arr[b], t_arr[b])
        acc[idx] += X
        cnt[idx] += 1
    exists = cnt > 0
    xyz = np.full((P, 3), np.nan)
    xyz[exists] = acc[exists] / cnt[exists, None]
    return xyz, exists, mask


def residuals_points_only(params, n_points, point_idx, cam_idx, k_arr, d_arr, r_arr, t_arr, points_2d):
    """cost_func_points_only, src/lib/sba.py:149-153 (vectorised per camera)."""
    X = np.asarray(params, np.float64).reshape(n_points, 3)
    point_idx = np.asarray(point_idx).astype(np.int64)
    cam_idx = np.asarray(cam_idx).astype(np.int64)
    proj = np.empty((point_idx.size, 2))
    n_cams = len(k_arr)
    d_arr = np.asarray(d_arr, np.float64).reshape(n_cams, 4)
    for j in range(n_cams):
        sel = cam_idx == j
        if sel.any():
            proj[sel] = project_points_fisheye(X[point_idx[sel]], k_arr[j], d_arr[j], r_arr[j], t_arr[j])
    return (proj - np.asarray(points_2d, np.float64)).ravel()


def residuals_points_extrinsics(params, n_cams, n_points, point_idx, cam_idx, k_arr, d_arr, points_2d):
    """cost_func_points_extrinsics + params_to_points_extrinsics, src/lib/sba.py:25-32,142-146."""
    params = np.asarray(params, np.float64)
    rv = params[:3 * n_cams].reshape(n_cams, 3)
    tv = params[3 * n_cams:6 * n_cams].reshape(n_cams, 3)
    r_arr = np.array([rodrigues_vec_to_mat(r) for r in rv])
    return residuals_points_only(params[6 * n_cams:], n_points, point_idx, cam_idx, k_arr, d_arr, r_arr, tv, points_2d)


def cauchy_cost(f, f_scale):
    """scipy least_squares cost for loss='cauchy' (least_squares.py: 0.5 * C**2 * sum(log1p((f/C)**2)))."""
    f = np.asarray(f, np.float64)
    return 0.5 * f_scale ** 2 * np.sum(np.log1p((f / f_scale) ** 2))
