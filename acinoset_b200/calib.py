"""Drop-in for the fisheye camera functions of the reference's lib/calib.py (src/lib/calib.py:120-136).

Same names, argument order, accepted shapes and return shapes; the arithmetic runs in the CUDA
library (K1 / K2). fp32 inputs are up-cast to fp64 (the reference's board path passes fp32 pixels,
src/lib/sba.py:88-89).
"""
from typing import Dict
import numpy as np
from . import _capi


def project_points_fisheye(obj_pts, k, d, r, t, shutter_delay_param: Dict = None):
    """src/lib/calib.py:132-136. obj_pts (N,3) or (3,) -> (N,2) float64 pixels.
    ``shutter_delay_param`` is accepted and ignored, as in the reference."""
    _capi.require_gpu()
    X = _capi.f64(obj_pts).reshape(-1, 3)
    uv = np.empty((X.shape[0], 2))
    cam = _capi.pack_cam(k, d, r, t)
    _capi.check(_capi.lib().acino_project(_capi.ptr(X), X.shape[0], _capi.ptr(cam), _capi.ptr(uv)))
    return uv


def project_points_fisheye_all(obj_pts, k_arr, d_arr, r_arr, t_arr, cam_res=None):
    """project_points_fisheye for every camera of the rig in ONE launch (the per-camera loop of save_3d_cheetah_as_2d,
    src/lib/utils.py:255-263). obj_pts (..., 3) -> (n_cams, N, 2). With cam_res = (width, height) the reference's range rule is
    applied on the device: a pixel outside [0, width] x [0, height] becomes (NaN, NaN) (utils.py:262-263)."""
    _capi.require_gpu()
    X = _capi.f64(obj_pts).reshape(-1, 3)
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    n_cams = cams.shape[0]
    uv = np.empty((n_cams, X.shape[0], 2))
    w, h = (float(cam_res[0]), float(cam_res[1])) if cam_res is not None else (0.0, 0.0)
    _capi.check(_capi.lib().acino_project_all(_capi.ptr(X), X.shape[0], _capi.ptr(cams), n_cams, w, h, _capi.ptr(uv)))
    return uv


def triangulate_points_fisheye(img_pts_1, img_pts_2, k1, d1, r1, t1, k2, d2, r2, t2):
    """src/lib/calib.py:120-129. Pixels (M,1,2) / (M,2) / (2,) per view -> (M,3) float64."""
    _capi.require_gpu()
    p1 = _capi.f64(img_pts_1).reshape(-1, 2)
    p2 = _capi.f64(img_pts_2).reshape(-1, 2)
    if p1.shape[0] != p2.shape[0]:
        raise ValueError('img_pts_1 and img_pts_2 must hold the same number of points')
    xyz = np.empty((p1.shape[0], 3))
    ca = _capi.pack_cam(k1, d1, r1, t1)
    cb = _capi.pack_cam(k2, d2, r2, t2)
    _capi.check(_capi.lib().acino_triangulate_pair(_capi.ptr(p1), _capi.ptr(p2), p1.shape[0], _capi.ptr(ca), _capi.ptr(cb),
                                                   _capi.ptr(xyz)))
    return xyz


def calibrate_pair_extrinsics_fisheye(*args, **kwargs):
    """Selector token of the reference (src/lib/calib.py:257: `calib_func == calibrate_pair_extrinsics_fisheye` picks the
    fisheye model). The one-off OpenCV stereo calibration itself is outside the hot path."""
    raise NotImplementedError('pairwise OpenCV stereo calibration is outside this package (SURVEY.md §8: out of scope)')


def _is_fisheye_calib(func):
    return callable(func) and getattr(func, '__name__', '') == 'calibrate_pair_extrinsics_fisheye'


def manual_points_residuals(img_pts_arr, cam_idxs_to_correct, cam_pairs, k_arr, d_arr, r_arr, t_arr, params, loss=(3.0, 10.0, 20.0)):
    """Residual vector(s) of adjust_extrinsics_manual_points (src/lib/calib.py:224-256) for one or several 6-vectors
    (rvec, t) in ONE launch: triangulation of every camera pair (K2), reprojection into every camera (K1) and the
    redescending loss (src/lib/misc.py:329-343) of every error component. params (6,) -> (M,), params (B, 6) -> (B, M),
    M = n_cams * n_pairs * n_points * 2 in the reference's order."""
    _capi.require_gpu()
    pts = _capi.f64(img_pts_arr)
    n_points, n_cams = pts.shape[0], pts.shape[1]
    prm = _capi.f64(params)
    single = prm.ndim == 1
    prm = np.ascontiguousarray(prm.reshape(-1, 6))
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    corr = np.ascontiguousarray(cam_idxs_to_correct, dtype=np.int32).reshape(-1)
    pairs = np.ascontiguousarray(cam_pairs, dtype=np.int32).reshape(-1, 2)
    f = np.empty((prm.shape[0], n_cams * len(pairs) * n_points * 2))
    _capi.check(_capi.lib().acino_manual_residuals(_capi.ptr(pts), n_points, _capi.ptr(cams), n_cams, _capi.ptr(corr), len(corr),
                                                   _capi.ptr(pairs), len(pairs), _capi.ptr(prm), prm.shape[0],
                                                   float(loss[0]), float(loss[1]), float(loss[2]), _capi.ptr(f)))
    return f[0] if single else f


def adjust_extrinsics_manual_points(calib_func, img_pts_arr, cam_idxs_to_correct, k_arr, d_arr, r_arr, t_arr):
    """src/lib/calib.py:216-297, same arguments, prints and return value (r_arr, t_arr corrected in place for the cameras
    listed). The optimiser is the reference's: scipy least_squares with its defaults (trf, linear loss, forward
    differences); what changes is that one residual evaluation is one kernel launch and the six forward differences of the
    Jacobian are one more, with scipy's own step rule (sqrt(eps) * sign(x) * max(1, |x|)), instead of
    (2 + n_cams) x n_pairs OpenCV calls and a Python loop over every error per evaluation."""
    from scipy.optimize import least_squares
    from .sba import rotation_matrix_to_rvec
    if not _is_fisheye_calib(calib_func):
        raise NotImplementedError('only the fisheye camera model is implemented')
    calib_type = 'fisheye'
    if type(cam_idxs_to_correct) is int:
        cam_idxs_to_correct = [cam_idxs_to_correct]
        msg = f'Cam with index {cam_idxs_to_correct} is to have its pose corrected.'
    else:
        msg = f'Cams with indices {cam_idxs_to_correct} are to have their poses corrected'
    n_cams = len(k_arr)
    img_pts_arr = np.asarray(img_pts_arr)
    assert n_cams == img_pts_arr.shape[1], 'Number of cams in intrinsic file differs from number of cams in manual points file'
    cam_pairs = []
    for i in cam_idxs_to_correct:
        cam_pairs.append(sorted([(i - 1) % n_cams, i]))
        cam_pairs.append(sorted([i, (i + 1) % n_cams]))
    cam_pairs = np.unique(cam_pairs, axis=0).tolist()
    print(msg, '\nMinimizing error for', calib_type, 'cam pairs with indices', cam_pairs, '...\n')

    def residual_arr(R_t):
        return manual_points_residuals(img_pts_arr, cam_idxs_to_correct, cam_pairs, k_arr, d_arr, r_arr, t_arr, R_t)

    def jacobian(R_t):
        # scipy's 2-point scheme (_numdiff.approx_derivative): h = sqrt(eps) * sign(x) * max(1, |x|), dx = (x + h) - x
        x0 = np.asarray(R_t, dtype=np.float64)
        h = np.sqrt(np.finfo(np.float64).eps) * ((x0 >= 0).astype(float) * 2 - 1) * np.maximum(1.0, np.abs(x0))
        batch = np.tile(x0, (7, 1))
        for i in range(6):
            batch[i + 1, i] += h[i]
        f = manual_points_residuals(img_pts_arr, cam_idxs_to_correct, cam_pairs, k_arr, d_arr, r_arr, t_arr, batch)
        dx = batch[np.arange(1, 7), np.arange(6)] - x0
        return ((f[1:] - f[0]) / dx[:, None]).T

    R0_t0 = np.zeros(6)                      # cv.Rodrigues(identity) = 0, t0 = 0
    res = least_squares(residual_arr, R0_t0, jac=jacobian)
    print(res.message, f'success: {res.success}', f'func evals: {res.nfev}', f'cost: {res.cost}\n', sep='\n')
    R = rodrigues(res.x[0:3])
    t = res.x[3:6].reshape((3, 1))
    for cam_idx in cam_idxs_to_correct:
        t_arr[cam_idx] -= r_arr[cam_idx] @ t
        r_arr[cam_idx] = r_arr[cam_idx] @ R.T
    adjust_extrinsics_manual_points.last_result = res
    return r_arr, t_arr


def rodrigues(rvec):
    """cv::Rodrigues(vector -> matrix) on the device routine the solver uses (acino_rodrigues)."""
    rv = _capi.f64(rvec).reshape(-1, 3)
    R = np.empty((rv.shape[0], 3, 3))
    _capi.check(_capi.lib().acino_rodrigues(_capi.ptr(rv), rv.shape[0], _capi.ptr(R)))
    return R[0] if np.ndim(rvec) == 1 or np.shape(rvec) in ((3, 1), (1, 3)) else R
