"""Drop-in for the reference's lib/sba.py (src/lib/sba.py) — same function names, argument order, array
layouts and return values; the residual / Jacobian / normal-equation / LM arithmetic runs in CUDA.

Differences that are deliberate: the optimiser is an analytic-Jacobian
Levenberg-Marquardt on the same Cauchy objective scipy minimises (not scipy's finite-difference TRF),
so iterates differ while residuals at equal parameters and the converged minimiser agree; the broken
sparsity pattern of create_bundle_adjustment_jacobian_sparsity_matrix is not needed and not reproduced.
"""
from time import time
import numpy as np
from . import _capi, dense
from .utils import load_scene, save_scene, get_pairwise_3d_points_from_df, _is_fisheye_triangulate


def _is_fisheye_project(func):
    return callable(func) and getattr(func, '__name__', '') == 'project_points_fisheye'


def _require_fisheye(project_func):
    if not _is_fisheye_project(project_func):
        raise NotImplementedError('only project_points_fisheye is supported (CUDA path, no fallback)')


def rotation_matrix_to_rvec(R):
    """cv::Rodrigues(matrix -> vector) as used at src/lib/sba.py:162 to seed the extrinsic parameters."""
    R = np.asarray(R, dtype=np.float64).reshape(3, 3)
    u, _, vt = np.linalg.svd(R)
    R = u @ vt
    rx, ry, rz = R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]
    s = np.sqrt((rx * rx + ry * ry + rz * rz) * 0.25)
    c = np.clip((R[0, 0] + R[1, 1] + R[2, 2] - 1.0) * 0.5, -1.0, 1.0)
    theta = np.arccos(c)
    if s < 1e-5:
        if c > 0:
            return np.zeros(3)
        t = (R[0, 0] + 1) * 0.5
        rx = np.sqrt(max(t, 0.0))
        t = (R[1, 1] + 1) * 0.5
        ry = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 1] < 0 else 1.0)
        t = (R[2, 2] + 1) * 0.5
        rz = np.sqrt(max(t, 0.0)) * (-1.0 if R[0, 2] < 0 else 1.0)
        if abs(rx) < abs(ry) and abs(rx) < abs(rz) and (R[1, 2] > 0) != (ry * rz > 0):
            rz = -rz
        v = np.array([rx, ry, rz])
        return v * (theta / np.linalg.norm(v))
    return np.array([rx, ry, rz]) * (theta / (2.0 * s))


def _lists_to_dense(points_2d, point_3d_indices, camera_indices, n_points, n_cams):
    pidx = np.asarray(point_3d_indices).astype(np.int64).ravel()
    cidx = np.asarray(camera_indices).astype(np.int64).ravel()
    obs = _capi.f64(points_2d).reshape(-1, 2)
    if not (len(pidx) == len(cidx) == len(obs)):
        raise ValueError('points_2d, point_3d_indices and camera_indices must have one entry per observation')
    if len(pidx) and (pidx.min() < 0 or pidx.max() >= n_points or cidx.min() < 0 or cidx.max() >= n_cams):
        raise ValueError('observation index out of range')
    if n_cams > _capi.MAX_CAMS:
        raise NotImplementedError(f'at most {_capi.MAX_CAMS} cameras per scene')
    slot = pidx * n_cams + cidx
    if np.unique(slot).size != slot.size:
        raise NotImplementedError('a camera observes a point more than once; the dense slot layout needs unique (point, camera) pairs')
    uv = np.zeros((n_points * n_cams, 2))
    uv[slot] = obs
    mask = np.zeros(n_points, dtype=np.uint8)
    np.bitwise_or.at(mask, pidx, (1 << cidx).astype(np.uint8))
    return uv.reshape(n_points, n_cams, 2), mask, slot


# ========== COST FUNCTIONS ==========

def cost_func_points_only(params, n_points, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr, points_2d, project_func):
    """src/lib/sba.py:149-153: [du0, dv0, du1, dv1, ...] = projected - observed."""
    _require_fisheye(project_func)
    _capi.require_gpu()
    X = _capi.f64(params).reshape(n_points, 3)
    pidx = np.ascontiguousarray(point_3d_indices, dtype=np.int32).ravel()
    cidx = np.ascontiguousarray(camera_indices, dtype=np.int32).ravel()
    obs = _capi.f64(points_2d).reshape(-1, 2)
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    f = np.empty(2 * len(pidx))
    _capi.check(_capi.lib().acino_residuals_list(_capi.ptr(X), n_points, _capi.ptr(pidx), _capi.ptr(cidx), _capi.ptr(obs),
                                                 len(pidx), _capi.ptr(cams), len(cams), _capi.ptr(f)))
    return f


def params_to_points_extrinsics(params, n_cams, n_points):
    """src/lib/sba.py:25-32; the Rodrigues conversion runs in the CUDA library."""
    params = _capi.f64(params)
    r_end, t_end = n_cams * 3, n_cams * 6
    r_vecs = np.ascontiguousarray(params[:r_end].reshape(n_cams, 3))
    r_arr = np.empty((n_cams, 3, 3))
    _capi.require_gpu()
    _capi.check(_capi.lib().acino_rodrigues(_capi.ptr(r_vecs), n_cams, _capi.ptr(r_arr)))
    t_arr = params[r_end:t_end].reshape((n_cams, 3, 1))
    obj_pts = params[t_end:].reshape((n_points, 3))
    return obj_pts, r_arr, t_arr


def cost_func_points_extrinsics(params, n_cams, n_points, point_3d_indices, camera_indices, k_arr, d_arr, points_2d, project_func):
    """src/lib/sba.py:142-146."""
    obj_pts, r_arr, t_arr = params_to_points_extrinsics(params, n_cams, n_points)
    return cost_func_points_only(obj_pts, n_points, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr,
                                 points_2d, project_func)


# ========== SBA ==========

def _print_stats(st):
    print(f"   LM iterations: {st['iterations']}, kernel launches: {st['launches']}, "
          f"cost {st['cost_initial']:.6e} -> {st['cost_final']:.6e}, unconverged: {st['n_unconverged']}, "
          f"device time {st['solve_ms']:.3f} ms")


def bundle_adjust_points_only(points_2d, points_3d, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr,
                              project_func, f_scale=50, **solver_opts):
    """src/lib/sba.py:181-195. Returns (obj_pts (P,3), {'before': f0, 'after': f}) with residual vectors
    in input observation order."""
    _require_fisheye(project_func)
    points_3d = _capi.f64(points_3d).reshape(-1, 3)
    n_points, n_cams = len(points_3d), len(k_arr)
    uv, mask, slot = _lists_to_dense(points_2d, point_3d_indices, camera_indices, n_points, n_cams)
    solver = dense.SbaSolver(n_cams, n_points)
    try:
        solver.set_cameras(k_arr, d_arr, r_arr, t_arr)
        solver.set_observations(uv, mask)
        solver.set_points(points_3d)
        f0 = solver.residuals().reshape(-1, 2)[slot].ravel()
        t0 = time()
        st = solver.solve(dense.default_opts(_capi.MODE_POINTS_ONLY, f_scale=float(f_scale), **solver_opts))
        t1 = time()
        _print_stats(st)
        print(f'\nOptimization took {t1-t0:.2f} seconds')
        obj_pts = solver.get_points()
        f1 = solver.residuals().reshape(-1, 2)[slot].ravel()
    finally:
        solver.close()
    bundle_adjust_points_only.last_stats = st
    return obj_pts, dict(before=f0, after=f1)


def bundle_adjust_points_and_extrinsics(points_2d, points_3d, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr,
                                        project_func, **solver_opts):
    """src/lib/sba.py:158-178. Parameters [rvecs | tvecs | points] (additive rvec update, as scipy's x + step),
    Cauchy loss with scipy's default f_scale = 1.0. Returns (obj_pts, r_arr (n,3,3), t_arr (n,3,1), residuals)."""
    _require_fisheye(project_func)
    points_3d = _capi.f64(points_3d).reshape(-1, 3)
    n_points, n_cams = len(points_3d), len(k_arr)
    uv, mask, slot = _lists_to_dense(points_2d, point_3d_indices, camera_indices, n_points, n_cams)
    r_vecs = np.array([rotation_matrix_to_rvec(r) for r in r_arr])
    t_vecs = _capi.f64(t_arr).reshape(n_cams, 3)
    solver = dense.SbaSolver(n_cams, n_points)
    try:
        solver.set_cameras(k_arr, d_arr, r_arr, t_arr)
        solver.set_rvec_t(r_vecs, t_vecs)
        solver.set_observations(uv, mask)
        solver.set_points(points_3d)
        f0 = solver.residuals().reshape(-1, 2)[slot].ravel()
        t0 = time()
        st = solver.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, **solver_opts))
        t1 = time()
        _print_stats(st)
        print(f'\nOptimization took {t1-t0:.2f} seconds')
        obj_pts = solver.get_points()
        cams = solver.get_cameras()
        f1 = solver.residuals().reshape(-1, 2)[slot].ravel()
    finally:
        solver.close()
    r_out = cams[:, 9:18].reshape(n_cams, 3, 3).copy()
    t_out = cams[:, 18:21].reshape(n_cams, 3, 1).copy()
    bundle_adjust_points_and_extrinsics.last_stats = st
    return obj_pts, r_out, t_out, dict(before=f0, after=f1)
