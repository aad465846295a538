"""The two facade entry points of the reference's lib/app.py that sit on the hot path (src/lib/app.py:123-138):
they bind the fisheye callables and hand over to the managers in sba.py. core.sba calls sba_points_fisheye
(src/core/sba.py:43); the calibration notebook calls sba_board_points_fisheye (calib_with_gui.ipynb cell 30)."""
import numpy as np
from . import _capi, utils
from .calib import triangulate_points_fisheye, project_points_fisheye
from .sba import _sba_board_points, _sba_points


def sba_board_points_fisheye(scene_fpath, points_fpaths, out_fpath, manual_points_fpath=None, manual_points_only=False,
                             camera_indices=None):
    """src/lib/app.py:123-126."""
    return _sba_board_points(scene_fpath, points_fpaths, manual_points_fpath, out_fpath, triangulate_points_fisheye,
                             project_points_fisheye, camera_indices, manual_points_only)


def sba_points_fisheye(scene_fpath, points_2d_df):
    """src/lib/app.py:135-138."""
    return _sba_points(scene_fpath, points_2d_df, triangulate_points_fisheye, project_points_fisheye)


def ring_critical_likelihood(lik):
    """lik (P, n_cams) -> (P,): the supremum of thresholds t at which `likelihood > t` still leaves a ring pair
    (c, c+1 mod n) of cameras for the point, i.e. the point is triangulated iff result > t (kernel ring_critical)."""
    _capi.require_gpu()
    lik = _capi.f64(lik)
    out = np.empty(lik.shape[0])
    _capi.check(_capi.lib().acino_ring_critical(_capi.ptr(lik), lik.shape[0], lik.shape[1], _capi.ptr(out)))
    return out


def EOM_curve_fit(pts_3d, frames=None, fit_order=3):
    """src/lib/points.py:94-132 (host logic, unchanged in substance): per axis, polynomial equation-of-motion fits of
    increasing order with scipy curve_fit (trf, Cauchy loss), each initialised by the previous one. Returns (fit, fit_deriv).
    The polynomial is evaluated directly instead of through sympy.lambdify. Like the reference, `frames` is needed."""
    from scipy.optimize import curve_fit
    assert 0 < fit_order < 19 and isinstance(fit_order, int), 'fit_order must be an integer from 1 to 18'
    num_axes = pts_3d.shape[1]

    def fit_func(t, *p):
        out = p[0]
        for order in range(1, len(p)):
            out = out + p[order] * t ** order
        return out

    def fit_func_deriv(t, *p):                 # p = (b, c, ...): derivative b + 2 c t + 3 d t^2 ...
        out = p[0] + 0 * t
        for order in range(2, len(p) + 1):
            out = out + order * p[order - 1] * t ** (order - 1)
        return out

    func_params = np.zeros((num_axes, 1))
    for order in range(1, fit_order + 1):
        func_params = np.append(func_params, np.zeros((num_axes, 1)), axis=1)
        for ax in range(num_axes):
            func_params[ax, :], _ = curve_fit(fit_func, frames, pts_3d[:, ax], p0=func_params[ax, :], method='trf', loss='cauchy')
    fit, fit_deriv = np.full(pts_3d.shape, np.nan), np.full(pts_3d.shape, np.nan)
    for ax in range(num_axes):
        fit[:, ax] = fit_func(frames, *func_params[ax, :])
        fit_deriv[:, ax] = fit_func_deriv(frames, *func_params[ax, 1:])
    return fit, fit_deriv


def initialize_marker_3d(pts_2d_df, marker, k_arr, d_arr, r_arr, t_arr, dlc_thresh_step=0.01, plot=False, **kwargs):
    """src/lib/app.py:39-84, same arguments and return value (fit, fit_deriv). The reference finds the highest usable
    likelihood threshold by re-triangulating the whole table at every 0.01 step; here one kernel gives every frame's
    critical likelihood for the marker (max over ring pairs of the smaller likelihood) and the same threshold sequence
    (same float additions, same strict comparison) is walked over that array. One triangulation at the final threshold."""
    if plot:
        raise NotImplementedError('plotting is outside this package')
    n_cams = len(k_arr)
    dlc_thresh = dlc_thresh_step
    frames = pts_2d_df['frame'].unique()
    tot_frames = num_frames = len(frames)
    # the marker's likelihoods as [frame][camera]; -1 where the table has no row (never above a threshold)
    order = np.argsort(frames, kind='stable')
    sorted_frames = frames[order]
    rows = pts_2d_df[pts_2d_df['marker'] == marker]
    lik = np.full((tot_frames, n_cams), -1.0)
    if len(rows):
        fi = np.searchsorted(sorted_frames, rows['frame'].to_numpy())
        cam = rows['camera'].to_numpy().astype(np.int64)
        lik[fi, cam] = rows['likelihood'].to_numpy(dtype=np.float64)
    crit = ring_critical_likelihood(lik) if tot_frames else np.zeros(0)
    while num_frames == tot_frames:
        num_frames = int(np.count_nonzero(crit > dlc_thresh))      # frames in which the marker is triangulated at this threshold
        dlc_thresh += dlc_thresh_step
    dlc_thresh -= 2 * dlc_thresh_step
    print(f"Initializing {marker}'s 3D points using an interim dlc_thresh of {dlc_thresh:.2f}")
    pts_3d_df = utils.get_pairwise_3d_points_from_df(pts_2d_df[pts_2d_df['likelihood'] > dlc_thresh], k_arr, d_arr, r_arr, t_arr,
                                                     triangulate_points_fisheye)
    pts_3d = np.full((tot_frames, 3), np.nan)
    for frame, *pt_3d in pts_3d_df[pts_3d_df['marker'] == marker][['frame', 'x', 'y', 'z']].values:
        pts_3d[int(frame) - frames[0], :] = pt_3d
    initialize_marker_3d.last_dlc_thresh = dlc_thresh
    return EOM_curve_fit(pts_3d, frames=frames, **kwargs)


def install(lib_calib=None, lib_utils=None, lib_sba=None, lib_app=None, lib_metric=None):
    """Drop-in integration: rebinds the hot functions of already-imported reference modules
    (lib.calib, lib.utils, lib.sba, lib.app, lib.metric) to this package's CUDA-backed ones; reference files are not
    edited. lib.app binds `_sba_points` / `_sba_board_points` by name at import (src/lib/app.py:14), so pass it too when
    it is already imported. Callables the reference took with `from lib.calib import ...` before this call keep working:
    they are only used as selector tokens (their `__name__` is checked), never called."""
    from . import calib, utils, sba, metric
    if lib_calib is not None:
        lib_calib.project_points_fisheye = calib.project_points_fisheye
        lib_calib.triangulate_points_fisheye = calib.triangulate_points_fisheye
        lib_calib.adjust_extrinsics_manual_points = calib.adjust_extrinsics_manual_points
    if lib_utils is not None:
        lib_utils.get_pairwise_3d_points_from_df = utils.get_pairwise_3d_points_from_df
        lib_utils.save_3d_cheetah_as_2d = utils.save_3d_cheetah_as_2d
        lib_utils.load_dlc_points_as_df = utils.load_dlc_points_as_df
    if lib_sba is not None:
        for name in ('cost_func_points_only', 'cost_func_points_extrinsics', 'params_to_points_extrinsics',
                     'bundle_adjust_points_only', 'bundle_adjust_points_and_extrinsics', 'bundle_adjust_board_points_only',
                     'prepare_calib_board_data_for_bundle_adjustment', 'prepare_manual_points_for_bundle_adjustment',
                     '_sba_board_points', '_sba_points'):
            setattr(lib_sba, name, getattr(sba, name))
        lib_sba.get_pairwise_3d_points_from_df = utils.get_pairwise_3d_points_from_df
    if lib_app is not None:
        lib_app._sba_points = sba._sba_points
        lib_app._sba_board_points = sba._sba_board_points
        lib_app.project_points_fisheye = calib.project_points_fisheye
        lib_app.triangulate_points_fisheye = calib.triangulate_points_fisheye
        lib_app.initialize_marker_3d = initialize_marker_3d
    if lib_metric is not None:
        lib_metric.residual_error = metric.residual_error
