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
    present = np.zeros(n_points * n_cams, dtype=bool)
    present[slot] = True
    if int(present.sum()) != slot.size:
        raise NotImplementedError('a camera observes a point more than once; the dense slot layout needs unique (point, camera) pairs')
    uv = np.zeros((n_points * n_cams, 2))
    uv[slot] = obs
    mask = (present.reshape(n_points, n_cams) * (1 << np.arange(n_cams, dtype=np.uint16))).sum(axis=1).astype(np.uint8)
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


def jacobian_points_extrinsics(points_3d, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr):
    """The analytic Jacobian behind bundle_adjust_points_and_extrinsics (the reference differentiates
    cost_func_points_extrinsics, src/lib/sba.py:142-146, numerically): (n_obs, 2, 9) = d(u, v)/d(w, tau, X) per observation,
    (w, tau) the world-frame twist of the observing camera (R <- R Exp(w), t <- t + R tau), X the observed point."""
    _capi.require_gpu()
    X = _capi.f64(points_3d).reshape(-1, 3)
    pidx = np.ascontiguousarray(point_3d_indices, dtype=np.int32).ravel()
    cidx = np.ascontiguousarray(camera_indices, dtype=np.int32).ravel()
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    J = np.empty((len(pidx), 2, 9))
    _capi.check(_capi.lib().acino_jacobian_list(_capi.ptr(X), len(X), _capi.ptr(pidx), _capi.ptr(cidx), len(pidx),
                                                _capi.ptr(cams), len(cams), _capi.ptr(J)))
    return J


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

_TERMINATION = {1.0: '`gtol` termination condition is satisfied.', 2.0: '`ftol` termination condition is satisfied.',
                3.0: '`xtol` termination condition is satisfied.', 4.0: 'The damping is exhausted: no descent left at double precision.',
                0.0: 'The maximum number of iterations is exceeded.'}


def _print_iteration_table(hist, st):
    """What scipy.optimize.least_squares prints with verbose=2 (src/lib/sba.py:170,189; core/sba.py tees it into sba.log):
    one row per LM iteration in scipy's column layout, then its two closing lines. `Step norm` is max |step| (the
    quantity the xtol rule of this solver tests), `Total nfev` the linearisations so far."""
    print('{0:^15}{1:^15}{2:^15}{3:^15}{4:^15}{5:^15}'.format('Iteration', 'Total nfev', 'Cost', 'Cost reduction', 'Step norm', 'Optimality'))
    for i, row in enumerate(hist):
        if i == 0:
            print('{0:^15}{1:^15}{2:^15.4e}{3:^15}{4:^15}{5:^15.2e}'.format(0, 1, row[1], '', '', row[4]))
        else:
            print('{0:^15}{1:^15}{2:^15.4e}{3:^15.2e}{4:^15.2e}{5:^15.2e}'.format(int(row[0]), int(row[0]) + 1, row[1], row[2], row[3], row[4]))
    print(_TERMINATION.get(st.get('status'), 'The solver stopped.'))
    print('Function evaluations {0}, initial cost {1:.4e}, final cost {2:.4e}, first-order optimality {3:.2e}.'.format(
        st['iterations'] + 1, st['cost_initial'], st['cost_final'], st.get('gmax', float('nan'))))


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
                                        project_func, verbose=2, **solver_opts):
    """src/lib/sba.py:158-178. The reference's parameters are [rvecs | tvecs | points] with scipy's additive update; the
    solver here steps every camera by a world-frame twist (R <- R Exp(w), t <- t + R tau: the same set of (R, t), the same
    minimiser) and converts back. Cauchy loss with scipy's default f_scale = 1.0. verbose=2 (the reference's literal)
    prints scipy's iteration table. Returns (obj_pts, r_arr (n,3,3), t_arr (n,3,1), residuals)."""
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
        if verbose and verbose >= 2:
            _print_iteration_table(solver.history(), st)
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


# ========== SBA PREPARATION (host side, runs once per calibration; src/lib/sba.py:37-137) ==========

def prepare_calib_board_data_for_bundle_adjustment(img_pts_arr, fnames_arr, board_shape, k_arr, d_arr, r_arr, t_arr,
                                                   triangulate_func):
    """src/lib/sba.py:37-93. Images seen by >= 2 cameras, in first-seen order over the cameras' file lists (the reference
    iterates a dict built from a set, whose order is arbitrary; first-seen order is one of its possible orders and is
    deterministic). Every kept image contributes board_shape[0]*board_shape[1] new point ids and one observation per
    seeing camera and corner; the initial 3-D estimate comes from the first two seeing cameras. Outputs are float32 /
    int arrays like the reference's (sba.py:88-91)."""
    if not _is_fisheye_triangulate(triangulate_func):
        raise NotImplementedError('only triangulate_points_fisheye is supported (CUDA path, no fallback)')
    n_cam = len(img_pts_arr)
    seen = {}
    for cam_idx in range(n_cam):
        for name in fnames_arr[cam_idx]:
            seen[name] = seen.get(name, 0) + 1
    keep = [name for name, v in seen.items() if v >= 2]
    index_of = [{name: i for i, name in enumerate(fnames_arr[c])} for c in range(n_cam)]
    per_img = board_shape[0] * board_shape[1]
    points_2d, point_3d_indices, camera_indices = [], [], []
    tri_a, tri_b, tri_cams = [], [], []
    for n_img, name in enumerate(keep):
        cams = [c for c in range(n_cam) if name in index_of[c]]
        ids = np.arange(n_img * per_img, (n_img + 1) * per_img)
        for c in cams:
            points_2d.append(np.asarray(img_pts_arr[c][index_of[c][name]]).reshape(per_img, 2))
            point_3d_indices.append(ids)
            camera_indices.append(np.full(per_img, c))
        a, b = cams[0], cams[1]
        tri_a.append(np.asarray(img_pts_arr[a][index_of[a][name]]).reshape(per_img, 2))
        tri_b.append(np.asarray(img_pts_arr[b][index_of[b][name]]).reshape(per_img, 2))
        tri_cams.append((a, b))
    if not keep:
        return (np.zeros((0, 2), np.float32), np.zeros((0, 3), np.float32), np.zeros(0, dtype=int), np.zeros(0, dtype=int))
    # one CUDA launch per distinct camera pair instead of one per image
    points_3d = np.empty((len(keep) * per_img, 3))
    pairs = sorted(set(tri_cams))
    for a, b in pairs:
        sel = [i for i, ab in enumerate(tri_cams) if ab == (a, b)]
        pa = np.concatenate([tri_a[i] for i in sel]); pb = np.concatenate([tri_b[i] for i in sel])
        X = triangulate_func(pa, pb, k_arr[a], d_arr[a], r_arr[a], t_arr[a], k_arr[b], d_arr[b], r_arr[b], t_arr[b])
        for j, i in enumerate(sel):
            points_3d[i * per_img:(i + 1) * per_img] = X[j * per_img:(j + 1) * per_img]
    return (np.concatenate(points_2d).astype(np.float32), points_3d.astype(np.float32),
            np.concatenate(point_3d_indices).astype(int), np.concatenate(camera_indices).astype(int))


def prepare_manual_points_for_bundle_adjustment(img_pts_arr, k_arr, d_arr, r_arr, t_arr, triangulate_func):
    """src/lib/sba.py:96-137. img_pts_arr (n_points, n_cams, 2) with NaN where a camera did not see the point; points
    seen by >= 2 cameras are kept and initialised from their first two cameras."""
    if not _is_fisheye_triangulate(triangulate_func):
        raise NotImplementedError('only triangulate_points_fisheye is supported (CUDA path, no fallback)')
    pts = np.asarray(img_pts_arr, dtype=np.float64)
    n_pts, n_cam = pts.shape[0], pts.shape[1]
    points_2d, point_3d_indices, camera_indices, first_two = [], [], [], []
    idx = 0
    for i in range(n_pts):
        cams = [c for c in range(n_cam) if not np.isnan(pts[i, c]).any()]
        if len(cams) < 2:
            continue
        for c in cams:
            points_2d.append(pts[i, c]); camera_indices.append(c); point_3d_indices.append(idx)
        first_two.append((i, cams[0], cams[1]))
        idx += 1
    # one CUDA launch per distinct camera pair instead of one per point
    points_3d = np.empty((idx, 3))
    for a, b in sorted({(a, b) for _, a, b in first_two}):
        sel = [j for j, (_, a_, b_) in enumerate(first_two) if (a_, b_) == (a, b)]
        src = [first_two[j][0] for j in sel]
        points_3d[sel] = triangulate_func(pts[src, a], pts[src, b], k_arr[a], d_arr[a], r_arr[a], t_arr[a],
                                          k_arr[b], d_arr[b], r_arr[b], t_arr[b])
    return (np.array(points_2d, dtype=np.float32).reshape(-1, 2), points_3d.astype(np.float32).reshape(-1, 3),
            np.array(point_3d_indices, dtype=int), np.array(camera_indices, dtype=int))


def bundle_adjust_board_points_only(img_pts_arr, fnames_arr, board_shape, k_arr, d_arr, r_arr, t_arr, triangulate_func,
                                    project_func):
    """src/lib/sba.py:198-206."""
    points_2d, points_3d, point_3d_indices, camera_indices = prepare_calib_board_data_for_bundle_adjustment(
        img_pts_arr, fnames_arr, board_shape, k_arr, d_arr, r_arr, t_arr, triangulate_func)
    return bundle_adjust_points_only(points_2d, points_3d, point_3d_indices, camera_indices, k_arr, d_arr, r_arr, t_arr,
                                     project_func)


# ==========  BACK-END FUNCTION MANAGERS (src/lib/sba.py:211-313) ==========

def merge_board_and_manual_points(board, manual, reference_ids=False):
    """The merge of src/lib/sba.py:254-258: board = (points_2d, points_3d, point_3d_indices, camera_indices) of the
    checkerboard corners, manual = the same of the manually labelled points; returns the four merged arrays.

    reference_ids=True reproduces the reference bit for bit: the manual ids are offset by ``point_3d_indices.max()``
    (sba.py:257), so the first manual point shares its id - i.e. its 3-D parameter - with the last board corner and
    the last manual row of points_3d is referenced by nothing. reference_ids=False (the default of this package) offsets
    by ``max() + 1``: no two physical points are merged. With no manual points both agree."""
    p2d, p3d, pidx, pcam = board
    m2d, m3d, midx, mcam = manual
    offset = (pidx.max() + (0 if reference_ids else 1)) if len(pidx) else 0
    return (np.append(p2d, m2d, axis=0), np.append(p3d, m3d, axis=0), np.append(pidx, midx + offset, axis=0),
            np.append(pcam, mcam, axis=0))


def _sba_board_points(scene_fpath, points_fpaths, manual_points_fpath, out_fpath, triangulate_func, project_func,
                      camera_indices=None, manual_points_only=False, reference_ids=False):
    """src/lib/sba.py:211-282: extrinsic calibration refinement from checkerboard (and optional manual) points;
    writes the refined scene JSON and returns the residual dict. ``reference_ids``: see merge_board_and_manual_points."""
    from .utils import load_points, load_manual_points
    img_pts_arr, fnames_arr, board_shape = [], [], None
    if camera_indices is None:
        camera_indices = range(len(points_fpaths))
    for i in camera_indices:
        points, fnames, board_shape, *_ = load_points(points_fpaths[i])
        img_pts_arr.append(points)
        fnames_arr.append(fnames)
    k_arr, d_arr, r_arr, t_arr, cam_res = load_scene(scene_fpath)
    assert len(k_arr) == len(img_pts_arr)
    if manual_points_fpath is not None:
        manual_points, _manual_fnames, *_ = load_manual_points(manual_points_fpath)
        manual = prepare_manual_points_for_bundle_adjustment(manual_points, k_arr, d_arr, r_arr, t_arr, triangulate_func)
        if manual_points_only:
            print('bundle_adjust_board_points_and_extrinsics (with manual points only)')
            obj_pts, r_arr, t_arr, res = bundle_adjust_points_and_extrinsics(*manual, k_arr, d_arr, r_arr, t_arr, project_func)
        else:
            print('bundle_adjust_board_points_and_extrinsics (with manual points)')
            board = prepare_calib_board_data_for_bundle_adjustment(img_pts_arr, fnames_arr, board_shape, k_arr, d_arr, r_arr,
                                                                   t_arr, triangulate_func)
            obj_pts, r_arr, t_arr, res = bundle_adjust_points_and_extrinsics(
                *merge_board_and_manual_points(board, manual, reference_ids), k_arr, d_arr, r_arr, t_arr, project_func)
    else:
        print('bundle_adjust_board_points_and_extrinsics')
        p2d, p3d, pidx, pcam = prepare_calib_board_data_for_bundle_adjustment(img_pts_arr, fnames_arr, board_shape, k_arr,
                                                                              d_arr, r_arr, t_arr, triangulate_func)
        obj_pts, r_arr, t_arr, res = bundle_adjust_points_and_extrinsics(p2d, p3d, pidx, pcam, k_arr, d_arr, r_arr, t_arr,
                                                                         project_func)
    print(f"\nBefore: mean: {np.mean(res['before'])}, std: {np.std(res['before'])}")
    print(f"After: mean: {np.mean(res['after'])}, std: {np.std(res['after'])}\n")
    save_scene(out_fpath, k_arr, d_arr, r_arr, t_arr, cam_res)
    return res


def _sba_points(scene_fpath, points_2d_df, triangulate_func, project_func):
    """src/lib/sba.py:285-313: scene + thresholded DeepLabCut frame -> (DataFrame[frame, marker, x, y, z, point_index],
    residuals). TRI initialises, the index arrays come from the same inner merge as the reference's (sba.py:296-302),
    the points-only solve runs with f_scale=50."""
    k_arr, d_arr, r_arr, t_arr, _cam_res = load_scene(scene_fpath)
    if not _is_fisheye_triangulate(triangulate_func):
        raise NotImplementedError('only triangulate_points_fisheye is supported (CUDA path, no fallback)')
    from .utils import _pairwise_3d_with_index
    points_3d_df, row3d = _pairwise_3d_with_index(points_2d_df, k_arr, d_arr.reshape((-1, 4)), r_arr, t_arr)
    points_3d_df['point_index'] = points_3d_df.index
    points_3d = np.array(points_3d_df[['x', 'y', 'z']], dtype=np.float64)
    # The reference builds the index arrays with an inner merge of the 2-D frame with the 3-D frame on (frame, marker)
    # (sba.py:296-302): the rows of the 2-D frame whose point was triangulated, in the 2-D frame's order, each with the
    # row number of its point. The lowering above already knows that row for every 2-D row: same arrays, no merge.
    keep = row3d >= 0
    points_2d = np.stack([points_2d_df['x'].to_numpy(dtype=np.float64)[keep], points_2d_df['y'].to_numpy(dtype=np.float64)[keep]], axis=1)
    point_indices = row3d[keep]
    camera_indices = points_2d_df['camera'].to_numpy()[keep]
    print('bundle_adjust_points_only')
    pts_3d, res = bundle_adjust_points_only(points_2d, points_3d, point_indices, camera_indices, k_arr, d_arr, r_arr, t_arr,
                                            project_func, f_scale=50)
    print(f"\nBefore: mean: {np.mean(res['before'])}, std: {np.std(res['before'])}")
    print(f"After: mean: {np.mean(res['after'])}, std: {np.std(res['after'])}\n")
    new_points_3d_df = points_3d_df.copy()
    new_points_3d_df[['x', 'y', 'z']] = pts_3d
    return new_points_3d_df, res
