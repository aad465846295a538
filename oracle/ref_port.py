"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) — CPU port of the reference's TRI + SBA *algorithm*,
built on the same third-party numerics the reference executes (OpenCV + scipy, un-vendored, whatever the image
has). This is the timed CPU baseline of bench.py (`cpu_baseline.kind = "port"`, `--impl reference`)
and the converged-solution checker of the tests at sizes where scipy finishes in seconds. /root/reference does not
exist on the GPU box, so this restatement is what travels; tests/test_oracle_golden.py pins it against fixtures
produced by the real reference in the build container.

What is restated, and where it lives in the reference:
  * one cv2.fisheye.projectPoints call per observation from a Python loop      src/lib/sba.py:149-153 -> calib.py:132-136
  * the same loop with cameras unpacked from the parameter vector              src/lib/sba.py:25-32,142-146
  * scipy least_squares(trf, cauchy, x_scale='jac', 2-point FD on a sparsity pattern), literals of
    src/lib/sba.py:170-173 (extrinsics: ftol 1e-10, max_nfev 1000, f_scale 1.0 default) and
    src/lib/sba.py:189-190 (points only: ftol 1e-15, max_nfev 500, f_scale 50)
  * the 0/1 sparsity pattern                                                   src/lib/sba.py:11-22
    (as shipped, and the layout-consistent variant of SURVEY.md F6 — selectable, labelled)
  * ring-pair two-view triangulation + unweighted mean over pairs             src/lib/utils.py:319-349, calib.py:120-129
"""
import time
import numpy as np


def _project_one_cv(X, k, d, R, t):
    """calib.py:132-136 on a single point: matrix -> rvec -> cv2.fisheye.projectPoints."""
    import cv2
    rvec = cv2.Rodrigues(R)[0]
    return cv2.fisheye.projectPoints(np.asarray(X, np.float64).reshape(1, 1, 3), rvec, t, k, d)[0].reshape(2)


class CostCounter:
    """Residual function in the reference's per-observation form; counts sweeps."""

    def __init__(self, pidx, cidx, obs, k_arr, d_arr, r_arr=None, t_arr=None, n_cams=None):
        self.pidx = np.asarray(pidx).astype(np.int64)
        self.cidx = np.asarray(cidx).astype(np.int64)
        self.obs = np.asarray(obs, np.float64).reshape(-1, 2)
        self.k, self.d, self.r, self.t = k_arr, d_arr, r_arr, t_arr
        self.n_cams = n_cams if n_cams is not None else len(k_arr)
        self.calls = 0

    def points_only(self, params):
        self.calls += 1
        X = params.reshape(-1, 3)
        proj = np.empty_like(self.obs)
        for o, (i, j) in enumerate(zip(self.pidx, self.cidx)):
            proj[o] = _project_one_cv(X[i], self.k[j], self.d[j], self.r[j], self.t[j])
        return (proj - self.obs).ravel()

    def points_extrinsics(self, params):
        import cv2
        self.calls += 1
        n = self.n_cams
        R = [cv2.Rodrigues(params[3 * j:3 * j + 3].copy())[0] for j in range(n)]
        T = params[3 * n:6 * n].reshape(n, 3, 1)
        X = params[6 * n:].reshape(-1, 3)
        proj = np.empty_like(self.obs)
        for o, (i, j) in enumerate(zip(self.pidx, self.cidx)):
            proj[o] = _project_one_cv(X[i], self.k[j], self.d[j], R[j], T[j])
        return (proj - self.obs).ravel()


def sparsity_pattern(n_cams, n_params_per_camera, cidx, n_points, pidx, layout='shipped'):
    """src/lib/sba.py:11-22. layout='shipped': camera columns at cam*npc+s (the reference as written, which for the
    [rvecs | tvecs | points] vector marks the wrong columns); layout='consistent': rvec columns 3j+s and
    tvec columns 3n+3j+s, the pattern that matches the parameter layout."""
    from scipy.sparse import coo_matrix
    cidx = np.asarray(cidx).astype(np.int64)
    pidx = np.asarray(pidx).astype(np.int64)
    n_obs = cidx.size
    rows, cols = [], []
    o = np.arange(n_obs)
    cam_cols = []
    if n_params_per_camera:
        if layout == 'shipped':
            cam_cols = [cidx * n_params_per_camera + s for s in range(n_params_per_camera)]
        else:
            cam_cols = [3 * cidx + s for s in range(3)] + [3 * n_cams + 3 * cidx + s for s in range(3)]
    pt_cols = [n_cams * n_params_per_camera + 3 * pidx + s for s in range(3)]
    for c in cam_cols + pt_cols:
        for r in (2 * o, 2 * o + 1):
            rows.append(r); cols.append(c)
    rows = np.concatenate(rows); cols = np.concatenate(cols)
    A = coo_matrix((np.ones(rows.size, dtype=np.int64), (rows, cols)),
                   shape=(2 * n_obs, n_cams * n_params_per_camera + 3 * n_points))
    A.sum_duplicates()
    A.data[:] = 1
    return A.tolil()


def bundle_adjust_points_only(points_2d, points_3d, pidx, cidx, k_arr, d_arr, r_arr, t_arr, f_scale=50, verbose=0):
    """src/lib/sba.py:181-195."""
    from scipy.optimize import least_squares
    points_3d = np.asarray(points_3d, np.float64)
    n_points = len(points_3d)
    fn = CostCounter(pidx, cidx, points_2d, k_arr, d_arr, r_arr, t_arr)
    x0 = points_3d.ravel().copy()
    f0 = fn.points_only(x0)
    A = sparsity_pattern(len(k_arr), 0, cidx, n_points, pidx)
    t0 = time.perf_counter()
    res = least_squares(fn.points_only, x0, jac_sparsity=A, verbose=verbose, x_scale='jac', ftol=1e-15, method='trf',
                        loss='cauchy', f_scale=f_scale, max_nfev=500)
    dt = time.perf_counter() - t0
    info = dict(seconds=dt, nfev=int(res.nfev), njev=int(res.njev), cost_calls=fn.calls, status=int(res.status), cost=float(res.cost),
                optimality=float(res.optimality))
    return res.x.reshape(n_points, 3), dict(before=f0, after=res.fun), info


def bundle_adjust_points_and_extrinsics(points_2d, points_3d, pidx, cidx, k_arr, d_arr, r_arr, t_arr, layout='shipped',
                                        verbose=0, max_nfev=1000):
    """src/lib/sba.py:158-178."""
    import cv2
    from scipy.optimize import least_squares
    points_3d = np.asarray(points_3d, np.float64)
    n_points, n_cams = len(points_3d), len(k_arr)
    rv = np.array([cv2.Rodrigues(np.asarray(R, np.float64))[0] for R in r_arr], dtype=np.float64).ravel()
    x0 = np.concatenate([rv, np.asarray(t_arr, np.float64).ravel(), points_3d.ravel()])
    fn = CostCounter(pidx, cidx, points_2d, k_arr, d_arr, n_cams=n_cams)
    f0 = fn.points_extrinsics(x0)
    A = sparsity_pattern(n_cams, 6, cidx, n_points, pidx, layout=layout)
    t0 = time.perf_counter()
    res = least_squares(fn.points_extrinsics, x0, jac_sparsity=A, verbose=verbose, x_scale='jac', ftol=1e-10, method='trf',
                        loss='cauchy', max_nfev=max_nfev)
    dt = time.perf_counter() - t0
    R = np.array([cv2.Rodrigues(res.x[3 * j:3 * j + 3].copy())[0] for j in range(n_cams)])
    T = res.x[3 * n_cams:6 * n_cams].reshape(n_cams, 3, 1)
    info = dict(seconds=dt, nfev=int(res.nfev), njev=int(res.njev), cost_calls=fn.calls, status=int(res.status), cost=float(res.cost),
                optimality=float(res.optimality), x=res.x)
    return res.x[6 * n_cams:].reshape(n_points, 3), R, T, dict(before=f0, after=res.fun), info


def triangulate_pair_cv(p1, p2, k1, d1, r1, t1, k2, d2, r2, t2):
    """calib.py:120-129 with OpenCV itself."""
    import cv2
    a = cv2.fisheye.undistortPoints(np.asarray(p1, np.float64).reshape(-1, 1, 2), k1, d1)
    b = cv2.fisheye.undistortPoints(np.asarray(p2, np.float64).reshape(-1, 1, 2), k2, d2)
    P1 = np.hstack((r1, np.asarray(t1).reshape(3, 1)))
    P2 = np.hstack((r2, np.asarray(t2).reshape(3, 1)))
    h = cv2.triangulatePoints(P1, P2, a, b)
    return (h[:3] / h[3]).T


def ring_triangulate_df(df, k_arr, d_arr, r_arr, t_arr):
    """utils.py:319-349 with pandas joins, as the reference does it (this is where its TRI time goes)."""
    import pandas as pd
    n = len(k_arr)
    d_arr = np.asarray(d_arr, np.float64).reshape(n, 4)
    parts = []
    for a in range(n):
        b = (a + 1) % n
        both = df[df['camera'] == a].merge(df[df['camera'] == b], how='inner', on=['frame', 'marker'], suffixes=('_a', '_b'))
        if len(both) == 0:
            continue
        X = triangulate_pair_cv(both[['x_a', 'y_a']].to_numpy(np.float64), both[['x_b', 'y_b']].to_numpy(np.float64),
                                k_arr[a], d_arr[a], r_arr[a], t_arr[a], k_arr[b], d_arr[b], r_arr[b], t_arr[b])
        parts.append(pd.DataFrame({'frame': both['frame'].to_numpy(), 'marker': both['marker'].to_numpy(),
                                   'x': X[:, 0], 'y': X[:, 1], 'z': X[:, 2]}))
    if not parts:
        return pd.DataFrame({'frame': [], 'marker': [], 'x': [], 'y': [], 'z': []})
    return pd.concat(parts, ignore_index=True).groupby(['frame', 'marker']).mean().reset_index()


def observation_lists(df, p3):
    """sba.py:296-302: point_index = row id of the triangulated frame; inner merge keeps the 2-D frame's order."""
    p3 = p3.copy()
    p3['point_index'] = p3.index
    m = df.merge(p3, how='inner', on=['frame', 'marker'], suffixes=('_cam', ''))
    return (m[['x_cam', 'y_cam']].to_numpy(np.float64), p3[['x', 'y', 'z']].to_numpy(np.float64),
            m['point_index'].to_numpy().astype(np.int64), m['camera'].to_numpy().astype(np.int64))


def sba_points_scene(uv, lik, thresh, k_arr, d_arr, r_arr, t_arr, f_scale=50.0, n_kpts=20):
    """core.tri + core.sba on a dense synthetic scene: threshold (src/core/sba.py:41), TRI, index merge, points-only SBA."""
    from acinoset_b200 import synth      # data generation helper only (dense slots -> DLC-style DataFrame)
    df = synth.dense_to_df(uv, lik, n_kpts)
    df = df[df['likelihood'] > thresh]
    t0 = time.perf_counter()
    p3 = ring_triangulate_df(df, k_arr, d_arr, r_arr, t_arr)
    t_tri = time.perf_counter() - t0
    obs, X0, pidx, cidx = observation_lists(df, p3)
    X, res, info = bundle_adjust_points_only(obs, X0, pidx, cidx, k_arr, np.asarray(d_arr).reshape(len(k_arr), 4, 1), r_arr, t_arr,
                                             f_scale=f_scale)
    return dict(points=X, frame=p3['frame'].to_numpy(), marker=p3['marker'].to_numpy(), residuals=res, n_obs=len(pidx),
                n_points=len(X0), tri_seconds=t_tri, sba_seconds=info['seconds'], iterations=info['njev'], nfev=info['nfev'],
                cost_calls=info['cost_calls'], cost=info['cost'], pidx=pidx, cidx=cidx, obs=obs, X0=X0)


def sba_extrinsics_scene(uv, lik, thresh, k_arr, d_arr, r0, t0, n_kpts=20, layout='consistent', max_nfev=25):
    """bundle_adjust_points_and_extrinsics (src/lib/sba.py:158-178) fed with DLC-derived arrays: threshold,
    TRI, index merge, then the points+extrinsics solve. ``layout='consistent'`` uses the sparsity pattern that matches the
    parameter layout (the as-shipped pattern stalls at once); ``max_nfev`` bounds the sample — the rate
    (TRF iterations per second) is what the caller wants, not convergence."""
    from acinoset_b200 import synth
    df = synth.dense_to_df(uv, lik, n_kpts)
    df = df[df['likelihood'] > thresh]
    p3 = ring_triangulate_df(df, k_arr, d_arr, r0, t0)
    obs, X0, pidx, cidx = observation_lists(df, p3)
    n = len(k_arr)
    X, R, T, res, info = bundle_adjust_points_and_extrinsics(obs, X0, pidx, cidx, k_arr, np.asarray(d_arr).reshape(n, 4, 1),
                                                             np.asarray(r0), np.asarray(t0).reshape(n, 3, 1), layout=layout,
                                                             max_nfev=max_nfev)
    return dict(points=X, r=R, t=T, residuals=res, n_obs=len(pidx), n_points=len(X0), sba_seconds=info['seconds'],
                iterations=info['njev'], nfev=info['nfev'], cost_calls=info['cost_calls'], cost=info['cost'])
