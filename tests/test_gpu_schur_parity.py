"""GPU parity tests of the points+extrinsics path beyond cost / RMSE:

* the analytic Jacobian the kernels are built from against CENTRAL differences of the reference's residual function
  (oracle/fisheye_np.py, itself pinned by fixtures of the real reference);
* one LM iteration and a full solve of a masked + outlier rig scene with more than a thousand points (every mask class,
  many rounds, several CTAs and tiles of K4b) against the block-sparse numpy LM of oracle/sba_lm_np.py;
* the converged calibration scene of tests/golden/golden_calib.npz (real reference, layout-consistent pattern) on
  GEOMETRY: points and camera centres after a 7-DoF similarity alignment;
* rows of points without observations survive a solve untouched whatever the number of accepted steps.
"""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from acinoset_b200 import calib, sba, dense, synth, _capi
from oracle import fisheye_np as fe
from oracle import sba_lm_np as lm_np


def _umeyama(src, dst):
    """Similarity (s, R, t) minimising |s R src + t - dst|^2 (7 DoF: the gauge of the problem)."""
    mu_s, mu_d = src.mean(0), dst.mean(0)
    a, b = src - mu_s, dst - mu_d
    U, S, Vt = np.linalg.svd(b.T @ a / len(src))
    D = np.eye(3)
    if np.linalg.det(U) * np.linalg.det(Vt) < 0:
        D[2, 2] = -1
    R = U @ D @ Vt
    s = np.trace(np.diag(S) @ D) / (a ** 2).sum() * len(src)
    return s, R, mu_d - s * R @ mu_s


def _rig_problem(n_frames, seed, masked_frac=0.3, outlier_frac=0.05):
    """A rig scene as lists: triangulated start, perturbed extrinsics (the C4 / C5 recipe of SURVEY.md §8d in small)."""
    sc = synth.make_rig_scene(n_frames, 20, seed=seed, masked_frac=masked_frac, outlier_frac=outlier_frac)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(seed + 100)
    r0 = np.array([r[c] @ synth._rot_from_axis_angle(rng.normal(0, 8e-4, 3)) for c in range(C)])
    t0 = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    sel = ((mask[:, None] >> np.arange(C)) & 1).astype(bool)
    pidx, cidx = np.nonzero(sel)
    obs = sc['uv'][pidx, cidx]
    return dict(k=k, d=d, r0=r0, t0=t0, X0=np.nan_to_num(xyz), uv=sc['uv'], mask=mask, sel=sel, pidx=pidx, cidx=cidx, obs=obs,
                P=P, C=C)


def test_analytic_jacobian_vs_central_differences_of_the_reference_residual(rig):
    """d(u,v)/d(w, tau, X) of the kernels against central differences of the restated reference
    residual (project_points_fisheye per observation, src/lib/sba.py:142-153) with the camera moved as R Exp(w), t + R tau."""
    k, d, r, t, _ = rig
    C = len(k)
    rng = np.random.default_rng(7)
    n = 600
    X = np.stack([rng.uniform(-1, 5, n), rng.uniform(5, 8, n), rng.uniform(0, 1, n)], 1)
    pidx = np.arange(n)
    cidx = rng.integers(0, C, n)
    J = sba.jacobian_points_extrinsics(X, pidx, cidx, k, d, r, t)
    tt = np.asarray(t, np.float64).reshape(C, 3)

    def proj(Xq, Rq, tq):
        return np.stack([fe.project_points_fisheye(Xq[i], k[cidx[i]], d[cidx[i]], Rq[i], tq[i])[0] for i in range(n)])

    R_o, t_o = np.asarray(r, np.float64)[cidx], tt[cidx]
    Jn = np.zeros((n, 2, 9))
    for col in range(9):
        h = 1e-6 if col >= 3 else 1e-7
        out = []
        for sgn in (+1, -1):
            Xq, Rq, tq = X.copy(), R_o.copy(), t_o.copy()
            if col < 3:
                w = np.zeros(3); w[col] = sgn * h
                Rq = np.einsum('nij,jk->nik', R_o, fe.rodrigues_vec_to_mat(w))
            elif col < 6:
                tau = np.zeros(3); tau[col - 3] = sgn * h
                tq = t_o + R_o @ tau
            else:
                Xq[:, col - 6] += sgn * h
            out.append(proj(Xq, Rq, tq))
        Jn[:, :, col] = (out[0] - out[1]) / (2 * h)
    scale = np.maximum(1.0, np.abs(Jn).max(axis=(1, 2), keepdims=True))
    assert (np.abs(J - Jn) / scale).max() < 2e-6          # central differences: O(h^2) truncation + 1e-16/h rounding of ~1e3 px
    # and the residual whose derivative this is, is the reference's (1e-9 relative, as everywhere)
    f = sba.cost_func_points_only(X.ravel(), n, pidx, cidx, k, d, r, t, np.zeros((n, 2)), calib.project_points_fisheye)
    assert (np.abs(f.reshape(n, 2) - proj(X, R_o, t_o)) / np.maximum(1.0, np.abs(f.reshape(n, 2)))).max() < 1e-9


def test_one_lm_iteration_matches_block_oracle_on_a_multi_class_scene():
    """K4a + K4b + K5 + K6 + K7 once, at several fixed dampings, against the block-sparse numpy statement of the same step:
    cost, gradient maximum, predicted reduction, trial cost, the accept / reject decision and - where the step is
    accepted - the new points and cameras. 90 frames, 30 % masked, 5 % gross outliers: > 1,400 existing points in every
    mask class (multi-round, multi-CTA, multi-tile). (At damping 1 the robust cost of this outlier scene RISES along the
    step: both sides must reject it and leave the parameters alone.)"""
    pr = _rig_problem(90, seed=11)
    C, P = pr['C'], pr['P']
    assert (pr['mask'] != 0).sum() > 1400 and len(np.unique(pr['mask'])) > 40
    s = dense.SbaSolver(C, P)
    s.set_cameras(pr['k'], pr['d'], pr['r0'], pr['t0'])
    s.set_observations(pr['uv'], pr['mask'])
    s.set_points(pr['X0'])
    prob = lm_np.BlockProblem(pr['obs'], pr['pidx'], pr['cidx'], pr['k'], pr['d'], P, f_scale=1.0, hess_floor=0.3)
    L = prob.linearise(pr['r0'], pr['t0'], pr['X0'])
    ex = pr['mask'] != 0
    decisions = set()
    for lam in (1e4, 1e2, 1.0, 1e-3):
        s.set_cameras(pr['k'], pr['d'], pr['r0'], pr['t0'])
        s.set_points(pr['X0'])
        s.iterate(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, force_iters=1, point_refine_iters=0, lambda0=lam))
        g = s.lm_state()
        st = prob.step(L, lam)
        Rn, tn = lm_np.apply_twists(pr['r0'], pr['t0'], st['dc'])
        Xn = pr['X0'] + st['dp']
        Fn = prob.cost(Rn, tn, Xn)
        assert abs(g['cost_first'] - L['F']) <= 1e-12 * L['F']
        assert abs(g['pred'] - st['pred']) <= 1e-8 * abs(st['pred']), (lam, g['pred'], st['pred'])
        assert abs(g['cost_trial'] - Fn) <= 1e-9 * Fn, (lam, g['cost_trial'], Fn)
        assert abs(g['gmax'] - st['gmax']) <= 1e-9 * st['gmax']
        accept = bool(Fn < L['F'] and st['pred'] > 0)
        decisions.add(accept)
        assert g['accepted'] == int(accept), (lam, Fn, L['F'])
        cams = s.get_cameras().reshape(C, 24)
        R_want, t_want, X_want = (Rn, tn, Xn) if accept else (pr['r0'], pr['t0'], pr['X0'])
        assert np.abs(cams[:, 9:18].reshape(C, 3, 3) - R_want).max() < 1e-10
        assert np.abs(cams[:, 18:21] - t_want).max() < 1e-9
        assert np.abs(s.get_points()[ex] - X_want[ex]).max() < 1e-9          # metres
        assert np.array_equal(s.get_points()[~ex], pr['X0'][~ex])
        assert g['cost'] == (g['cost_trial'] if accept else g['cost_first'])
    assert decisions == {True, False}          # both branches of K7 were compared
    s.close()


def test_full_solve_matches_block_oracle_on_a_multi_class_scene():
    """The converged solve (joint steps without per-point refinement or polishing, i.e. the plain LM both sides run)
    against the numpy LM from the same start: cost 1e-9 relative; residual vector (gauge invariant) 1e-4 px, the north
    star's RMSE tolerance - two LMs that stop by their own rules near a minimum with weakly determined directions agree
    to about 1e-5 px here, not to rounding; and the CUDA endpoint is a stationary point of the ORACLE's objective: the
    oracle's own damped step from it predicts a reduction below 1e-10 of the cost."""
    pr = _rig_problem(70, seed=12, masked_frac=0.3, outlier_frac=0.0)
    C, P = pr['C'], pr['P']
    s = dense.SbaSolver(C, P)
    s.set_cameras(pr['k'], pr['d'], pr['r0'], pr['t0'])
    s.set_observations(pr['uv'], pr['mask'])
    s.set_points(pr['X0'])
    st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, point_refine_iters=0, polish_tol=0.0, max_iter=300,
                                    ftol=1e-15, gtol=1e-10))
    f_gpu = s.residuals()[pr['sel']].ravel()
    prob = lm_np.BlockProblem(pr['obs'], pr['pidx'], pr['cidx'], pr['k'], pr['d'], P, f_scale=1.0, hess_floor=0.3)
    R, t, X, info = lm_np.solve_block(prob, pr['r0'], pr['t0'], pr['X0'], lam=1.0, max_iter=300, gtol=1e-10, ftol=1e-15)
    f_np = prob.residuals(R, t, X)
    assert abs(st['cost_final'] - info['cost']) <= 1e-9 * info['cost'], (st, info)
    assert np.abs(f_gpu - f_np).max() < 1e-4
    assert abs(np.sqrt(np.mean(f_gpu ** 2)) - np.sqrt(np.mean(f_np ** 2))) < 1e-7
    cams = s.get_cameras().reshape(C, 24)
    Xg = s.get_points()
    Lg = prob.linearise(cams[:, 9:18].reshape(C, 3, 3), cams[:, 18:21], Xg)
    assert abs(Lg['F'] - st['cost_final']) <= 1e-12 * Lg['F']
    stp = prob.step(Lg, 1e-4)
    print('oracle step at the CUDA endpoint: predicted reduction %.3e of cost %.6f, max |dX| %.3e m' % (stp['pred'], Lg['F'], np.abs(stp['dp']).max()))
    assert 0.0 <= stp['pred'] <= 1e-10 * Lg['F']
    s.close()


def test_calibration_geometry_after_similarity_alignment(rig, golden_dir):
    """the converged calibration of the real reference (layout-consistent pattern,
    tests/golden/golden_calib.npz) against the CUDA solve on GEOMETRY. The problem has a 7-DoF gauge (F7): all cameras
    and points are free, so raw coordinates differ by a similarity, removed here by a Umeyama alignment on the points.

    Achieved bound against the reference's endpoint: points 1.4 mm, camera centres 1.5 cm, residual vector 0.04 px -
    NOT the 0.1 mm of the north star, and the test shows why that is the reference's distance from the minimum and not
    this solver's: the reference (finite-difference TRF) stops by ftol = 1e-10 while this 8-board scene leaves the
    camera centres weakly determined. (a) The CUDA solve reaches a LOWER cost than the reference's endpoint. (b) Started
    AT the reference's endpoint, the CUDA solve keeps descending and lands on the same minimum as the solve from the
    initial guess: points and camera centres within 0.1 mm after alignment, residual vector within 1e-4 px."""
    g = np.load(os.path.join(golden_dir, 'golden_calib.npz'))
    k, d, r, t, _ = rig
    args = (g['points_2d'], g['points_3d'], g['point_indices'], g['camera_indices'], k, d)
    kw = dict(ftol=1e-15, max_iter=600)
    obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(*args, g['r0'], g['t0'], calib.project_points_fisheye, **kw)
    cen = lambda rr, tt: np.stack([-(rr[c].T @ np.asarray(tt[c]).reshape(3)) for c in range(len(rr))])

    def aligned_distance(obj_a, r_a, t_a, obj_b, r_b, t_b):
        s_, R_, tr_ = _umeyama(obj_a, obj_b)
        d_pts = np.linalg.norm(s_ * obj_a @ R_.T + tr_ - obj_b, axis=1)
        d_cen = np.linalg.norm(s_ * cen(r_a, t_a) @ R_.T + tr_ - cen(r_b, t_b), axis=1)
        return s_, d_pts, d_cen

    s_, d_pts, d_cen = aligned_distance(obj, r1, t1, g['fixed_obj'], g['fixed_r'], g['fixed_t'])
    print('vs the reference endpoint: scale %.9f; points max %.3e m (rms %.3e); camera centres max %.3e m; residual vector max diff %.3e px'
          % (s_, d_pts.max(), np.sqrt((d_pts ** 2).mean()), d_cen.max(), np.abs(res['after'] - g['fixed_after']).max()))
    assert abs(s_ - 1.0) < 1e-2
    assert d_pts.max() < 3e-3 and d_cen.max() < 3e-2
    assert np.abs(res['after'] - g['fixed_after']).max() < 0.1
    cost_ref, cost_new = fe.cauchy_cost(g['fixed_after'], 1.0), fe.cauchy_cost(res['after'], 1.0)
    assert cost_new < cost_ref                                        # (a)
    # (b) from the reference's endpoint
    obj2, r2, t2, res2 = sba.bundle_adjust_points_and_extrinsics(g['points_2d'], g['fixed_obj'], g['point_indices'], g['camera_indices'],
                                                                 k, d, g['fixed_r'], g['fixed_t'], calib.project_points_fisheye, **kw)
    assert np.abs(res2['before'] - g['fixed_after']).max() < 1e-9 * np.abs(g['fixed_after']).max() + 1e-9
    cost2 = fe.cauchy_cost(res2['after'], 1.0)
    s2, d_pts2, d_cen2 = aligned_distance(obj2, r2, t2, obj, r1, t1)
    print('from the reference endpoint: cost %.9f -> %.9f (from the initial guess: %.9f); points max %.3e m; camera centres max %.3e m; '
          'residual vector max diff %.3e px' % (cost_ref, cost2, cost_new, d_pts2.max(), d_cen2.max(), np.abs(res2['after'] - res['after']).max()))
    assert cost2 < cost_ref and abs(cost2 - cost_new) <= 1e-9 * cost_new
    assert d_pts2.max() < 1e-4 and d_cen2.max() < 1e-4
    assert np.abs(res2['after'] - res['after']).max() < 1e-4


def test_rows_of_absent_points_survive_every_accept_parity():
    """a point without observations is not a parameter of the problem (scipy leaves parameters without
    residuals where they are); its row must come back as it was set - NaN here - after an odd and after an even number
    of accepted steps (the trial buffer becomes the current one with every accepted step). A second solve on the same
    handle with other masks: rows that stop being observed keep what the caller set."""
    pr = _rig_problem(40, seed=5, masked_frac=0.5, outlier_frac=0.0)
    C, P = pr['C'], pr['P']
    ex = pr['mask'] != 0
    assert (~ex).sum() > 50
    X0 = pr['X0'].copy()
    X0[~ex] = np.nan
    s = dense.SbaSolver(C, P)
    parities = set()
    for max_iter in (1, 2, 3, 4, 5):
        s.set_cameras(pr['k'], pr['d'], pr['r0'], pr['t0'])
        s.set_observations(pr['uv'], pr['mask'])
        s.set_points(X0)
        # a large initial damping: short steps that are accepted from the first iteration on
        opts = dict(max_iter=max_iter, lambda0=1e3, ftol=0.0, gtol=0.0, xtol=0.0, polish_tol=0.0)
        st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, **opts))
        out = s.get_points()
        assert st['n_accept'] >= 1, st
        parities.add(int(st['n_accept']) % 2)
        assert np.isnan(out[~ex]).all()
        assert np.isfinite(out[ex]).all() and np.isfinite(st['cost_final'])
        assert st['cost_final'] < st['cost_initial']
        mask2 = pr['mask'].copy()
        drop = np.nonzero(ex)[0][::7]
        mask2[drop] = 0
        s.set_observations(pr['uv'], mask2)
        s.set_points(out)
        st2 = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, **opts))
        out2 = s.get_points()
        assert st2['n_accept'] >= 1
        assert np.array_equal(out2[drop], out[drop])
        assert np.isnan(out2[~ex]).all()
        keep = ex & (mask2 != 0)
        assert np.isfinite(out2[keep]).all() and not np.array_equal(out2[keep], out[keep])
    assert parities == {0, 1}
    s.close()
