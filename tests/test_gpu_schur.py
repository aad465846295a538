"""GPU parity tests of the points+extrinsics solver (bundle_adjust_points_and_extrinsics, src/lib/sba.py:158-178).

The reference optimises the same Cauchy objective with a different algorithm (finite-difference TRF) and
the problem has a 7-DoF gauge (F7), so parity is asserted on gauge-invariant quantities: the residual vector at equal
parameters (1e-9 relative), the converged cost / RMSE (1e-4 px) against the reference run with the layout-consistent
sparsity pattern (tests/golden/golden_calib.npz, made by oracle/make_golden_calib.py from the REAL reference), and
against the dense numpy LM of oracle/sba_lm_np.py on a small case.
"""
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from acinoset_b200 import calib, sba, dense, synth, _capi
from oracle import fisheye_np as fe
from oracle import sba_lm_np as lm_np


def _rel(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


def _small_problem(seed=0, P=40, noise=0.3):
    k, d, r, t, _ = synth.load_rig()
    rng = np.random.default_rng(seed)
    Xg = np.stack([rng.uniform(-1, 5, P), rng.uniform(5, 8, P), rng.uniform(0, 1, P)], 1)
    pidx, cidx, obs = [], [], []
    for p in range(P):
        cams = [c for c in range(6) if rng.uniform() < 0.6]
        if len(cams) < 2:
            cams = [0, 1]
        for c in cams:
            pidx.append(p); cidx.append(c)
            obs.append(fe.project_points_fisheye(Xg[p], k[c], d[c], r[c], t[c])[0] + rng.normal(0, noise, 2))
    r0 = np.array([r[c] @ synth._rot_from_axis_angle(rng.normal(0, 8e-4, 3)) for c in range(6)])
    t0 = t + rng.normal(0, 3e-3, t.shape)
    X0 = Xg + rng.normal(0, 0.01, Xg.shape)
    return k, d, r0, t0, X0, np.array(obs), np.array(pidx), np.array(cidx)


def test_extrinsics_residuals_vs_oracle():
    k, d, r0, t0, X0, obs, pidx, cidx = _small_problem()
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r0])
    x0 = np.concatenate([rv.ravel(), t0.ravel(), X0.ravel()])
    f = sba.cost_func_points_extrinsics(x0, 6, len(X0), pidx, cidx, k, d, obs, calib.project_points_fisheye)
    f_o = fe.residuals_points_extrinsics(x0, 6, len(X0), pidx, cidx, k, d, obs)
    assert _rel(f, f_o).max() < 1e-9


def test_schur_lm_matches_dense_numpy_lm():
    """Same objective, same damping policy: the CUDA Schur LM and the dense numpy LM land on the same minimum."""
    k, d, r0, t0, X0, obs, pidx, cidx = _small_problem()
    prob = lm_np.Problem(obs, pidx, cidx, k, d, len(X0), f_scale=1.0)
    R, T, X, info = lm_np.solve(prob, r0, t0, X0, max_iter=200)
    obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(obs, X0, pidx, cidx, k, d, r0, t0, calib.project_points_fisheye)
    st = sba.bundle_adjust_points_and_extrinsics.last_stats
    cost_gpu = fe.cauchy_cost(res['after'], 1.0)
    assert abs(st['cost_final'] - cost_gpu) <= 1e-9 * cost_gpu          # device bookkeeping == oracle cost of device result
    assert abs(st['cost_initial'] - fe.cauchy_cost(res['before'], 1.0)) <= 1e-9 * st['cost_initial']
    assert cost_gpu <= info['cost'] * (1 + 1e-7)
    assert abs(cost_gpu - info['cost']) <= 1e-6 * info['cost']
    # residual vectors are gauge invariant
    f_np = prob.residuals(R, T, X)
    assert np.abs(res['after'] - f_np).max() < 1e-4
    # returned rotations are rotations
    for m in r1:
        assert np.abs(m @ m.T - np.eye(3)).max() < 1e-12 and abs(np.linalg.det(m) - 1) < 1e-12
    assert r1.shape == (6, 3, 3) and t1.shape == (6, 3, 1) and obj.shape == X0.shape


def test_schur_vs_reference_golden_calib(rig, golden_dir):
    """Config C2 (small): against the real reference run with the layout-consistent sparsity pattern."""
    g = np.load(os.path.join(golden_dir, 'golden_calib.npz'))
    k, d, r, t, _ = rig
    obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(g['points_2d'], g['points_3d'], g['point_indices'],
                                                               g['camera_indices'], k, d, g['r0'], g['t0'],
                                                               calib.project_points_fisheye)
    assert _rel(res['before'], g['fixed_before']).max() < 1e-9
    cost_ref = fe.cauchy_cost(g['fixed_after'], 1.0)
    cost_new = fe.cauchy_cost(res['after'], 1.0)
    rmse_ref = np.sqrt(np.mean(g['fixed_after'] ** 2))
    rmse_new = np.sqrt(np.mean(res['after'] ** 2))
    assert cost_new <= cost_ref * (1 + 1e-9), (cost_new, cost_ref)
    assert abs(rmse_new - rmse_ref) < 1e-4, (rmse_new, rmse_ref)
    # the reference as shipped stalls; recorded, and we must be no worse than it
    assert cost_new < fe.cauchy_cost(g['shipped_after'], 1.0)


def test_schur_dense_rig_scene_and_reset():
    """Dense rig layout (the bench path): solve, check cost bookkeeping, reset_state restores points and cameras."""
    sc = synth.make_rig_scene(30, 20, seed=9)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(1)
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
    tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    s = dense.SbaSolver(C, P)
    s.set_cameras(k, d, r, t)
    s.set_rvec_t(rv, tv)
    s.set_observations(sc['uv'], mask)
    s.set_points(np.nan_to_num(xyz))
    cams0 = s.get_cameras()
    f0 = s.residuals()
    pm = (mask[:, None] >> np.arange(C)) & 1
    sel = pm.astype(bool)
    c0 = fe.cauchy_cost(f0[sel], 1.0)
    st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS))
    assert st['n_unconverged'] == 0, st
    assert abs(st['cost_initial'] - c0) <= 1e-9 * c0
    c1 = fe.cauchy_cost(s.residuals()[sel], 1.0)
    assert abs(st['cost_final'] - c1) <= 1e-9 * c1 and c1 < c0
    # gradient at the solution is (numerically) zero: one more linearisation reports it
    assert st['gmax'] < 1e-3 * max(1.0, st['cost_final'])
    s.reset_state()
    assert np.array_equal(s.get_cameras(), cams0)
    assert np.array_equal(s.get_points(), np.nan_to_num(xyz))
    # points-only on the same handle still works after an extrinsics solve
    st2 = s.solve(dense.default_opts(_capi.MODE_POINTS_ONLY, f_scale=1.0))
    assert st2['n_unconverged'] == 0
    s.close()


def test_schur_iterate_matches_solve():
    """acino_sba_iterate (no host sync, the bench path) walks the same iterates as acino_sba_solve."""
    k, d, r0, t0, X0, obs, pidx, cidx = _small_problem(seed=3, P=200)
    uv, mask, slot = sba._lists_to_dense(obs, pidx, cidx, len(X0), 6)
    a = dense.SbaSolver(6, len(X0)); b = dense.SbaSolver(6, len(X0))
    for s in (a, b):
        s.set_cameras(k, d, r0, t0)
        s.set_observations(uv, mask)
        s.set_points(X0)
    for refine in (0, 2):
        for s in (a, b):
            s.set_cameras(k, d, r0, t0)
            s.set_points(X0)
        st = a.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, max_iter=7, ftol=0.0, gtol=0.0, xtol=0.0,
                                        point_refine_iters=refine))
        o = dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, force_iters=1, point_refine_iters=refine)
        for _ in range(7):
            b.iterate(o)
        lb = b.lm_state()
        assert lb['iterations'] == 7 == st['iterations']
        assert np.array_equal(a.get_points(), b.get_points())
        assert np.array_equal(a.get_cameras(), b.get_cameras())
        if refine == 0:
            assert lb['cost'] == st['cost_final']
        else:       # iterate leaves the cost of the last refinement to the next linearisation; solve refreshes it
            assert st['cost_final'] <= lb['cost']
    a.close(); b.close()


def test_polishing_reaches_the_same_minimum_in_fewer_iterations():
    """polish_tol (points solved to convergence with the cameras fixed once the joint steps creep) is an
    acceleration, not a different problem: same stopping status, a final cost that is not worse than without it beyond
    the noise between runs of one nonconvex problem (1e-5 relative), fewer joint iterations, cost bookkeeping still exact."""
    sc = synth.make_rig_scene(3000, 20, seed=21)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(5)
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
    tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    s = dense.SbaSolver(C, P)
    s.set_observations(sc['uv'], mask)
    sel = ((mask[:, None] >> np.arange(C)) & 1).astype(bool)
    res = {}
    for tol in (0.0, 3e-4):
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
        st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, ftol=1e-10, max_iter=400, polish_tol=tol))
        c = fe.cauchy_cost(s.residuals()[sel], 1.0)
        assert abs(st['cost_final'] - c) <= 1e-9 * c
        assert st['n_unconverged'] == 0
        res[tol] = st
    assert res[3e-4]['cost_final'] <= res[0.0]['cost_final'] * (1 + 1e-5)
    assert res[3e-4]['iterations'] < res[0.0]['iterations']
    s.close()


def test_polish_paths_agree(monkeypatch):
    """The three ways a polish can run - the points the linearisation listed by one warp per point (the default for short
    lists), the same list by one thread per point (ACINO_POLISH_WARP=0), and a first pass over all points
    (ACINO_POLISH_LIST=0, the path a list that overflowed takes) - are the same LM on the same objective: they end with
    the same status, costs within 1e-8 relative of each other (the sums run in another order; which stragglers a pass
    sees differs by a few points) and iteration counts within two. The switches are read when a handle is created."""
    sc = synth.make_rig_scene(3000, 20, seed=21, outlier_frac=0.03)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(5)
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
    tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    sel = ((mask[:, None] >> np.arange(C)) & 1).astype(bool)
    res = {}
    for name, env in (('warp', {}), ('thread', {'ACINO_POLISH_WARP': '0'}), ('all', {'ACINO_POLISH_LIST': '0'})):
        for key in ('ACINO_POLISH_WARP', 'ACINO_POLISH_LIST'):
            monkeypatch.delenv(key, raising=False)
        for key, v in env.items():
            monkeypatch.setenv(key, v)
        s = dense.SbaSolver(C, P)
        s.set_observations(sc['uv'], mask)
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
        st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, ftol=1e-10, max_iter=400))
        c = fe.cauchy_cost(s.residuals()[sel], 1.0)
        assert abs(st['cost_final'] - c) <= 1e-9 * c, name           # the reported cost is the cost of what comes back
        assert st['n_unconverged'] == 0, name
        res[name] = st
        s.close()
    for name in ('thread', 'all'):
        assert res[name]['status'] == res['warp']['status']
        assert abs(res[name]['cost_final'] - res['warp']['cost_final']) <= 1e-8 * res['warp']['cost_final'], (name, res[name]['cost_final'], res['warp']['cost_final'])
        assert abs(res[name]['iterations'] - res['warp']['iterations']) <= 2, name


def test_board_calibration_ends_by_a_stopping_rule():
    """Config C2 in small: a checkerboard scene converges within a few tens of joint iterations, and the
    solve has to END there by one of scipy's rules. (Polishing used to reset the ftol / stall status each time it ran,
    so that on small scenes polish -> rejected steps -> polish cycled until max_iter.)"""
    b = synth.make_board_scene(60, seed=3)
    p2, p3, pi, ci = sba.prepare_calib_board_data_for_bundle_adjustment(b['img_pts_arr'], b['fnames_arr'], (9, 6), b['k'], b['d'],
                                                                        b['r0'], b['t0'], calib.triangulate_points_fisheye)
    obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(p2, p3, pi, ci, b['k'], b['d'], b['r0'].copy(), b['t0'].copy(),
                                                               calib.project_points_fisheye)
    st = sba.bundle_adjust_points_and_extrinsics.last_stats
    assert st['status'] in (1.0, 2.0, 3.0, 4.0), st
    assert st['iterations'] < 80, st
    rmse0, rmse1 = np.sqrt(np.mean(res['before'] ** 2)), np.sqrt(np.mean(res['after'] ** 2))
    assert rmse1 < rmse0 and rmse1 < 0.35          # 0.3 px pixel noise: the solve reaches the noise floor
    # the reference's own ftol ends no later, at a cost that is the same to 1e-8
    obj2, r2, t2, res2 = sba.bundle_adjust_points_and_extrinsics(p2, p3, pi, ci, b['k'], b['d'], b['r0'].copy(), b['t0'].copy(),
                                                                 calib.project_points_fisheye, ftol=1e-10)
    st2 = sba.bundle_adjust_points_and_extrinsics.last_stats
    assert st2['status'] in (1.0, 2.0, 3.0, 4.0) and st2['iterations'] <= st['iterations']
    assert abs(st2['cost_final'] - st['cost_final']) <= 1e-8 * st['cost_final']


def test_point_order_does_not_change_results():
    """K3 visits the points sorted by observation count (point_order_kernel); every point is still solved by the same
    arithmetic, so a scene and its frame-reversed copy give bit-identical points, also when P is not a multiple of the
    4096-point order tile and when whole tiles are unobserved."""
    sc = synth.make_rig_scene(517, 20, seed=33)            # 10,340 points: 2 full tiles + a ragged one
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    lik = sc['lik'].copy()
    lik[4096:8192] = 0.0                                    # one whole tile without observations
    a = dense.solve_scene_host(sc['uv'], lik, 0.5, k, d, r, t)
    b = dense.solve_scene_host(sc['uv'][::-1].copy(), lik[::-1].copy(), 0.5, k, d, r, t)
    assert np.isnan(a['points'][4096:8192]).all()
    assert np.array_equal(a['points'], b['points'][::-1], equal_nan=True)
    assert a['n_unconverged'] == 0 and b['n_unconverged'] == 0


def test_fp64_peak_is_sane():
    tf = dense.fp64_peak_tflops()
    assert 5.0 < tf < 80.0, tf
