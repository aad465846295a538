"""CPU tests: pin the oracle (oracle/fisheye_np.py restatement and oracle/ref_port.py port) against fixtures produced
by the REAL reference (oracle/make_golden.py, run in the build container), and check the host-side logic.
No GPU needed."""
import os
import numpy as np
import pytest

from oracle import fisheye_np as oracle
from oracle import ref_port
from acinoset_b200 import synth


def _rel(a, b):
    return np.abs(a - b) / np.maximum(1.0, np.abs(b))


@pytest.fixture(scope='module')
def gcam(golden_dir):
    return np.load(os.path.join(golden_dir, 'golden_camera.npz'))


@pytest.fixture(scope='module')
def gc1(golden_dir):
    return np.load(os.path.join(golden_dir, 'golden_c1.npz'))


def test_projection_restatement_vs_reference(rig, gcam):
    k, d, r, t, _ = rig
    for c in range(6):
        assert _rel(oracle.project_points_fisheye(gcam['X'], k[c], d[c], r[c], t[c]), gcam['uv'][c]).max() < 1e-10
        assert _rel(oracle.project_points_fisheye(gcam['Xw'], k[c], d[c], r[c], t[c]), gcam['uvw'][c]).max() < 1e-9


def test_undistort_restatement_vs_opencv(rig, gcam):
    k, d, r, t, _ = rig
    for c in range(6):
        assert np.abs(oracle.undistort_points_fisheye(gcam['noisy'][c], k[c], d[c]) - gcam['und'][c]).max() < 1e-13
    e = oracle.undistort_points_fisheye(gcam['edge'], k[0], d[0])
    ref = gcam['und_edge']
    assert np.allclose(e, ref, rtol=1e-12, atol=1e-13)


def test_two_view_dlt_restatement_vs_reference(rig, gcam):
    k, d, r, t, _ = rig
    for c in range(6):
        b = (c + 1) % 6
        X = oracle.triangulate_points_fisheye(gcam['noisy'][c], gcam['noisy'][b], k[c], d[c], r[c], t[c], k[b], d[b], r[b], t[b])
        assert np.abs(X - gcam['tri'][c]).max() < 1e-10


def test_rodrigues_restatement_vs_opencv(gcam):
    R = np.stack([oracle.rodrigues_vec_to_mat(v) for v in gcam['rv']])
    assert np.abs(R - gcam['rmat']).max() < 1e-14


def test_ring_triangulation_restatement_c1(gc1):
    """Dense-slot restatement of get_pairwise_3d_points_from_df reproduces the reference's rows (order = frame, then
    marker name) and values on config C1."""
    sc = synth.make_rig_scene(102, 20, seed=0)
    xyz, exists, mask = oracle.ring_triangulate_dense(sc['uv'], sc['lik'], 0.5, sc['k'], sc['d'], sc['r'], sc['t'])
    frames = 69 + np.repeat(np.arange(102), 20)
    markers = np.tile(np.array(synth.MARKERS), 102)
    order = np.lexsort((markers[exists], frames[exists]))
    assert (frames[exists][order] == gc1['frame']).all()
    assert (markers[exists][order] == gc1['marker']).all()
    assert np.abs(xyz[exists][order] - gc1['points_3d']).max() < 1e-10


def test_restatements_on_the_outlier_scene(rig, golden_dir):
    """Config C5 in small (30 % of the slots masked, 5 % gross outliers; golden_c5small.npz from the real reference):
    the dense-slot triangulation reproduces the reference's rows and values, and the residual restatement its vector -
    also where undistortion fails or residuals are hundreds of pixels."""
    g = np.load(os.path.join(golden_dir, 'golden_c5small.npz'))
    k, d, r, t, _ = rig
    sc = synth.make_rig_scene(60, 20, seed=5, masked_frac=0.3, outlier_frac=0.05)
    xyz, exists, mask = oracle.ring_triangulate_dense(sc['uv'], sc['lik'], 0.5, sc['k'], sc['d'], sc['r'], sc['t'])
    frames = np.repeat(np.arange(60), 20)
    markers = np.tile(np.array(synth.MARKERS), 60)
    order = np.lexsort((markers[exists], frames[exists]))
    assert (frames[exists][order] == g['frame']).all()
    assert (markers[exists][order] == g['marker']).all()
    assert np.abs(xyz[exists][order] - g['points_3d']).max() < 1e-9
    f = oracle.residuals_points_only(g['points_3d'].ravel(), len(g['points_3d']), g['point_indices'], g['camera_indices'],
                                     k, d, r, t, g['points_2d'])
    assert _rel(f, g['f_before']).max() < 1e-10
    assert np.abs(g['f_before']).max() > 100.0          # the scene does contain gross outliers


def test_residual_restatement_c1(rig, gc1):
    k, d, r, t, _ = rig
    n = len(gc1['points_3d'])
    f = oracle.residuals_points_only(gc1['points_3d'].ravel(), n, gc1['point_indices'], gc1['camera_indices'], k, d, r, t, gc1['points_2d'])
    assert _rel(f, gc1['f_before']).max() < 1e-10
    f = oracle.residuals_points_only(gc1['x_pert'], n, gc1['point_indices'], gc1['camera_indices'], k, d, r, t, gc1['points_2d'])
    assert _rel(f, gc1['f_pert']).max() < 1e-10


def test_ref_port_cost_function_is_the_reference_loop(rig, gc1):
    """The per-observation OpenCV loop of the port gives the reference's residual vector bit-for-bit (same library calls)."""
    k, d, r, t, _ = rig
    sel = slice(0, 300)
    fn = ref_port.CostCounter(gc1['point_indices'][sel], gc1['camera_indices'][sel], gc1['points_2d'][sel], k, d, r, t)
    f = fn.points_only(gc1['points_3d'].ravel())
    assert np.array_equal(f, gc1['f_before'][:600])


def test_ref_port_scene_pipeline_matches_reference_on_c1_prefix(gc1):
    """TRI + merge + scipy solve of the port on the first frames of C1 lands on the reference's converged points.
    (Points are independent problems, SURVEY.md F4, so a prefix of the scene converges to the same per-point minimisers
    up to scipy's global stopping rule: agreement is asserted at 0.05 mm.)"""
    sc = synth.make_rig_scene(102, 20, seed=0)
    nf = 6
    out = ref_port.sba_points_scene(sc['uv'][:nf * 20], sc['lik'][:nf * 20], 0.5, sc['k'], sc['d'], sc['r'], sc['t'])
    n = out['n_points']
    assert (out['frame'] + 69 == gc1['frame'][:n]).all() and (out['marker'] == gc1['marker'][:n]).all()
    assert np.abs(out['X0'] - gc1['points_3d'][:n]).max() < 1e-12
    assert (out['pidx'].max() == n - 1)
    assert np.linalg.norm(out['points'] - gc1['obj_pts'][:n], axis=1).max() < 5e-5


def test_sparsity_pattern_shapes():
    cidx = np.array([0, 1, 1, 2]); pidx = np.array([0, 0, 1, 1])
    A = ref_port.sparsity_pattern(3, 0, cidx, 2, pidx).toarray()
    assert A.shape == (8, 6) and A.sum() == 8 * 3
    S = ref_port.sparsity_pattern(3, 6, cidx, 2, pidx, layout='shipped').toarray()
    Cn = ref_port.sparsity_pattern(3, 6, cidx, 2, pidx, layout='consistent').toarray()
    assert S.shape == Cn.shape == (8, 24) and S.sum() == Cn.sum() == 8 * 9
    # shipped: obs 0 (camera 0) marks columns 0..5; consistent: rvec 0..2 and tvec 9..11
    assert S[0, :6].all() and not S[0, 6:18].any()
    assert Cn[0, [0, 1, 2, 9, 10, 11]].all() and Cn[0, :18].sum() == 6


def test_oracle_projection_vs_reference_project2d_golden(golden_dir):
    """N2 oracle pin: numpy projection + the range rule of save_3d_cheetah_as_2d (src/lib/utils.py:261-266) against the
    tables the reference function itself produced."""
    import os
    import numpy as np
    from oracle import fisheye_np as fe
    from acinoset_b200 import synth
    g = np.load(os.path.join(golden_dir, 'golden_project2d.npz'))
    k, d, r, t, res = synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json'))
    pos = g['positions']
    for c in range(6):
        prj = fe.project_points_fisheye(pos.reshape(-1, 3), k[c], d[c], r[c], t[c])
        bad = np.where((prj > np.array(res)) | (prj < 0))[0]
        prj[bad] = np.nan
        want = g['tables'][c].reshape(102, 20, 3)[:, :, :2].reshape(-1, 2)
        assert np.array_equal(np.isnan(prj), np.isnan(want))
        m = ~np.isnan(want)
        assert np.abs(prj[m] - want[m]).max() < 1e-8
        assert np.isnan(g['tables'][c].reshape(102, 20, 3)[:, :, 2]).all()


def test_oracle_manual_adjust_residuals_vs_reference_golden(golden_dir):
    """N3 oracle pin: residual_arr of adjust_extrinsics_manual_points at two parameter points."""
    import os
    import numpy as np
    from oracle import calib_np
    from acinoset_b200 import synth
    g = np.load(os.path.join(golden_dir, 'golden_manual_adjust.npz'))
    case = synth.make_manual_points_case(14, seed=3, rig=synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json')))
    pairs = [[2, 3], [3, 4], [4, 5], [0, 5]]
    pairs = np.unique(pairs, axis=0).tolist()
    for x, want in ((np.zeros(6), g['f0']), (g['probe'], g['f_probe'])):
        f = calib_np.manual_points_residuals(x, case['img_pts'], case['bad_cams'], pairs, case['k'], case['d'], case['r_bad'], case['t_bad'])
        assert f.shape == want.shape
        assert np.abs(f - want).max() <= 1e-9 * max(1.0, np.abs(want).max())


def test_oracle_ring_critical_is_the_threshold_loop():
    """The closed form equals what the reference's loop observes: a point is triangulated at threshold t iff crit > t."""
    import numpy as np
    from oracle import calib_np
    rng = np.random.default_rng(0)
    lik = rng.uniform(size=(200, 6))
    lik[5, 2] = np.nan
    crit = calib_np.ring_critical(lik)
    for t in (0.01, 0.2, 0.5, 0.77):
        m = lik > t
        has_pair = (m & np.roll(m, -1, axis=1)).any(axis=1)
        assert np.array_equal(has_pair, crit > t)


def test_oracle_ekf_measurement_vs_reference_golden(golden_dir):
    """N4 oracle pin: skeleton forward kinematics, per-camera pixels and forward-difference Jacobian against the reference's
    own get_3d_marker_coords / h_function / numerical_jacobian (float64 poses)."""
    import os
    import numpy as np
    from oracle import ekf_np
    from acinoset_b200 import synth
    g = np.load(os.path.join(golden_dir, 'golden_ekf.npz'))
    k, d, r, t, _ = synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json'))
    for mode in ('default', 'head', 'upper_body', 'head_stabilize'):
        x = g[f'{mode}_x']
        assert np.abs(ekf_np.marker_coords(x, mode) - g[f'{mode}_coords']).max() < 1e-13
        h, H = ekf_np.measurement(x, k, d, r, t, mode)
        assert np.abs(h - g[f'{mode}_f64_h']).max() < 1e-8
        assert np.abs(H - g[f'{mode}_f64_H']).max() < 1e-6          # differences of ~1e3 px numbers divided by 1e-3
        # the reference's in-loop float32 pose gives the same Jacobian up to its own rounding noise
        assert np.abs(H - g[f'{mode}_f32_H']).max() < 2.0


def test_block_oracle_equals_dense_oracle():
    """oracle/sba_lm_np.BlockProblem (block-sparse, used at > 1000 points) states the same normal equations, reduced
    camera system and step as the dense Problem."""
    import numpy as np
    from oracle import sba_lm_np as lm, fisheye_np as fe
    from acinoset_b200 import synth
    k, d, r, t, _ = synth.load_rig()
    rng = np.random.default_rng(0)
    P = 14
    Xg = np.stack([rng.uniform(-1, 5, P), rng.uniform(5, 8, P), rng.uniform(0, 1, P)], 1)
    pidx, cidx, obs = [], [], []
    for p in range(P):
        cams = [c for c in range(6) if rng.uniform() < 0.6]
        if len(cams) < 2:
            cams = [0, 1]
        for c in cams:
            pidx.append(p); cidx.append(c)
            obs.append(fe.project_points_fisheye(Xg[p], k[c], d[c], r[c], t[c])[0] + rng.normal(0, 2.0, 2))
    obs, pidx, cidx = np.array(obs), np.array(pidx), np.array(cidx)
    X0 = Xg + rng.normal(0, 0.01, Xg.shape)
    tt = np.asarray(t).reshape(6, 3)
    dense_p = lm.Problem(obs, pidx, cidx, k, d, P, 1.0)
    block_p = lm.BlockProblem(obs, pidx, cidx, k, d, P, 1.0, hess_floor=0.1)
    H, g, F = dense_p.linearise(r, tt, X0)
    L = block_p.linearise(r, tt, X0)
    assert abs(F - L['F']) <= 1e-13 * F
    assert np.abs(dense_p.residuals(r, tt, X0) - block_p.residuals(r, tt, X0)).max() < 1e-10
    for lam in (1e-2, 1.0):
        dc, dp, pred = dense_p.step(H, g, lam)
        s = block_p.step(L, lam)
        S_dense = dense_p.reduced_system(H, g, lam)[0]
        assert np.abs(S_dense - s['S']).max() <= 1e-10 * np.abs(S_dense).max()
        assert np.abs(dc - s['dc']).max() < 1e-11 and np.abs(dp - s['dp'].ravel()).max() < 1e-11
        assert abs(pred - s['pred']) <= 1e-11 * abs(pred)


def test_board_manual_merge_ids_vs_reference_golden(golden_dir):
    """src/lib/sba.py:254-258 through the real _sba_board_points (oracle/make_golden_merge.py records what its two
    prepare_* calls returned and what it passed to the solver): reference_ids=True reproduces the merged arrays bit for
    bit, including the id the first manual point shares with the last board corner; the default keeps them apart."""
    import os
    import numpy as np
    from acinoset_b200 import sba
    g = np.load(os.path.join(golden_dir, 'golden_board_merge.npz'))
    board = tuple(g[f'board_{n}'] for n in ('points_2d', 'points_3d', 'point_indices', 'camera_indices'))
    manual = tuple(g[f'manual_{n}'] for n in ('points_2d', 'points_3d', 'point_indices', 'camera_indices'))
    ref = sba.merge_board_and_manual_points(board, manual, reference_ids=True)
    for a, n in zip(ref, ('points_2d', 'points_3d', 'point_indices', 'camera_indices')):
        assert a.dtype == g[f'merged_{n}'].dtype or n.endswith('indices')
        assert np.array_equal(a, g[f'merged_{n}'])
    assert ref[2][len(board[2])] == board[2].max()                       # the collision of the reference
    own = sba.merge_board_and_manual_points(board, manual)
    assert np.array_equal(own[2][:len(board[2])], board[2])
    assert np.array_equal(own[2][len(board[2]):], manual[2] + board[2].max() + 1)
    assert own[2].max() == len(own[1]) - 1                               # every row of points_3d is a point of the problem
    for a, b in zip(own[:2] + own[3:], ref[:2] + ref[3:]):
        assert np.array_equal(a, b)
