"""GPU tests of the rows SURVEY.md §8(f) marks next: N1 (array-level TRI / SBA pipelines without pandas joins) and N2
(metric.residual_error batched on K1), each against the reference."""
import os
import numpy as np
import pandas as pd
import pytest

pytestmark = pytest.mark.gpu

from acinoset_b200 import app, calib, metric, pipeline, synth, utils


def _c1():
    sc = synth.make_rig_scene(102, 20, seed=0)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=69)
    return sc, df


def test_metric_residual_error_vs_reference_golden(golden_dir):
    """a12 / N2: same tables as the reference's metric.residual_error on config C1 (golden from the real reference)."""
    g = np.load(os.path.join(golden_dir, 'golden_metric.npz'))
    c1 = np.load(os.path.join(golden_dir, 'golden_c1.npz'))
    sc, df = _c1()
    df_t = df.query('likelihood > 0.5')
    p3 = pd.DataFrame({'frame': c1['frame'], 'marker': c1['marker'], 'x': c1['points_3d'][:, 0], 'y': c1['points_3d'][:, 1],
                       'z': c1['points_3d'][:, 2]})
    err = metric.residual_error(df_t, p3, synth.MARKERS[:20], (sc['k'], sc['d'], sc['r'], sc['t'], None, None))
    assert sorted(err) == [str(i) for i in range(6)]
    for c in range(6):
        e = err[str(c)]
        assert list(e.columns) == metric.COLUMNS
        assert np.array_equal(e['frame'].to_numpy().astype(np.float64), g[f'cam{c}_frame'])
        assert np.array_equal(e['marker'].to_numpy().astype(str), g[f'cam{c}_marker'])
        for col, tol in (('camera_distance', 1e-12), ('pixel_residual', 1e-9), ('error_u', 1e-9), ('error_v', 1e-9)):
            a, b = e[col].to_numpy().astype(np.float64), g[f'cam{c}_{col}']
            assert np.abs(a - b).max() <= tol * max(1.0, np.abs(b).max()), (c, col)
        a, b = e['pck_threshold'].to_numpy().astype(np.float64), g[f'cam{c}_pck_threshold']
        assert np.array_equal(np.isnan(a), np.isnan(b)) and np.allclose(a[~np.isnan(a)], b[~np.isnan(b)], rtol=0, atol=1e-12)


def test_dense_pipeline_equals_dataframe_pipeline(golden_dir):
    """N1: the array-level TRI / SBA give exactly the positions array core.tri / core.sba would save."""
    sc, df = _c1()
    cam = (sc['k'], sc['d'], sc['r'], sc['t'])
    markers = synth.MARKERS[:20]
    start, end = 75, 160
    # reference-shaped path: threshold + window + DataFrame TRI + scatter (src/core/tri.py:41-60)
    d = df.query('likelihood > 0.5')
    d = d[d['frame'].between(start, end)]
    p3 = utils.get_pairwise_3d_points_from_df(d, sc['k'], sc['d'].reshape(-1, 4), sc['r'], sc['t'], calib.triangulate_points_fisheye,
                                              verbose=False)
    want = np.full((end - start + 1, len(markers), 3), np.nan)
    for i, m in enumerate(markers):
        rows = p3[p3['marker'] == m][['frame', 'x', 'y', 'z']].values
        for frame, *pt in rows:
            want[int(frame) - start, i] = pt
    uv, lik = pipeline.dlc_df_to_dense(df, 6, markers, start, end)
    got = pipeline.tri(uv, lik, len(markers), 0.5, cam)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.array_equal(got[~np.isnan(got)], want[~np.isnan(want)])          # same kernel, same inputs: bit-equal
    # the .h5-like array form lowers to the same slots
    per_cam = [np.concatenate([sc['uv'][:, c, :], sc['lik'][:, c:c + 1]], axis=1).reshape(102, 20, 3)[start - 69:end - 69 + 1] for c in range(6)]
    uv2, lik2 = pipeline.dlc_arrays_to_dense(per_cam)
    assert np.array_equal(uv2, uv) and np.array_equal(lik2, lik)
    # SBA: against the DataFrame manager
    import json
    pos, st = pipeline.sba(uv, lik, len(markers), 0.5, cam)
    out_df, res = app.sba_points_fisheye(os.path.join(golden_dir, 'scene_6cam.json'), d)
    want2 = np.full_like(want, np.nan)
    for i, m in enumerate(markers):
        rows = out_df[out_df['marker'] == m][['frame', 'x', 'y', 'z']].values
        for frame, *pt in rows:
            want2[int(frame) - start, i] = pt
    assert np.array_equal(np.isnan(pos), np.isnan(want2))
    assert np.array_equal(pos[~np.isnan(pos)], want2[~np.isnan(want2)])
    assert st['n_unconverged'] == 0


def test_project_all_cameras_vs_reference_golden(golden_dir, tmp_path):
    """N2: save_3d_cheetah_as_2d (src/lib/utils.py:237-283) — all cameras in one launch, reference range rule on the device."""
    g = np.load(os.path.join(golden_dir, 'golden_project2d.npz'))
    sc = synth.make_rig_scene(2, 2, seed=0)
    markers = synth.MARKERS[:20]
    k, d, r, t, res = utils.load_scene(os.path.join(golden_dir, 'scene_6cam.json'), verbose=False)
    dfs = utils.project_3d_cheetah_to_2d(g['positions'], k, d, r, t, res, markers, 69)
    assert len(dfs) == 6
    for c in range(6):
        a, b = dfs[c].to_numpy(), g['tables'][c]
        assert np.array_equal(np.isnan(a), np.isnan(b)), c
        m = ~np.isnan(b)
        assert np.abs(a[m] - b[m]).max() < 1e-9 * 2704
        assert list(dfs[c].index) == list(g['index'])
        assert list(dfs[c].columns.get_level_values(0)[::3]) == markers
    # a per-camera list of position arrays (utils.py:249-250) gives the same tables
    dfs2 = utils.project_3d_cheetah_to_2d([g['positions']] * 6, k, d, r, t, res, markers, 69)
    for c in range(6):
        assert np.array_equal(dfs2[c].to_numpy(), dfs[c].to_numpy(), equal_nan=True)
    # file-level drop-in: same directory contract, csv written
    data_dir = tmp_path / 'run'; out_dir = data_dir / 'sba'
    (data_dir / 'extrinsic_calib').mkdir(parents=True); out_dir.mkdir()
    scene = data_dir / 'extrinsic_calib' / '6_cam_scene_sba.json'
    scene.write_text(open(os.path.join(golden_dir, 'scene_6cam.json')).read())
    for c in range(6):
        (data_dir / f'cam{c + 1}.mp4').write_text('')
    out = utils.save_3d_cheetah_as_2d(g['positions'], str(out_dir), str(scene), markers, calib.project_points_fisheye, 69)
    assert len(out) == 6 and (out_dir / 'cam3_sba.csv').exists()
    back = pd.read_csv(out_dir / 'cam3_sba.csv', header=[0, 1], index_col=0)
    assert np.allclose(back.to_numpy(), g['tables'][2], rtol=0, atol=1e-6, equal_nan=True)
    # no videos -> empty list, as the reference
    assert utils.save_3d_cheetah_as_2d(g['positions'], str(tmp_path / 'run2' / 'sba'), str(tmp_path / 'run2' / 'x' / 's.json'), markers,
                                       calib.project_points_fisheye, 69) == []
    # plain projection of every camera equals the per-camera call bit for bit
    X = np.nan_to_num(g['positions']).reshape(-1, 3)
    all_ = calib.project_points_fisheye_all(X, k, d, r, t)
    for c in range(6):
        assert np.array_equal(all_[c], calib.project_points_fisheye(X, k[c], d[c], r[c], t[c]), equal_nan=True)


def test_pck_matches_reference_formula(golden_dir):
    g = np.load(os.path.join(golden_dir, 'golden_metric.npz'))
    errs = {str(c): pd.DataFrame({'pixel_residual': g[f'cam{c}_pixel_residual'].astype(str), 'pck_threshold': g[f'cam{c}_pck_threshold'].astype(str)})
            for c in range(6)}
    dist = np.concatenate([g[f'cam{c}_pixel_residual'] for c in range(6)])
    thr = np.concatenate([g[f'cam{c}_pck_threshold'] for c in range(6)])
    want = 100.0 * (np.sum(dist <= 0.5 * thr) / len(dist))           # src/core/metrics.py:45
    assert metric.pck(errs) == want
    assert np.isnan(metric.pck({}))


def test_manual_points_residuals_and_adjust_vs_reference_golden(golden_dir):
    """N3: adjust_extrinsics_manual_points (src/lib/calib.py:216-297): residual vector 1e-9 against the reference's closure,
    batch == single evaluations, and the corrected extrinsics against the reference's own scipy run."""
    from oracle import calib_np
    g = np.load(os.path.join(golden_dir, 'golden_manual_adjust.npz'))
    case = synth.make_manual_points_case(14, seed=3, rig=synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json')))
    pairs = [[0, 5], [2, 3], [3, 4], [4, 5]]
    args = (case['img_pts'], case['bad_cams'], pairs, case['k'], case['d'], case['r_bad'], case['t_bad'])
    f0 = calib.manual_points_residuals(*args, np.zeros(6))
    fp = calib.manual_points_residuals(*args, g['probe'])
    assert np.abs(f0 - g['f0']).max() <= 1e-9 * np.abs(g['f0']).max()
    assert np.abs(fp - g['f_probe']).max() <= 1e-9 * np.abs(g['f_probe']).max()
    fo = calib_np.manual_points_residuals(g['probe'], case['img_pts'], case['bad_cams'], pairs, case['k'], case['d'], case['r_bad'], case['t_bad'])
    assert np.abs(fp - fo).max() <= 1e-9 * np.abs(fo).max()
    batch = calib.manual_points_residuals(*args, np.stack([np.zeros(6), g['probe'], 0.5 * g['probe']]))
    assert batch.shape == (3, len(f0)) and np.array_equal(batch[0], f0) and np.array_equal(batch[1], fp)
    # full drop-in call, in place on copies like the reference
    r_in, t_in = case['r_bad'].copy(), case['t_bad'].copy()
    r_out, t_out = calib.adjust_extrinsics_manual_points(calib.calibrate_pair_extrinsics_fisheye, case['img_pts'], case['bad_cams'],
                                                         case['k'], case['d'], r_in, t_in)
    assert r_out is r_in and t_out is t_in
    res = calib.adjust_extrinsics_manual_points.last_result
    assert abs(res.cost - float(g['cost'])) <= 1e-8 * float(g['cost'])
    assert np.abs(res.x - g['x']).max() < 1e-6
    assert np.abs(r_out - g['r_out']).max() < 1e-6 and np.abs(t_out - g['t_out']).max() < 1e-6
    assert np.abs(r_out - case['r_true']).max() < np.abs(case['r_bad'] - case['r_true']).max() / 5
    for c in (0, 1, 2):
        assert np.array_equal(r_out[c], case['r_bad'][c]) and np.array_equal(t_out[c], case['t_bad'][c])
    with pytest.raises(NotImplementedError):
        calib.adjust_extrinsics_manual_points(lambda: None, case['img_pts'], 3, case['k'], case['d'], r_in, t_in)
    # a single integer index is accepted like a list of one
    r2, t2 = calib.adjust_extrinsics_manual_points(calib.calibrate_pair_extrinsics_fisheye, case['img_pts'], 3, case['k'], case['d'],
                                                   case['r_bad'].copy(), case['t_bad'].copy())
    assert r2.shape == (6, 3, 3)


def test_initialize_marker_3d_vs_reference_golden(golden_dir, capsys):
    """N3: initialize_marker_3d (src/lib/app.py:39-84): same interim threshold (one kernel instead of ~20 re-triangulations
    of the whole table) and the same fitted trajectory as the reference function."""
    from oracle import calib_np
    g = np.load(os.path.join(golden_dir, 'golden_init_marker.npz'))
    sc = synth.make_track_scene(60, 20, seed=2, rig=synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json')))
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=10)
    assert np.array_equal(app.ring_critical_likelihood(sc['lik']), calib_np.ring_critical(sc['lik']))
    lik = sc['lik'][:50].copy(); lik[3, 1] = np.nan; lik[4] = 0.0
    assert np.array_equal(app.ring_critical_likelihood(lik), calib_np.ring_critical(lik))
    for marker in ('nose', 'spine'):
        fit, fit_deriv = app.initialize_marker_3d(df, marker, sc['k'], sc['d'].reshape((-1, 4)), sc['r'], sc['t'])
        line = [l for l in capsys.readouterr().out.splitlines() if l.startswith('Initializing')][0]
        assert line == str(g[f'{marker}_thresh_line'])
        assert fit.shape == (60, 3) and fit_deriv.shape == (60, 3)
        assert np.abs(fit - g[f'{marker}_fit']).max() < 1e-6
        assert np.abs(fit_deriv - g[f'{marker}_fit_deriv']).max() < 1e-6


def test_ekf_measurement_vs_reference_golden(golden_dir):
    """N4: EKF measurement h and numerical Jacobian H (src/core/ekf.py:66-96,245-251) for all cameras and perturbed poses
    in two launches, against the reference's own functions (golden) and the numpy oracle; all four marker modes."""
    from acinoset_b200 import ekf
    from oracle import ekf_np
    g = np.load(os.path.join(golden_dir, 'golden_ekf.npz'))
    k, d, r, t, _ = synth.load_rig(os.path.join(golden_dir, 'scene_6cam.json'))
    for mode in ('default', 'head', 'upper_body', 'head_stabilize'):
        x = g[f'{mode}_x']
        h, H, coords = ekf.measurement(x, k, d, r, t, mode, with_coords=True)
        assert H.shape == g[f'{mode}_f64_H'].shape and h.shape == g[f'{mode}_f64_h'].shape
        assert np.abs(coords - g[f'{mode}_coords']).max() < 1e-13
        assert np.abs(h - g[f'{mode}_f64_h']).max() < 1e-8
        assert np.abs(H - g[f'{mode}_f64_H']).max() < 1e-6
        assert not H[:, len(x):].any()                                 # velocity / acceleration columns stay zero
        ho, Ho = ekf_np.measurement(x, k, d, r, t, mode)
        assert np.abs(H - Ho).max() < 1e-6 and np.abs(h - ho).max() < 1e-9
        # single-camera helpers of the reference's interface
        assert np.allclose(ekf.h_function(x, k[2], d[2], r[2], t[2], mode).ravel(), h.reshape(6, -1)[2], rtol=0, atol=1e-9)
        assert np.array_equal(ekf.get_3d_marker_coords({'x': x}, mode), coords)
    with pytest.raises(ValueError):
        ekf.measurement(np.zeros(5), k, d, r, t, 'head')
    with pytest.raises(ValueError):
        ekf.measurement(np.zeros(6), k, d, r, t, 'tail')
