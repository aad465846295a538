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
