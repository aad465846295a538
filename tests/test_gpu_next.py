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
