"""GPU tests of the reference-shaped managers: _sba_points (src/lib/sba.py:285-313), the board preparation
(src/lib/sba.py:37-93), _sba_board_points (src/lib/sba.py:211-282) and the app facade (src/lib/app.py:123-138)."""
import json
import os
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from acinoset_b200 import app, calib, sba, synth, utils
from oracle import fisheye_np as fe


def test_sba_points_manager_c1(golden_dir, tmp_path, capsys):
    """core.sba on config C1 through app.sba_points_fisheye: frame layout, index arrays and solution vs the reference."""
    g = np.load(os.path.join(golden_dir, 'golden_c1.npz'))
    sc = synth.make_rig_scene(102, 20, seed=0)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=69)
    df = df.query('likelihood > 0.5')
    df = df[df['frame'].between(69, 170)]
    out_df, res = app.sba_points_fisheye(os.path.join(golden_dir, 'scene_6cam.json'), df)
    out = capsys.readouterr().out
    assert 'bundle_adjust_points_only' in out and 'Before: mean' in out and 'After: mean' in out
    assert list(out_df.columns) == ['frame', 'marker', 'x', 'y', 'z', 'point_index']
    assert (out_df['frame'].to_numpy().astype(np.int64) == g['frame']).all()
    assert (out_df['marker'].to_numpy().astype(str) == g['marker']).all()
    assert (out_df['point_index'].to_numpy() == np.arange(len(g['frame']))).all()
    assert res['before'].shape == g['f_before'].shape
    assert np.abs(res['before'] - g['f_before']).max() < 1e-9 * np.abs(g['f_before']).max()   # same observation order
    assert np.linalg.norm(out_df[['x', 'y', 'z']].to_numpy() - g['obj_pts'], axis=1).max() < 1e-4
    assert abs(np.sqrt(np.mean(res['after'] ** 2)) - np.sqrt(np.mean(g['f_after'] ** 2))) < 1e-4


def test_board_preparation_vs_reference_golden(rig, golden_dir):
    """prepare_calib_board_data_for_bundle_adjustment on the golden calibration scene: same observations and initial
    points as the reference's own prep (image order is arbitrary in the reference — a set — so compare per image)."""
    g = np.load(os.path.join(golden_dir, 'golden_calib.npz'))
    k, d, r, t, _ = rig
    sc = synth.make_board_scene(int(g['n_boards']), seed=int(g['seed']), rig=rig)
    p2, p3, pidx, cidx = sba.prepare_calib_board_data_for_bundle_adjustment(sc['img_pts_arr'], sc['fnames_arr'], sc['board_shape'],
                                                                            k, d, g['r0'], g['t0'], calib.triangulate_points_fisheye)
    assert p2.dtype == np.float32 and p3.dtype == np.float32
    assert p2.shape == g['points_2d'].shape and p3.shape == g['points_3d'].shape
    assert pidx.shape == g['point_indices'].shape and (np.bincount(cidx) == np.bincount(g['camera_indices'])).all()

    def images(p2_, p3_, pidx_, cidx_):
        out = {}
        for img in range(len(p3_) // 54):
            sel = (pidx_ // 54) == img
            key = tuple(np.round(p2_[sel][:54].ravel().astype(np.float64), 3))       # first camera's corners identify the image
            out[key] = (p2_[sel], cidx_[sel], pidx_[sel] % 54, p3_[img * 54:(img + 1) * 54])
        return out
    mine, ref = images(p2, p3, pidx, cidx), images(g['points_2d'], g['points_3d'], g['point_indices'], g['camera_indices'])
    assert set(mine) == set(ref)
    for key in mine:
        a, b = mine[key], ref[key]
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
        assert np.abs(a[3].astype(np.float64) - b[3].astype(np.float64)).max() < 1e-5      # float32 storage of ~5 m coordinates


def test_sba_board_points_manager_files(rig, tmp_path, capsys):
    """notebook cell 30 path: JSON corner files + scene JSON in, refined scene JSON out."""
    k, d, r, t, res = rig
    sc = synth.make_board_scene(6, seed=11, rig=rig)
    scene_in = tmp_path / '6_cam_scene_in.json'
    utils.save_scene(str(scene_in), k, d, sc['r0'], sc['t0'], res)
    fpaths = []
    for c in range(6):
        fp = tmp_path / f'points_{c}.json'
        with open(fp, 'w') as f:
            json.dump({'points': {n: p.tolist() for n, p in zip(sc['fnames_arr'][c], sc['img_pts_arr'][c])},
                       'board_shape': list(sc['board_shape']), 'board_square_len': sc['square'], 'camera_resolution': list(res)}, f)
        fpaths.append(str(fp))
    scene_out = tmp_path / '6_cam_scene_sba.json'
    out = app.sba_board_points_fisheye(str(scene_in), fpaths, str(scene_out))
    assert set(out) == {'before', 'after'} and out['before'].shape == out['after'].shape
    assert fe.cauchy_cost(out['after'], 1.0) < fe.cauchy_cost(out['before'], 1.0)
    assert np.sqrt(np.mean(out['after'] ** 2)) < 0.6          # 0.3 px corner noise
    k2, d2, r2, t2, res2 = utils.load_scene(str(scene_out), verbose=False)
    assert np.array_equal(k2, k) and np.array_equal(d2, d) and tuple(res2) == tuple(res)
    assert r2.shape == (6, 3, 3) and t2.shape == (6, 3, 1)
    for m in r2:
        assert np.abs(m @ m.T - np.eye(3)).max() < 1e-12
    text = capsys.readouterr().out
    assert 'bundle_adjust_board_points_and_extrinsics' in text and 'Saved extrinsics' in text


def test_manual_points_preparation(rig):
    k, d, r, t, _ = rig
    rng = np.random.default_rng(2)
    X = np.stack([rng.uniform(0, 4, 12), rng.uniform(5.5, 7.5, 12), rng.uniform(0.2, 1, 12)], 1)
    pts = np.full((12, 6, 2), np.nan)
    for i in range(12):
        cams = [c for c in range(6) if rng.uniform() < 0.5]
        if i == 0:
            cams = [2]                      # seen once: dropped
        if i == 1:
            cams = []                       # never seen: dropped
        for c in cams:
            pts[i, c] = fe.project_points_fisheye(X[i], k[c], d[c], r[c], t[c])[0]
    p2, p3, pidx, cidx = sba.prepare_manual_points_for_bundle_adjustment(pts, k, d, r, t, calib.triangulate_points_fisheye)
    kept = [i for i in range(12) if (~np.isnan(pts[i, :, 0])).sum() >= 2]
    assert len(p3) == len(kept) and pidx.max() == len(kept) - 1 and len(p2) == len(pidx) == len(cidx)
    assert np.abs(p3 - X[kept]).max() < 1e-4                   # noise-free pixels triangulate to the truth (f32 storage)
    assert p2.dtype == np.float32 and p3.dtype == np.float32


def test_extrinsics_solve_prints_scipy_iteration_table(rig, golden_dir, capsys):
    """The reference calls least_squares(..., verbose=2) (src/lib/sba.py:170) and core/sba.py tees the table into sba.log:
    the same columns come out here, one row per LM iteration, with monotone costs on accepted rows."""
    g = np.load(os.path.join(golden_dir, 'golden_calib.npz'))
    k, d, r, t, _ = rig
    sba.bundle_adjust_points_and_extrinsics(g['points_2d'], g['points_3d'], g['point_indices'], g['camera_indices'], k, d,
                                            g['r0'], g['t0'], calib.project_points_fisheye)
    st = sba.bundle_adjust_points_and_extrinsics.last_stats
    lines = capsys.readouterr().out.splitlines()
    head = [i for i, l in enumerate(lines) if 'Iteration' in l and 'Total nfev' in l and 'Optimality' in l]
    assert len(head) == 1
    rows = []
    for l in lines[head[0] + 1:]:
        parts = l.split()
        if not parts or not parts[0].isdigit():
            break
        rows.append(parts)
    assert len(rows) == st['iterations'] + 1
    assert [int(p[0]) for p in rows] == list(range(st['iterations'] + 1))
    assert abs(float(rows[0][2]) - st['cost_initial']) <= 1e-4 * st['cost_initial']
    assert abs(float(rows[-1][2]) - st['cost_final']) <= 1e-3 * st['cost_final']
    costs = [float(p[2]) for p in rows]
    assert all(b <= a * (1 + 1e-12) for a, b in zip(costs, costs[1:]))
    assert any('termination condition is satisfied' in l for l in lines)
    assert any(l.startswith('Function evaluations') for l in lines)


def test_scene_pipeline_run_many_equals_run():
    """dense.ScenePipeline.run_many (uploads and downloads on copy streams beside the solves, second buffers) returns, scene by scene, bit for bit what run returns - in both modes, for scenes that differ, and run
    still works on the same pipeline afterwards."""
    from acinoset_b200 import _capi, dense, synth
    _capi.require_gpu()
    scs = [synth.make_rig_scene(150, 20, seed=s, masked_frac=0.4) for s in (21, 22, 23)]
    k, d, r, t = scs[0]['k'], scs[0]['d'], scs[0]['r'], scs[0]['t']
    P, C = scs[0]['lik'].shape
    for mode, kw in ((_capi.MODE_POINTS_ONLY, {}), (_capi.MODE_POINTS_AND_EXTRINSICS, dict(ftol=1e-10))):
        pipe = dense.ScenePipeline(P, C, k, d, r, t, mode)
        want = [pipe.run(sc['uv'], sc['lik'], 0.5, **kw) for sc in scs]
        want = [(w['points'].copy(), w['cost_final'], w['iterations']) for w in want]
        got = pipe.run_many([(sc['uv'], sc['lik']) for sc in scs], 0.5, **kw)
        assert len(got) == 3
        for (pw, cw, iw), g in zip(want, got):
            assert np.array_equal(np.isnan(pw), np.isnan(g['points']))
            assert np.array_equal(np.nan_to_num(pw), np.nan_to_num(g['points']))
            assert g['cost_final'] == cw and g['iterations'] == iw
        again = pipe.run(scs[1]['uv'], scs[1]['lik'], 0.5, **kw)
        assert np.array_equal(np.nan_to_num(again['points']), np.nan_to_num(want[1][0]))
        assert pipe.run_many([], 0.5) == []
        pipe.close()
