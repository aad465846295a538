"""TEST INFRASTRUCTURE ONLY — generates tests/golden/*.npz by running the REAL reference
(/root/reference/src/lib/{calib,utils,sba}.py, unmodified, loaded through oracle/ref_shim.py)
on seeded synthetic inputs. Run in the build container only:

    python -m oracle.make_golden [camera|c1|c5small|calib|metric|project2d|all]

Versions of record for the un-pinned third-party numerics: see the 'versions' entry stored in
every fixture (OpenCV / scipy / numpy / pandas of the build image).
"""
import json
import os
import sys
import time
import numpy as np

from . import ref_shim

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, 'tests', 'golden')


def _versions():
    import cv2, scipy, pandas
    return json.dumps(dict(cv2=cv2.__version__, scipy=scipy.__version__, numpy=np.__version__, pandas=pandas.__version__))


def gen_scene():
    calib, utils, sba = ref_shim.load()
    with ref_shim.quiet():
        k, d, r, t, res = utils.load_scene(ref_shim.REF_SCENE)
    data = {'timestamp': 'golden copy of the reference rig (configs/dummy_scene.json) via load_scene',
            'camera_resolution': list(res),
            'cameras': [{'k': k[i].tolist(), 'd': d[i].tolist(), 'r': r[i].tolist(), 't': t[i].tolist()} for i in range(len(k))]}
    with open(os.path.join(GOLD, 'scene_6cam.json'), 'w') as f:
        json.dump(data, f)


def gen_camera():
    """a1 / a7 / Rodrigues known-answer vectors straight from the reference callables."""
    import cv2
    calib, utils, sba = ref_shim.load()
    with ref_shim.quiet():
        k, d, r, t, _ = utils.load_scene(ref_shim.REF_SCENE)
    rng = np.random.default_rng(11)
    N = 2000
    X = np.stack([rng.uniform(-1, 5, N), rng.uniform(5, 8, N), rng.uniform(0, 1, N)], axis=1)
    uv = np.stack([calib.project_points_fisheye(X, k[c], d[c], r[c], t[c]) for c in range(6)])
    # wide field: points well off-axis and close to the image plane of each camera
    Xw = np.stack([rng.uniform(-8, 12, N), rng.uniform(-2, 15, N), rng.uniform(-1, 3, N)], axis=1)
    uvw = np.stack([calib.project_points_fisheye(Xw, k[c], d[c], r[c], t[c]) for c in range(6)])
    # two-view triangulation per ring pair on noisy pixels
    noisy = uv + rng.normal(0, 1.0, uv.shape)
    tri = np.stack([calib.triangulate_points_fisheye(noisy[c], noisy[(c + 1) % 6], k[c], d[c], r[c], t[c],
                                                     k[(c + 1) % 6], d[(c + 1) % 6], r[(c + 1) % 6], t[(c + 1) % 6])
                    for c in range(6)])
    und = np.stack([cv2.fisheye.undistortPoints(noisy[c].reshape(-1, 1, 2), k[c], d[c]).reshape(-1, 2) for c in range(6)])
    # undistort edge cases: principal point, far corners, way outside the image
    edge = np.array([[k[0][0, 2], k[0][1, 2]], [0, 0], [2703, 1519], [-5000, 4000], [9000, -3000], [1345.7, 772.9],
                     [k[0][0, 2] + 1e-7, k[0][1, 2]], [40000.0, 40000.0]])
    und_edge = cv2.fisheye.undistortPoints(edge.reshape(-1, 1, 2), k[0], d[0]).reshape(-1, 2)
    rv = rng.normal(0, 1.0, (64, 3))
    rv[0] = 0; rv[1] = [1e-20, 0, 0]; rv[2] = [1e-9, -1e-9, 2e-9]; rv[3] = [np.pi, 0, 0]; rv[4] = [0, 3.0, 0.2]
    rmat = np.stack([cv2.Rodrigues(v)[0] for v in rv])
    rv_back = np.stack([cv2.Rodrigues(m)[0].ravel() for m in rmat])
    rv_scene = np.stack([cv2.Rodrigues(m)[0].ravel() for m in r])
    np.savez_compressed(os.path.join(GOLD, 'golden_camera.npz'), X=X, uv=uv, Xw=Xw, uvw=uvw, noisy=noisy, tri=tri, und=und,
                        edge=edge, und_edge=und_edge, rv=rv, rmat=rmat, rv_back=rv_back, rv_scene=rv_scene,
                        versions=_versions())


def c1_inputs():
    from acinoset_b200 import synth
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    sc = synth.make_rig_scene(102, 20, seed=0, rig=rig)       # frames 69..170 inclusive (all_optimizations.py:115-116)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=69)
    return sc, df


def gen_c1():
    """Config C1: TRI + points-only SBA through the reference's own functions (src/lib/sba.py:285-313,
    with the pandas-3 float cast of SURVEY §8c)."""
    calib, utils, sba = ref_shim.load()
    sc, df = c1_inputs()
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    df_t = df.query('likelihood > 0.5')                        # src/core/sba.py:41
    df_t = df_t[df_t['frame'].between(69, 170)]                # src/core/sba.py:42
    t0 = time.time()
    with ref_shim.quiet():
        p3 = utils.get_pairwise_3d_points_from_df(df_t, k, d.reshape((-1, 4)), r, t, calib.triangulate_points_fisheye)
    t_tri = time.time() - t0
    p3[['x', 'y', 'z']] = p3[['x', 'y', 'z']].astype(np.float64)
    p3['point_index'] = p3.index                               # sba.py:296
    points_3d = np.array(p3[['x', 'y', 'z']], dtype=np.float64)
    pdf = df_t.merge(p3, how='inner', on=['frame', 'marker'], suffixes=('_cam', ''))   # sba.py:299
    points_2d = np.array(pdf[['x_cam', 'y_cam']], dtype=np.float64)
    pidx = np.array(pdf['point_index']).astype(np.int64)
    cidx = np.array(pdf['camera']).astype(np.int64)
    t0 = time.time()
    with ref_shim.quiet():
        obj, res = sba.bundle_adjust_points_only(points_2d, points_3d, pidx, cidx, k, d, r, t,
                                                 calib.project_points_fisheye, f_scale=50)
    t_sba = time.time() - t0
    rng = np.random.default_rng(5)
    xp = points_3d.ravel() + rng.normal(0, 0.01, points_3d.size)
    with ref_shim.quiet():
        f_pert = sba.cost_func_points_only(xp, len(points_3d), pidx, cidx, k, d, r, t, points_2d, calib.project_points_fisheye)
    np.savez_compressed(os.path.join(GOLD, 'golden_c1.npz'),
                        frame=np.array(p3['frame']).astype(np.int64), marker=np.array(p3['marker']).astype(str),
                        points_3d=points_3d, points_2d=points_2d, point_indices=pidx, camera_indices=cidx,
                        f_before=res['before'], f_after=res['after'], obj_pts=obj, x_pert=xp, f_pert=f_pert,
                        t_tri=t_tri, t_sba=t_sba, versions=_versions())
    print(f'c1: {len(points_3d)} points, {len(pidx)} obs, tri {t_tri:.2f}s sba {t_sba:.2f}s')


def c5small_inputs():
    from acinoset_b200 import synth
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    # config C5 in small: 30 % of the slots below the threshold, 5 % of the observations with N(0, 50 px)
    # outlier noise on top of N(0, 1 px), seed 5
    sc = synth.make_rig_scene(60, 20, seed=5, masked_frac=0.3, outlier_frac=0.05, rig=rig)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=0)
    return sc, df


def gen_c5small():
    """Config C5 in small: TRI + points-only SBA through the reference's own functions on a scene with gross outliers,
    where the Cauchy loss (f_scale 50) actually bends: pins the robust objective, not only the least-squares one."""
    calib, utils, sba = ref_shim.load()
    sc, df = c5small_inputs()
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    df_t = df.query('likelihood > 0.5')
    t0 = time.time()
    with ref_shim.quiet():
        p3 = utils.get_pairwise_3d_points_from_df(df_t, k, d.reshape((-1, 4)), r, t, calib.triangulate_points_fisheye)
    t_tri = time.time() - t0
    p3[['x', 'y', 'z']] = p3[['x', 'y', 'z']].astype(np.float64)
    p3['point_index'] = p3.index
    points_3d = np.array(p3[['x', 'y', 'z']], dtype=np.float64)
    pdf = df_t.merge(p3, how='inner', on=['frame', 'marker'], suffixes=('_cam', ''))
    points_2d = np.array(pdf[['x_cam', 'y_cam']], dtype=np.float64)
    pidx = np.array(pdf['point_index']).astype(np.int64)
    cidx = np.array(pdf['camera']).astype(np.int64)
    t0 = time.time()
    with ref_shim.quiet():
        obj, res = sba.bundle_adjust_points_only(points_2d, points_3d, pidx, cidx, k, d, r, t,
                                                 calib.project_points_fisheye, f_scale=50)
    t_sba = time.time() - t0
    np.savez_compressed(os.path.join(GOLD, 'golden_c5small.npz'),
                        frame=np.array(p3['frame']).astype(np.int64), marker=np.array(p3['marker']).astype(str),
                        points_3d=points_3d, points_2d=points_2d, point_indices=pidx, camera_indices=cidx,
                        f_before=res['before'], f_after=res['after'], obj_pts=obj, t_tri=t_tri, t_sba=t_sba, versions=_versions())
    print(f'c5small: {len(points_3d)} points, {len(pidx)} obs, tri {t_tri:.2f}s sba {t_sba:.2f}s')


def gen_metric():
    """a12: the reference's own metric.residual_error (src/lib/metric.py:36-101) on config C1 (thresholded 2-D frame vs the
    reference's triangulated points)."""
    calib, utils, sba = ref_shim.load()
    from lib import metric
    sc, df = c1_inputs()
    g = np.load(os.path.join(GOLD, 'golden_c1.npz'))
    import pandas as pd
    df_t = df.query('likelihood > 0.5')
    p3 = pd.DataFrame({'frame': g['frame'], 'marker': g['marker'], 'x': g['points_3d'][:, 0], 'y': g['points_3d'][:, 1],
                       'z': g['points_3d'][:, 2]})
    from acinoset_b200 import synth
    markers = synth.MARKERS[:20]
    cam = (sc['k'], sc['d'], sc['r'], sc['t'], None, None)
    with ref_shim.quiet():
        err = metric.residual_error(df_t, p3, markers, cam)
    out = {}
    for c, e in err.items():
        out[f'cam{c}_frame'] = e['frame'].to_numpy().astype(np.float64)
        out[f'cam{c}_marker'] = e['marker'].to_numpy().astype(str)
        for col in ('camera_distance', 'pixel_residual', 'pck_threshold', 'error_u', 'error_v'):
            out[f'cam{c}_{col}'] = e[col].to_numpy().astype(np.float64)
    np.savez_compressed(os.path.join(GOLD, 'golden_metric.npz'), versions=_versions(), **out)
    print('metric:', {c: len(e) for c, e in err.items()})


def gen_project2d():
    """N2: the reference's save_3d_cheetah_as_2d (src/lib/utils.py:237-283) on the C1 positions array. PyTables is not in the
    image, so DataFrame.to_hdf is stubbed for the call (harness shim, like np.int): the projection, range rule and table
    layout are the reference's own code."""
    import tempfile
    import pandas as pd
    calib, utils, sba = ref_shim.load()
    g = np.load(os.path.join(GOLD, 'golden_c1.npz'))
    from acinoset_b200 import synth
    markers = synth.MARKERS[:20]
    pos = np.full((102, 20, 3), np.nan)
    rank = {m: i for i, m in enumerate(markers)}
    for f, m, p in zip(g['frame'], g['marker'], g['obj_pts']):
        pos[int(f) - 69, rank[str(m)]] = p
    # some rows well outside every image, to exercise the range rule
    pos[3, 4] = [40.0, 6.0, 0.5]
    pos[7, 2] = [-30.0, 7.0, 0.2]
    saved = pd.DataFrame.to_hdf
    pd.DataFrame.to_hdf = lambda self, *a, **k: None
    try:
        with tempfile.TemporaryDirectory() as td:
            data_dir = os.path.join(td, 'run'); out_dir = os.path.join(data_dir, 'sba')
            os.makedirs(os.path.join(data_dir, 'extrinsic_calib')); os.makedirs(out_dir)
            scene = os.path.join(data_dir, 'extrinsic_calib', '6_cam_scene_sba.json')
            with open(os.path.join(GOLD, 'scene_6cam.json')) as fi, open(scene, 'w') as fo:
                fo.write(fi.read())
            for c in range(6):
                open(os.path.join(data_dir, f'cam{c + 1}.mp4'), 'w').close()
            with ref_shim.quiet():
                dfs = utils.save_3d_cheetah_as_2d(pos, out_dir, scene, markers, calib.project_points_fisheye, 69, save_as_csv=False)
    finally:
        pd.DataFrame.to_hdf = saved
    assert len(dfs) == 6
    np.savez_compressed(os.path.join(GOLD, 'golden_project2d.npz'), positions=pos, tables=np.stack([d.to_numpy() for d in dfs]),
                        index=np.array(dfs[0].index), versions=_versions())
    print('project2d:', np.stack([d.to_numpy() for d in dfs]).shape, 'nan rows', int(np.isnan(dfs[0].to_numpy()[:, 0::3]).sum()))


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else 'all'
    os.makedirs(GOLD, exist_ok=True)
    if not ref_shim.available():
        raise SystemExit('reference tree not present; fixtures can only be generated in the build container')
    gen_scene()
    if what in ('camera', 'all'):
        gen_camera()
    if what in ('c1', 'all'):
        gen_c1()
    if what in ('c5small', 'all'):
        gen_c5small()
    if what in ('calib', 'all'):
        from .make_golden_calib import gen_calib
        gen_calib()
    if what in ('metric', 'all'):
        gen_metric()
    if what in ('project2d', 'all'):
        gen_project2d()


if __name__ == '__main__':
    main()
