"""The reference-signature path at dataset scale: app.sba_points_fisheye (DataFrame in -> TRI -> points-only SBA ->
DataFrame out, what core.sba calls, src/core/sba.py:41-43) and utils.get_pairwise_3d_points_from_df (core.tri) on a
119,490-frame table, next to the array path (pipeline / dense.ScenePipeline) on the same scene.
Usage: python scripts/df_path_c4.py [frames]"""
import sys, os, time, json, contextlib, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, app, utils, calib, dense, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
sc = synth.make_rig_scene(frames, 20, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
t0 = time.perf_counter()
df = synth.dense_to_df(sc['uv'], sc['lik'], 20)
df = df[df['likelihood'] > 0.5]
t_df = time.perf_counter() - t0
scene = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'scene_6cam.json')
out = {}
for rep in range(2):
    t0 = time.perf_counter()
    p3 = utils.get_pairwise_3d_points_from_df(df, k, d.reshape(-1, 4), r, t, calib.triangulate_points_fisheye, verbose=False)
    out['tri_df_seconds'] = time.perf_counter() - t0
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        new_df, res = app.sba_points_fisheye(scene, df)
    out['sba_df_seconds'] = time.perf_counter() - t0
t0 = time.perf_counter()
utils.df_to_dense(df, 6)
out['df_to_dense_seconds'] = time.perf_counter() - t0
from acinoset_b200 import sba as sba_mod
st = sba_mod.bundle_adjust_points_only.last_stats
pipe = dense.ScenePipeline(sc['lik'].shape[0], 6, k, d, r, t, _capi.MODE_POINTS_ONLY)
for rep in range(2):
    t0 = time.perf_counter()
    s2 = pipe.run(sc['uv'], sc['lik'], 0.5)
    out['array_path_seconds'] = time.perf_counter() - t0
pipe.close()
out.update(frames=frames, rows_2d=int(len(df)), points=int(len(new_df)), device_solve_ms=st['solve_ms'], make_df_seconds=t_df,
           array_device_solve_ms=s2['solve_ms'], rmse_after=float(np.sqrt(np.mean(res['after'] ** 2))))
print(json.dumps(out))
