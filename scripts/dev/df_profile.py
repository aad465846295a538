"""Development aid: cProfile of the DataFrame-in / DataFrame-out path at C4 size (what scripts/df_path_c4.py times)."""
import sys, os, time, json, contextlib, io, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
from acinoset_b200 import synth, app, utils, calib, dense, _capi
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
sc = synth.make_rig_scene(frames, 20, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
df = synth.dense_to_df(sc['uv'], sc['lik'], 20)
df = df[df['likelihood'] > 0.5]
scene = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), 'tests', 'golden', 'scene_6cam.json')
utils.get_pairwise_3d_points_from_df(df, k, d.reshape(-1, 4), r, t, calib.triangulate_points_fisheye, verbose=False)
with contextlib.redirect_stdout(io.StringIO()):
    app.sba_points_fisheye(scene, df)
for name, fn in (('tri', lambda: utils.get_pairwise_3d_points_from_df(df, k, d.reshape(-1, 4), r, t, calib.triangulate_points_fisheye, verbose=False)),
                 ('sba', lambda: app.sba_points_fisheye(scene, df))):
    pr = cProfile.Profile()
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        pr.enable(); fn(); pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(28)
    print('=====', name)
    print('\n'.join(l[:170] for l in s.getvalue().splitlines()[4:44]))
