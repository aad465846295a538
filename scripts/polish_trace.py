"""One C4 points+extrinsics solve with ACINO_TRACE=1 (prints where polishing happened) and a timing of K3 polish launches."""
import sys, os
os.environ['ACINO_TRACE'] = '1'
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi
E = _capi.MODE_POINTS_AND_EXTRINSICS
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
sc = synth.make_rig_scene(frames, 20, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
P, C = sc['lik'].shape
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
s = dense.SbaSolver(C, P)
s.set_observations(sc['uv'], mask)
for rep in range(2):
    s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
    st = s.solve(dense.default_opts(E, ftol=1e-10, max_iter=400))
    print(f"solve {st['solve_ms']:.2f} ms {st['iterations']} it {st['launches']} launches cost {st['cost_final']:.6f}", flush=True)
s.close()
