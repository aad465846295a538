"""Debug aid: per-iteration LM state of the points+extrinsics solver on a dense rig scene (needs a GPU)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 60
fscale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
hfloor = float(sys.argv[4]) if len(sys.argv) > 4 else 0.3
refine = int(sys.argv[5]) if len(sys.argv) > 5 else 0
sc = synth.make_rig_scene(frames, 20, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
P, C = sc['lik'].shape
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
s = dense.SbaSolver(C, P)
s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv)
s.set_observations(sc['uv'], mask); s.set_points(np.nan_to_num(xyz))
o = dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, f_scale=fscale, hess_floor=hfloor, point_refine_iters=refine)
for i in range(iters):
    s.iterate(o)
    st = s.lm_state()
    if True: print(i, f"cost {st['cost']:.10e} lam {st['lam']:.3e} gmax {st['gmax']:.3e} acc {st['accepted']} status {st['status']} Ftrial {st['cost_trial']:.10e} pred {st['pred']:.3e} rho {st['rho']:.3f} dx {st['dxmax']:.2e} chol {st['chol_fail']}")
    if st['status']:
        break
