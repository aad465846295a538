"""Sweep of the initial damping and the curvature floor of the points+extrinsics solver (needs a GPU):
C4-sized dense scene and the 8-board calibration golden."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi
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
g = np.load(os.path.join(os.path.dirname(__file__), '..', 'tests', 'golden', 'golden_calib.npz'))
uvc, maskc, slot = sba._lists_to_dense(g['points_2d'].astype(np.float64), g['point_indices'], g['camera_indices'], len(g['points_3d']), 6)
sc2 = dense.SbaSolver(6, len(g['points_3d']))
sc2.set_observations(uvc, maskc)
for hf in (0.3, 0.6, 1.0):
    for lam0 in (1e-4, 1e-2, 1.0, 10.0):
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
        st = s.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, lambda0=lam0, hess_floor=hf, ftol=1e-10, max_iter=300))
        sc2.set_cameras(k, d, g['r0'], g['t0']); sc2.set_points(g['points_3d'].astype(np.float64))
        st2 = sc2.solve(dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, lambda0=lam0, hess_floor=hf, max_iter=300))
        print(f"hfloor {hf} lam0 {lam0:g}: C4 {st['solve_ms']:.1f} ms {st['iterations']} it status {st['status']} cost {st['cost_final']:.9e} | "
              f"calib {st2['solve_ms']:.1f} ms {st2['iterations']} it status {st2['status']} cost {st2['cost_final']:.9e}", flush=True)
