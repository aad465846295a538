"""Small workload for compute-sanitizer (memcheck / racecheck): config C1 (TRI + points-only SBA, 102 frames), the 8-board
calibration scene of tests/golden/golden_calib.npz (points+extrinsics SBA: K4a, K4b with its TMA tile loads and
mbarriers, K5, K6, K7, polishing) on a 60-frame rig scene too (all mask classes), K1 for all cameras, and the
allocation-free entry point. Usage: compute-sanitizer --tool memcheck python scripts/sanitize_target.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as ct
import numpy as np
from acinoset_b200 import synth, dense, sba, calib, _capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rig = synth.load_rig()
k, d, r, t, _ = rig
sc = synth.make_rig_scene(102, 20, seed=0)
st = dense.solve_scene_host(sc['uv'], sc['lik'], 0.5, k, d, r, t)
print('C1 points-only', st['iterations'], st['cost_final'])
g = np.load(os.path.join(ROOT, 'tests', 'golden', 'golden_calib.npz'))
obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(g['points_2d'], g['points_3d'], g['point_indices'], g['camera_indices'], k, d,
                                                           g['r0'], g['t0'], calib.project_points_fisheye, verbose=0)
print('calibration', sba.bundle_adjust_points_and_extrinsics.last_stats['iterations'])
sc = synth.make_rig_scene(60, 20, seed=3, masked_frac=0.3, outlier_frac=0.05)
rng = np.random.default_rng(1)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (6, 3))
tv = np.asarray(t).reshape(6, 3) + rng.normal(0, 3e-3, (6, 3))
st = dense.solve_scene_host(sc['uv'], sc['lik'], 0.5, k, d, r, t, mode=_capi.MODE_POINTS_AND_EXTRINSICS, rvec_t0=(rv, tv))
print('rig scene points+extrinsics', st['iterations'], st['status'])
X = np.nan_to_num(st['points'])
cams = _capi.pack_cams(k, d, r, t)
uv = np.empty((6, len(X), 2))
_capi.check(_capi.lib().acino_project_all(_capi.ptr(X), len(X), _capi.ptr(cams), 6, 2704.0, 1520.0, _capi.ptr(uv)))
n = 3000
ok_pts = np.flatnonzero(np.isfinite(st['points']).all(1))
pidx = rng.choice(ok_pts, n).astype(np.int32); cidx = rng.integers(0, 6, n).astype(np.int32); obs = rng.uniform(0, 1000, (n, 2))
Xd, pd_, cd, od = (dense.DeviceArray.from_host(a) for a in (X, pidx, cidx, obs))
fd, cost = dense.DeviceArray((n, 2), np.float64), dense.DeviceArray((1,), np.float64)
for _ in range(2):
    _capi.check(_capi.lib().acino_residual_cost_noalloc(Xd.ptr, len(X), pd_.ptr, cd.ptr, od.ptr, n, _capi.ptr(cams), 6, 50.0, fd.ptr, cost.ptr, None))
_capi.check(_capi.lib().acino_sync())
print('noalloc cost', float(cost.download()[0]))
print('done')
