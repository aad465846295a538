"""Small fixed workload for ncu: C4-sized scene (119,490 frames), TRI x2, K1 once, n_fixed + n_refine Schur LM iterations
(plain linearisation / embedded refinement), 2 points-only LM iterations.
Usage: python scripts/profile_target.py [frames] [n_fixed] [n_refine]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as ct
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
n_fixed = int(sys.argv[2]) if len(sys.argv) > 2 else 4
n_refine = int(sys.argv[3]) if len(sys.argv) > 3 else 4
sc = synth.make_rig_scene(frames, 20, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
P, C = sc['lik'].shape
L = _capi.lib()
cams = _capi.pack_cams(k, d, r, t)
uv_d = dense.DeviceArray.from_host(sc['uv']); lik_d = dense.DeviceArray.from_host(sc['lik'])
xyz_d = dense.DeviceArray((P, 3), np.float64); mask_d = dense.DeviceArray((P,), np.uint8); ex_d = dense.DeviceArray((P,), np.uint8)
s = dense.SbaSolver(C, P)
for _ in range(2):
    _capi.check(L.acino_triangulate_ring_dev(uv_d.ptr, lik_d.ptr, 0.5, P, C, _capi.ptr(cams), xyz_d.ptr, mask_d.ptr, ex_d.ptr,
                                             ct.c_void_p(s.stream)))
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv)
s.set_observations_dev(uv_d, mask_d)
s.set_points(np.nan_to_num(xyz_d.download()))
o = dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, force_iters=1)
f = s.residuals()       # K1 (residuals_dense_kernel) once
o_fixed = dense.default_opts(_capi.MODE_POINTS_AND_EXTRINSICS, force_iters=1, point_refine_iters=0)
for _ in range(n_fixed):      # the bench's fixed-work step: plain linearisation (schur_linearise_kernel<C, false>)
    s.iterate(o_fixed)
for _ in range(n_refine):      # a solve's step: embedded refinement on (schur_linearise_kernel<C, true>)
    s.iterate(o)
print('schur', s.lm_state())
s.reset_state()
o = dense.default_opts(_capi.MODE_POINTS_ONLY, force_iters=1, inner_iters=1)
for _ in range(2):
    s.iterate(o)
_capi.check(L.acino_sync())
print('done')
