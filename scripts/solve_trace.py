"""Development aid: one full points+extrinsics solve of the C4-sized scene per library build, three times each (the first
call of a process pays one-off costs), with ACINO_TRACE lines of the last one. Usage: python scripts/solve_trace.py"""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as ct
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
sc = synth.make_rig_scene(frames, 20, seed=4, masked_frac=0.5)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
P, C = sc['lik'].shape
L = _capi.lib()
cams = _capi.pack_cams(k, d, r, t)
uv_d = dense.DeviceArray.from_host(sc['uv']); lik_d = dense.DeviceArray.from_host(sc['lik'])
xyz_d = dense.DeviceArray((P, 3), np.float64); mask_d = dense.DeviceArray((P,), np.uint8); ex_d = dense.DeviceArray((P,), np.uint8)
s = dense.SbaSolver(C, P)
_capi.check(L.acino_triangulate_ring_dev(uv_d.ptr, lik_d.ptr, 0.5, P, C, _capi.ptr(cams), xyz_d.ptr, mask_d.ptr, ex_d.ptr,
                                         ct.c_void_p(s.stream)))
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
X0 = np.nan_to_num(xyz_d.download())
M = _capi.MODE_POINTS_AND_EXTRINSICS
for rep in range(4):
    s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv)
    s.set_observations_dev(uv_d, mask_d)
    s.set_points(X0)
    st = s.solve(dense.default_opts(M, ftol=1e-10))
    print(json.dumps({k2: st[k2] for k2 in ('solve_ms', 'iterations', 'launches', 'init_ms', 'polish_ms', 'syncs', 'cost_final')}))
