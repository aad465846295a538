"""Per-kernel device times of the points+extrinsics LM iteration (K4a, K4b, fold+K5, K6) at a C4-sized scene, and a full
solve from the TRI start for the cost trajectory. Usage: python scripts/k4_probe.py [frames] [steps] [masked_frac]"""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as ct
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
mf = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
sc = synth.make_rig_scene(frames, 20, seed=4, masked_frac=mf)
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
s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv)
s.set_observations_dev(uv_d, mask_d)
s.set_points(np.nan_to_num(xyz_d.download()))
M = _capi.MODE_POINTS_AND_EXTRINSICS
t0 = time.perf_counter()
st = s.solve(dense.default_opts(M, ftol=1e-10))
wall = time.perf_counter() - t0
print(json.dumps(dict(solve=st, wall_ms=wall * 1e3)))
s.reset_state()
o = dense.default_opts(M, inner_iters=1, force_iters=1, point_refine_iters=0)
for _ in range(5):
    s.iterate(o)
s.profile(True)
_capi.check(L.acino_sync())
t0 = time.perf_counter()
for _ in range(steps):
    s.iterate(o)
_capi.check(L.acino_sync())
wall = time.perf_counter() - t0
pr = s.profile_read()
s.profile(False)
print(json.dumps(dict(frames=frames, steps=steps, ms_per_step_wall=wall / steps * 1e3, kernels=pr, lm=s.lm_state())))
# the same iterations replayed as a CUDA graph (no profiling events between the kernels)
for _ in range(5):
    s.iterate(o)
_capi.check(L.acino_sync())
t0 = time.perf_counter()
for _ in range(steps):
    s.iterate(o)
_capi.check(L.acino_sync())
print(json.dumps(dict(ms_per_step_graph=(time.perf_counter() - t0) / steps * 1e3)))
