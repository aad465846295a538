"""Sweep of the embedded point refinement (point_refine_iters x point_refine_skip, optionally polish_tol) on a C4-sized
scene: joint iterations, device time of the solve, final cost. Usage: python scripts/refine_sweep.py [frames] [outlier_frac]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as ct
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 119490
of = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
sc = synth.make_rig_scene(frames, 20, seed=4, outlier_frac=of)
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
x0 = np.nan_to_num(xyz_d.download())
s.set_observations_dev(uv_d, mask_d)
M = _capi.MODE_POINTS_AND_EXTRINSICS
combos = [(it, sk, pt) for it in (1, 2, 3) for sk in (1e-11, 1e-8, 1e-6, 1e-4) for pt in (3e-4,)]
combos += [(3, 1e-6, 0.0), (2, 1e-6, 0.0), (0, 0.0, 3e-4), (0, 0.0, 0.0), (2, 1e-6, 1e-3), (2, 1e-5, 3e-4), (2, 1e-3, 3e-4)]
for it, sk, pt in combos:
    best = None
    for rep in range(2):
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(x0)
        st = s.solve(dense.default_opts(M, ftol=1e-10, point_refine_iters=it, point_refine_skip=sk, polish_tol=pt, max_iter=400))
        if best is None or st['solve_ms'] < best['solve_ms']:
            best = st
    print(json.dumps(dict(refine=it, skip=sk, polish_tol=pt, iterations=best['iterations'], solve_ms=round(best['solve_ms'], 2),
                          polish_ms=round(best['polish_ms'], 2), cost_final=best['cost_final'], status=best['status'])), flush=True)
