"""Debug aid: where does the slow tail of the points+extrinsics LM come from? After N joint iterations, solve the points
alone to convergence (cameras fixed) and report how much of the remaining cost decrease that recovers. Needs a GPU."""
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
E = _capi.MODE_POINTS_AND_EXTRINSICS
for n in (15, 25, 40):
    s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
    a = s.solve(dense.default_opts(E, max_iter=n, ftol=0.0, gtol=0.0, xtol=0.0))
    cams = s.get_cameras(); X = s.get_points()
    b = s.solve(dense.default_opts(_capi.MODE_POINTS_ONLY, f_scale=1.0, max_iter=400))
    print(f"after {n} joint iterations: cost {a['cost_final']:.6f}; points-only to convergence ({b['iterations']} it, {b['solve_ms']:.1f} ms): "
          f"{b['cost_initial']:.6f} -> {b['cost_final']:.6f} (decrease {b['cost_initial'] - b['cost_final']:.4f}, unconverged {b['n_unconverged']})", flush=True)
    # continue the joint solve from there
    c = s.solve(dense.default_opts(E, ftol=1e-10, max_iter=300))
    print(f"   joint solve continued from the refined points: {c['iterations']} it, {c['solve_ms']:.1f} ms, cost {c['cost_final']:.6f}", flush=True)
