"""Grid over lambda0 / point_refine_iters / polish_tol for the points+extrinsics solve on a few synthetic scenes. Needs a GPU."""
import sys, os, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi
E = _capi.MODE_POINTS_AND_EXTRINSICS
grid = list(itertools.product((1.0, 10.0, 100.0), (3, 4, 6), (3e-5, 3e-4, 3e-3), (0, 1)))
tot = {g: 0.0 for g in grid}
for frames, seed, kw in ((119490, 4, {}), (119490, 6, dict(outlier_frac=0.05)), (119490, 10, dict(masked_frac=0.2)), (20000, 7, {}), (2000, 8, {})):
    sc = synth.make_rig_scene(frames, 20, seed=seed, **kw)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(100 + seed)
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
    tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    s = dense.SbaSolver(C, P)
    s.set_observations(sc['uv'], mask)
    best = None
    rows = []
    for g in grid:
        lam0, refine, tol, pre = g
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
        pre_ms = 0.0
        if pre:     # the points solved with the cameras fixed before the first joint step
            sp = s.solve(dense.default_opts(_capi.MODE_POINTS_ONLY, f_scale=1.0, max_iter=100))
            pre_ms = sp['solve_ms']
        st = s.solve(dense.default_opts(E, ftol=1e-10, max_iter=400, lambda0=lam0, point_refine_iters=refine, polish_tol=tol))
        rows.append((g, st['solve_ms'] + pre_ms, st['iterations'], st['cost_final']))
        tot[g] += st['solve_ms'] + pre_ms
    cmin = min(r[3] for r in rows)
    print(f'--- {frames} frames seed {seed} {kw}: best cost {cmin:.6f}')
    for g, ms, it, cost in rows:
        print(f'  lambda0 {g[0]:g} refine {g[1]} polish_tol {g[2]:g} pre {g[3]}: {ms:7.2f} ms {it:3d} it  cost-best {cost - cmin:.3e}', flush=True)
    s.close()
print('--- total ms per setting')
for g in sorted(tot, key=tot.get):
    print(f'  lambda0 {g[0]:g} refine {g[1]} polish_tol {g[2]:g} pre {g[3]}: {tot[g]:.1f}')
