"""The factor by which an accepted step may shrink the damping (acino_sba_opts.lambda_dec; Nielsen's rule uses 1/3)
against solve time, iterations and final cost on five synthetic scenes (profiles/r02_lambda_dec_sweep.txt). Needs a GPU."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi
E = _capi.MODE_POINTS_AND_EXTRINSICS
decs = [float(a) for a in sys.argv[1:]] or [1 / 3, 0.2, 0.1, 0.05]
tot = {d: 0.0 for d in decs}
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
    rows = []
    for dec in decs:
        best = None
        for rep in range(2):
            s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
            st = s.solve(dense.default_opts(E, ftol=1e-10, max_iter=400, lambda_dec=dec))
            best = st if best is None or st['solve_ms'] < best['solve_ms'] else best
        rows.append((dec, best['solve_ms'], best['iterations'], best['cost_final'], best['n_accept']))
        tot[dec] += best['solve_ms']
    cmin = min(r[3] for r in rows)
    print(f'--- {frames} frames seed {seed} {kw}: best cost {cmin:.6f}')
    for dec, ms, it, cost, acc in rows:
        print(f'  dec {dec:.3f}: {ms:7.2f} ms {it:3d} it ({acc} accepted)  cost-best {cost - cmin:.3e}', flush=True)
    s.close()
print('--- total ms per setting')
for dec in sorted(tot, key=tot.get):
    print(f'  dec {dec:.3f}: {tot[dec]:.1f}')
