"""Points+extrinsics solve with and without polishing (acino_sba_opts.polish_tol) over sizes and seeds. Needs a GPU."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from acinoset_b200 import synth, dense, sba, _capi
E = _capi.MODE_POINTS_AND_EXTRINSICS
for frames, seed, kw in ((119490, 4, {}), (119490, 5, {}), (119490, 6, dict(outlier_frac=0.05)), (20000, 7, {}), (2000, 8, {}), (200, 9, {}),
                         (119490, 10, dict(masked_frac=0.2)), (119490, 11, dict(noise_px=0.3))):
    sc = synth.make_rig_scene(frames, 20, seed=seed, **kw)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    xyz, exists, mask = dense.triangulate_ring(sc['uv'], sc['lik'], 0.5, k, d, r, t)
    P, C = sc['lik'].shape
    rng = np.random.default_rng(100 + seed)
    rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
    tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
    s = dense.SbaSolver(C, P)
    s.set_observations(sc['uv'], mask)
    out = []
    for tol in (0.0, 3e-5):
        s.set_cameras(k, d, r, t); s.set_rvec_t(rv, tv); s.set_points(np.nan_to_num(xyz))
        st = s.solve(dense.default_opts(E, ftol=1e-10, max_iter=400, polish_tol=tol))
        out.append(f"polish {tol:g}: {st['solve_ms']:.1f} ms {st['iterations']} it status {int(st['status'])} cost {st['cost_final']:.6f}")
    print(f'{frames} frames seed {seed} {kw}: ' + ' | '.join(out), flush=True)
    s.close()
