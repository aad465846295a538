"""TEST INFRASTRUCTURE ONLY — golden fixture for two-view triangulation of badly conditioned pairs: the REAL reference
(src/lib/calib.py:120-129, cv2.fisheye.undistortPoints + cv2.triangulatePoints) on points 30-100 m from the shortest and
the longest ring baseline of the rig and on points within 5 degrees of a baseline (where the two rays are nearly
parallel and the 4x4 DLT system is closest to rank 2). Run in the build container:  python -m oracle.make_golden_far
"""
import os
import numpy as np
from . import ref_shim


def gen_far():
    from acinoset_b200 import synth
    from .make_golden import GOLD, _versions
    calib, utils, sba = ref_shim.load()
    k, d, r, t, _ = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    C = len(k)
    cen = np.stack([-(r[c].T @ np.asarray(t[c]).reshape(3)) for c in range(C)])
    rng = np.random.default_rng(23)
    base = [np.linalg.norm(cen[c] - cen[(c + 1) % C]) for c in range(C)]
    pairs = [int(np.argmin(base)), int(np.argmax(base))]
    out = dict(versions=_versions(), pairs=np.array([[a, (a + 1) % C] for a in pairs]))
    for i, a in enumerate(pairs):
        b = (a + 1) % C
        mid = 0.5 * (cen[a] + cen[b])
        ax = (cen[b] - cen[a]) / np.linalg.norm(cen[b] - cen[a])
        fwd = 0.5 * (r[a][2] + r[b][2])                                  # mean optical axis (world)
        fwd = fwd - (fwd @ ax) * ax
        fwd /= np.linalg.norm(fwd)
        n = 400
        X = np.empty((2 * n, 3))
        dist = rng.uniform(30.0, 100.0, n)
        X[:n] = mid + fwd * dist[:, None] + rng.normal(0, 0.1, (n, 3)) * dist[:, None]          # far away, in front
        ang = np.deg2rad(rng.uniform(1.0, 5.0, n))                                               # near the baseline, beyond b
        X[n:] = cen[b] + (np.cos(ang)[:, None] * ax + np.sin(ang)[:, None] * fwd) * rng.uniform(2.0, 10.0, n)[:, None]
        pa = calib.project_points_fisheye(X, k[a], d[a], r[a], t[a]) + rng.normal(0, 1.0, (2 * n, 2))
        pb = calib.project_points_fisheye(X, k[b], d[b], r[b], t[b]) + rng.normal(0, 1.0, (2 * n, 2))
        tri = calib.triangulate_points_fisheye(pa, pb, k[a], d[a], r[a], t[a], k[b], d[b], r[b], t[b])
        ok = np.isfinite(tri).all(1) & (np.abs(tri).max(1) < 1e5)
        out.update({f'X{i}': X[ok], f'pa{i}': pa[ok], f'pb{i}': pb[ok], f'tri{i}': tri[ok]})
        print('pair', (a, b), 'baseline %.2f m' % base[a], 'kept', int(ok.sum()), 'max |tri - X| %.2f m' % np.abs(tri[ok] - X[ok]).max())
    np.savez_compressed(os.path.join(GOLD, 'golden_tri_far.npz'), **out)


if __name__ == '__main__':
    gen_far()
