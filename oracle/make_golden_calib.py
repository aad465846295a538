"""TEST INFRASTRUCTURE ONLY — golden fixture for the extrinsic-calibration SBA (config C2, small): runs the REAL
reference (src/lib/sba.py: prepare_calib_board_data_for_bundle_adjustment + bundle_adjust_points_and_extrinsics)
on a seeded synthetic board scene, twice:
  * 'fixed'   — create_bundle_adjustment_jacobian_sparsity_matrix monkey-patched IN THE HARNESS to the pattern that
                matches the [rvecs | tvecs | points] parameter layout. This run reaches a
                stationary point of the reference's own objective and is the parity target.
  * 'shipped' — the reference exactly as written (stalls before convergence; recorded for the record).
Reference files are never edited. Run in the build container:  python -m oracle.make_golden calib
"""
import os
import time
import numpy as np
from . import ref_shim

N_BOARDS = 8
SEED = 3


def _consistent_pattern(n_cams, n_params_per_camera, camera_indices, n_points, point_indices):
    from scipy.sparse import lil_matrix
    m = camera_indices.size * 2
    n = n_cams * n_params_per_camera + n_points * 3
    A = lil_matrix((m, n), dtype=int)
    i = np.arange(camera_indices.size)
    for s in range(3):
        for row in (2 * i, 2 * i + 1):
            A[row, camera_indices * 3 + s] = 1                       # rvec_j
            A[row, n_cams * 3 + camera_indices * 3 + s] = 1          # t_j
            A[row, n_cams * 6 + point_indices * 3 + s] = 1           # X_p
    return A


def gen_calib(run_shipped=True):
    from acinoset_b200 import synth
    from .make_golden import GOLD, _versions
    calib, utils, sba = ref_shim.load()
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    sc = synth.make_board_scene(N_BOARDS, seed=SEED, rig=rig)
    k, d = sc['k'], sc['d']
    r0, t0 = sc['r0'], sc['t0']
    with ref_shim.quiet():
        p2, p3, pidx, cidx = sba.prepare_calib_board_data_for_bundle_adjustment(
            sc['img_pts_arr'], sc['fnames_arr'], sc['board_shape'], k, d, r0, t0, calib.triangulate_points_fisheye)
    out = dict(points_2d=p2, points_3d=p3, point_indices=pidx.astype(np.int64), camera_indices=cidx.astype(np.int64),
               r0=r0, t0=t0, n_boards=N_BOARDS, seed=SEED, versions=_versions())
    orig = sba.create_bundle_adjustment_jacobian_sparsity_matrix
    sba.create_bundle_adjustment_jacobian_sparsity_matrix = _consistent_pattern
    try:
        t_0 = time.time()
        with ref_shim.quiet():
            obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(p2, p3, pidx, cidx, k, d, r0, t0, calib.project_points_fisheye)
        out.update(fixed_obj=obj, fixed_r=r1, fixed_t=t1, fixed_before=res['before'], fixed_after=res['after'],
                   fixed_seconds=time.time() - t_0)
        print('fixed-pattern run:', time.time() - t_0, 's; rmse', np.sqrt(np.mean(res['before'] ** 2)), '->', np.sqrt(np.mean(res['after'] ** 2)))
    finally:
        sba.create_bundle_adjustment_jacobian_sparsity_matrix = orig
    np.savez_compressed(os.path.join(GOLD, 'golden_calib.npz'), **out)
    if run_shipped:
        t_0 = time.time()
        with ref_shim.quiet():
            obj, r1, t1, res = sba.bundle_adjust_points_and_extrinsics(p2, p3, pidx, cidx, k, d, r0, t0, calib.project_points_fisheye)
        out.update(shipped_obj=obj, shipped_r=r1, shipped_t=t1, shipped_after=res['after'], shipped_seconds=time.time() - t_0)
        print('as-shipped run:', time.time() - t_0, 's; rmse ->', np.sqrt(np.mean(res['after'] ** 2)))
        np.savez_compressed(os.path.join(GOLD, 'golden_calib.npz'), **out)


if __name__ == '__main__':
    gen_calib()
