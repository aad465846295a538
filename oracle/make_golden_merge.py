"""TEST INFRASTRUCTURE ONLY — golden fixture for the board + manual-point merge of _sba_board_points
(src/lib/sba.py:246-258): runs the REAL reference manager on a seeded synthetic scene with its loaders, its solver and
save_scene replaced IN THE HARNESS by recorders, and stores what the reference's own two prepare_* functions returned and
what it handed to bundle_adjust_points_and_extrinsics after merging them. The merged ids are offset by
point_3d_indices.max() (not max() + 1), i.e. the first manual point shares an id with the last board corner: recorded as
is. Reference files are never edited. Run in the build container:  python -m oracle.make_golden_merge
"""
import os
import numpy as np
from . import ref_shim

N_BOARDS, SEED, N_MANUAL = 5, 17, 9


def gen_merge():
    from acinoset_b200 import synth
    from .make_golden import GOLD, _versions
    calib, utils, sba = ref_shim.load()
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    k, d, r, t, res = rig
    sc = synth.make_board_scene(N_BOARDS, seed=SEED, rig=rig)
    mp = synth.make_manual_points_case(N_MANUAL, seed=SEED + 1, rig=rig)
    manual = mp['img_pts'].copy()
    rng = np.random.default_rng(SEED + 2)
    manual[rng.uniform(size=manual.shape[:2]) < 0.45] = np.nan          # unlabelled views
    manual[0, 1:] = np.nan                                               # a point seen once: dropped by the reference's prep
    rec = {}
    saved = {n: getattr(sba, n) for n in ('load_points', 'load_scene', 'load_manual_points', 'save_scene',
                                           'prepare_calib_board_data_for_bundle_adjustment',
                                           'prepare_manual_points_for_bundle_adjustment', 'bundle_adjust_points_and_extrinsics')}

    def rec_board(*a, **kw):
        out = saved['prepare_calib_board_data_for_bundle_adjustment'](*a, **kw)
        rec['board'] = out
        return out

    def rec_manual(*a, **kw):
        out = saved['prepare_manual_points_for_bundle_adjustment'](*a, **kw)
        rec['manual'] = out
        return out

    def rec_solve(p2, p3, pidx, cidx, k_, d_, r_, t_, project_func):
        rec['merged'] = (p2, p3, pidx, cidx)
        return p3, r_, t_, dict(before=np.zeros(2 * len(p2)), after=np.zeros(2 * len(p2)))

    try:
        sba.load_points = lambda fp: (sc['img_pts_arr'][fp], sc['fnames_arr'][fp], sc['board_shape'], sc['square'], res)
        sba.load_scene = lambda fp: (k, d, sc['r0'], sc['t0'], res)
        sba.load_manual_points = lambda fp: (manual, ['manual'], res)
        sba.save_scene = lambda *a, **kw: None
        sba.prepare_calib_board_data_for_bundle_adjustment = rec_board
        sba.prepare_manual_points_for_bundle_adjustment = rec_manual
        sba.bundle_adjust_points_and_extrinsics = rec_solve
        with ref_shim.quiet():
            sba._sba_board_points('scene', list(range(len(k))), 'manual', 'out', calib.triangulate_points_fisheye,
                                  calib.project_points_fisheye)
    finally:
        for n, f in saved.items():
            setattr(sba, n, f)
    out = dict(versions=_versions(), n_boards=N_BOARDS, seed=SEED, manual_points=manual, r0=sc['r0'], t0=sc['t0'])
    for name in ('board', 'manual', 'merged'):
        p2, p3, pidx, cidx = rec[name]
        out.update({f'{name}_points_2d': p2, f'{name}_points_3d': p3, f'{name}_point_indices': np.asarray(pidx).astype(np.int64),
                    f'{name}_camera_indices': np.asarray(cidx).astype(np.int64)})
    np.savez_compressed(os.path.join(GOLD, 'golden_board_merge.npz'), **out)
    print('board', rec['board'][1].shape, 'manual', rec['manual'][1].shape, 'merged', rec['merged'][1].shape,
          'ids', rec['merged'][2].min(), rec['merged'][2].max())


if __name__ == '__main__':
    gen_merge()
