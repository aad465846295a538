"""The two facade entry points of the reference's lib/app.py that sit on the hot path (src/lib/app.py:123-138):
they bind the fisheye callables and hand over to the managers in sba.py. core.sba calls sba_points_fisheye
(src/core/sba.py:43); the calibration notebook calls sba_board_points_fisheye (calib_with_gui.ipynb cell 30)."""
from .calib import triangulate_points_fisheye, project_points_fisheye
from .sba import _sba_board_points, _sba_points


def sba_board_points_fisheye(scene_fpath, points_fpaths, out_fpath, manual_points_fpath=None, manual_points_only=False,
                             camera_indices=None):
    """src/lib/app.py:123-126."""
    return _sba_board_points(scene_fpath, points_fpaths, manual_points_fpath, out_fpath, triangulate_points_fisheye,
                             project_points_fisheye, camera_indices, manual_points_only)


def sba_points_fisheye(scene_fpath, points_2d_df):
    """src/lib/app.py:135-138."""
    return _sba_points(scene_fpath, points_2d_df, triangulate_points_fisheye, project_points_fisheye)


def install(lib_calib=None, lib_utils=None, lib_sba=None):
    """Drop-in integration: rebinds the hot functions of already-imported reference modules
    (lib.calib, lib.utils, lib.sba) to this package's CUDA-backed ones; reference files are not edited."""
    from . import calib, utils, sba
    if lib_calib is not None:
        lib_calib.project_points_fisheye = calib.project_points_fisheye
        lib_calib.triangulate_points_fisheye = calib.triangulate_points_fisheye
    if lib_utils is not None:
        lib_utils.get_pairwise_3d_points_from_df = utils.get_pairwise_3d_points_from_df
    if lib_sba is not None:
        for name in ('cost_func_points_only', 'cost_func_points_extrinsics', 'params_to_points_extrinsics',
                     'bundle_adjust_points_only', 'bundle_adjust_points_and_extrinsics', 'bundle_adjust_board_points_only',
                     'prepare_calib_board_data_for_bundle_adjustment', 'prepare_manual_points_for_bundle_adjustment',
                     '_sba_board_points', '_sba_points'):
            setattr(lib_sba, name, getattr(sba, name))
        lib_sba.get_pairwise_3d_points_from_df = utils.get_pairwise_3d_points_from_df
