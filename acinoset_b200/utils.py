"""Drop-in for get_pairwise_3d_points_from_df of the reference's lib/utils.py (src/lib/utils.py:319-349),
plus the scene IO helpers the hot path's callers use (same JSON format, src/lib/utils.py:55-74,186-203).
"""
import json
from datetime import datetime
import numpy as np
import pandas as pd
from . import dense


def load_scene(fpath, verbose=True):
    """src/lib/utils.py:55-74."""
    with open(fpath, 'r') as f:
        data = json.load(f)
    cam_res = tuple(data['camera_resolution'])
    k_arr = np.array([c['k'] for c in data['cameras']], dtype=np.float64)
    d_arr = np.array([c['d'] for c in data['cameras']], dtype=np.float64)
    r_arr = np.array([c['r'] for c in data['cameras']], dtype=np.float64)
    t_arr = np.array([c['t'] for c in data['cameras']], dtype=np.float64)
    if verbose:
        print(f'Loaded extrinsics from {fpath}\n')
    return k_arr, d_arr, r_arr, t_arr, cam_res


def save_scene(out_fpath, k_arr, d_arr, r_arr, t_arr, cam_res):
    """src/lib/utils.py:186-203."""
    cameras = [{'k': np.asarray(k).tolist(), 'd': np.asarray(d).tolist(), 'r': np.asarray(r).tolist(), 't': np.asarray(t).tolist()}
               for k, d, r, t in zip(k_arr, d_arr, r_arr, t_arr)]
    data = {'timestamp': str(datetime.now()), 'camera_resolution': cam_res, 'cameras': cameras}
    with open(out_fpath, 'w') as f:
        json.dump(data, f)
    print(f'Saved extrinsics to {out_fpath}\n')


def _is_fisheye_triangulate(func):
    return callable(func) and getattr(func, '__name__', '') == 'triangulate_points_fisheye'


def df_to_dense(points_2d_df, n_cams):
    """Lowers the (already thresholded) DataFrame[frame, camera, marker, x, y, ...] to dense slots.
    Point order = sorted (frame, marker), i.e. the row order of groupby(['frame','marker']) at
    src/lib/utils.py:348. Returns uv (P,C,2), present (P,C) bool, frames (P,), markers (P,), slot (n_rows,)."""
    df = points_2d_df
    if len(df) == 0:
        return (np.zeros((0, n_cams, 2)), np.zeros((0, n_cams), bool), np.zeros(0), np.zeros(0, dtype=object),
                np.zeros(0, dtype=np.int64))
    keys = pd.MultiIndex.from_arrays([df['frame'].to_numpy(), df['marker'].to_numpy()])
    codes, uniq = keys.factorize(sort=True)
    P = len(uniq)
    cam = df['camera'].to_numpy().astype(np.int64)
    if cam.min() < 0 or cam.max() >= n_cams:
        raise ValueError('camera index outside the scene')
    slot = codes.astype(np.int64) * n_cams + cam
    if np.unique(slot).size != slot.size:
        raise ValueError('duplicate (frame, marker, camera) rows: each camera may observe a point once')
    uv = np.zeros((P * n_cams, 2))
    uv[slot, 0] = df['x'].to_numpy(dtype=np.float64)
    uv[slot, 1] = df['y'].to_numpy(dtype=np.float64)
    present = np.zeros(P * n_cams, dtype=bool)
    present[slot] = True
    frames = np.asarray(uniq.get_level_values(0))
    markers = np.asarray(uniq.get_level_values(1))
    return uv.reshape(P, n_cams, 2), present.reshape(P, n_cams), frames, markers, slot


def get_pairwise_3d_points_from_df(points_2d_df, k_arr, d_arr, r_arr, t_arr, triangulate_func, verbose=True):
    """src/lib/utils.py:319-349 — ring-pair two-view DLT + unweighted mean, one CUDA launch (K2).

    ``points_2d_df`` is the caller-thresholded frame (src/core/tri.py:41-42). ``triangulate_func`` must be
    a fisheye triangulation (the reference's or this package's); other camera models are not built."""
    if not _is_fisheye_triangulate(triangulate_func):
        raise NotImplementedError('only triangulate_points_fisheye is supported (CUDA path, no fallback)')
    n_cams = len(k_arr)
    uv, present, frames, markers, _slot = df_to_dense(points_2d_df, n_cams)
    if verbose:
        for a in range(n_cams):
            b = (a + 1) % n_cams
            n = int((present[:, a] & present[:, b]).sum()) if len(present) else 0
            if n > 0:
                print(f'Found {n} pairwise points between camera {a} and {b}')
            else:
                print(f'No pairwise points between camera {a} and {b}')
        print()
    if uv.shape[0] == 0:
        return pd.DataFrame({'frame': [], 'marker': [], 'x': [], 'y': [], 'z': []})
    xyz, exists, _mask = dense.triangulate_ring(uv, present.astype(np.float64), 0.5, k_arr, d_arr, r_arr, t_arr)
    return pd.DataFrame({'frame': frames[exists], 'marker': markers[exists],
                         'x': xyz[exists, 0], 'y': xyz[exists, 1], 'z': xyz[exists, 2]})


def load_points(fpath, verbose=False):
    """Checkerboard corner file of one camera (src/lib/utils.py:18-28): float32 points, file names, board geometry."""
    with open(fpath, 'r') as f:
        data = json.load(f)
    fnames = list(data['points'].keys())
    points = np.array(list(data['points'].values()), dtype=np.float32)
    if verbose:
        print(f'Loaded checkerboard points from {fpath}\n')
    return points, fnames, tuple(data['board_shape']), data['board_square_len'], tuple(data['camera_resolution'])


def load_manual_points(fpath, verbose=True):
    """Manually labelled points (src/lib/utils.py:31-41): (n_points, n_cams, 2) with NaN where unseen."""
    with open(fpath, 'r') as f:
        data = json.load(f)
    points = np.array(data['points'])
    fnames = [f'img{str(i).zfill(5)}.jpg' for i in data['frame_idx']]
    if verbose:
        print(f'Loaded manual points from {fpath}\n')
    return points, fnames, tuple(data['camera_resolution'])


def create_board_object_pts(board_shape, square_edge_length):
    """src/lib/utils.py:313-316."""
    object_pts = np.zeros((board_shape[0] * board_shape[1], 3), np.float32)
    object_pts[:, :2] = np.mgrid[0:board_shape[0], 0:board_shape[1]].T.reshape(-1, 2) * square_edge_length
    return object_pts
