"""Drop-in for get_pairwise_3d_points_from_df of the reference's lib/utils.py (src/lib/utils.py:319-349),
plus the scene IO helpers the hot path's callers use (same JSON format, src/lib/utils.py:55-74,186-203).
"""
import json
import os
from glob import glob
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


def _point_keys(frame, marker):
    """(frame, marker) columns -> one int64 key per row whose order is the order of sorted (frame, marker-name), i.e. of
    groupby(['frame', 'marker']) at src/lib/utils.py:348, plus what is needed to map keys back. Integer arithmetic on
    factorised columns instead of a MultiIndex of (int, str) tuples: 0.5 s instead of 19 s for the 7 M rows of config C4.
    The marker column is factorised as the pandas column it is (an Arrow-backed string column is dictionary-encoded in
    50 ms; as 7 M Python objects it took 0.5 s), and integer frames within a range comparable to the number of rows
    are their own codes."""
    m_codes, m_uniq = pd.factorize(marker, sort=True)
    m_uniq = np.asarray(m_uniq, dtype=object)
    f_arr = np.asarray(frame)
    if f_arr.dtype.kind in 'iu' and f_arr.size:
        f_min, f_max = int(f_arr.min()), int(f_arr.max())
        if f_max - f_min + 1 <= 8 * f_arr.size + 1024:
            f_codes = f_arr.astype(np.int64) - f_min
            f_uniq = np.arange(f_min, f_max + 1, dtype=f_arr.dtype)
            n_m = max(len(m_uniq), 1)
            return f_codes * n_m + m_codes.astype(np.int64), f_uniq, m_uniq, n_m
    f_codes, f_uniq = pd.factorize(f_arr, sort=True)
    n_m = max(len(m_uniq), 1)
    return f_codes.astype(np.int64) * n_m + m_codes.astype(np.int64), f_uniq, m_uniq, n_m


def df_to_dense(points_2d_df, n_cams, return_keys=False):
    """Lowers the (already thresholded) DataFrame[frame, camera, marker, x, y, ...] to dense slots.
    Point order = sorted (frame, marker), i.e. the row order of groupby(['frame','marker']) at
    src/lib/utils.py:348. Returns uv (P,C,2), present (P,C) bool, frames (P,), markers (P,), slot (n_rows,)
    (+ the rows' point ids when return_keys)."""
    df = points_2d_df
    if len(df) == 0:
        out = (np.zeros((0, n_cams, 2)), np.zeros((0, n_cams), bool), np.zeros(0), np.zeros(0, dtype=object),
               np.zeros(0, dtype=np.int64))
        return out + (np.zeros(0, dtype=np.int64),) if return_keys else out
    key, f_uniq, m_uniq, n_m = _point_keys(df['frame'].to_numpy(), df['marker'])
    # the keys are bounded by (frames x markers): a presence bitmap and its running count give the sorted unique keys
    # and every row's point id without a sort
    seen = np.zeros(len(f_uniq) * n_m, dtype=bool)
    seen[key] = True
    uniq_keys = np.flatnonzero(seen)
    codes = (np.cumsum(seen, dtype=np.int64) - 1)[key]
    P = len(uniq_keys)
    cam = df['camera'].to_numpy().astype(np.int64, copy=False)
    if cam.min() < 0 or cam.max() >= n_cams:
        raise ValueError('camera index outside the scene')
    slot = codes * n_cams + cam
    present = np.zeros(P * n_cams, dtype=bool)
    present[slot] = True
    if int(present.sum()) != slot.size:
        raise ValueError('duplicate (frame, marker, camera) rows: each camera may observe a point once')
    uv = np.zeros((P * n_cams, 2))
    uv[slot] = np.column_stack((df['x'].to_numpy(dtype=np.float64), df['y'].to_numpy(dtype=np.float64)))      # one scatter of rows
    frames = np.asarray(f_uniq)[uniq_keys // n_m]
    markers = m_uniq[uniq_keys % n_m]
    out = (uv.reshape(P, n_cams, 2), present.reshape(P, n_cams), frames, markers, slot)
    return out + (codes,) if return_keys else out


def get_pairwise_3d_points_from_df(points_2d_df, k_arr, d_arr, r_arr, t_arr, triangulate_func, verbose=True):
    """src/lib/utils.py:319-349 — ring-pair two-view DLT + unweighted mean, one CUDA launch (K2).

    ``points_2d_df`` is the caller-thresholded frame (src/core/tri.py:41-42). ``triangulate_func`` must be
    a fisheye triangulation (the reference's or this package's); other camera models are not built."""
    if not _is_fisheye_triangulate(triangulate_func):
        raise NotImplementedError('only triangulate_points_fisheye is supported (CUDA path, no fallback)')
    out_df, _extra = _pairwise_3d_with_index(points_2d_df, k_arr, d_arr, r_arr, t_arr, verbose)
    return out_df


def _pairwise_3d_with_index(points_2d_df, k_arr, d_arr, r_arr, t_arr, verbose=True):
    """get_pairwise_3d_points_from_df + what _sba_points needs to build its index arrays without a second relational
    pass: per row of points_2d_df the row of the returned frame its (frame, marker) went to (-1: not triangulated)."""
    n_cams = len(k_arr)
    uv, present, frames, markers, _slot, codes = df_to_dense(points_2d_df, n_cams, return_keys=True)
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
        return pd.DataFrame({'frame': [], 'marker': [], 'x': [], 'y': [], 'z': []}), np.zeros(0, dtype=np.int64)
    xyz, exists, _mask = dense.triangulate_ring(uv, present.astype(np.float64), 0.5, k_arr, d_arr, r_arr, t_arr)
    row_of_point = np.full(len(exists), -1, dtype=np.int64)
    row_of_point[exists] = np.arange(int(exists.sum()))
    out_df = pd.DataFrame({'frame': frames[exists], 'marker': markers[exists],
                           'x': xyz[exists, 0], 'y': xyz[exists, 1], 'z': xyz[exists, 2]})
    return out_df, row_of_point[codes]


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


def project_3d_cheetah_to_2d(position3d_arr, k_arr, d_arr, r_arr, t_arr, cam_res, bodyparts, start_frame):
    """The numeric core of save_3d_cheetah_as_2d (src/lib/utils.py:248-268): per camera a DLC-shaped DataFrame (columns
    MultiIndex bodyparts x [x, y, likelihood], index start_frame..) of the reprojected positions, pixels outside the image
    set to NaN, likelihood NaN. A single positions array is projected into all cameras by ONE kernel launch; a per-camera
    list (the reference accepts one, utils.py:249-250) takes one launch per entry."""
    from . import calib
    n_cams = len(k_arr)
    pdindex = pd.MultiIndex.from_product([bodyparts, ['x', 'y', 'likelihood']], names=['bodyparts', 'coords'])
    if isinstance(position3d_arr, list):
        pos = [np.asarray(p, dtype=np.float64) for p in position3d_arr]
        prj = [calib.project_points_fisheye_all(pos[i], k_arr[i:i + 1], np.asarray(d_arr)[i:i + 1], r_arr[i:i + 1], t_arr[i:i + 1],
                                                cam_res)[0] for i in range(n_cams)]
    else:
        pos = [np.asarray(position3d_arr, dtype=np.float64)] * n_cams
        prj = calib.project_points_fisheye_all(pos[0], k_arr, d_arr, r_arr, t_arr, cam_res)
    out = []
    for i in range(n_cams):
        n_frames = len(pos[i])
        data = np.full(pos[i].shape, np.nan)
        data[:, :, 0:2] = prj[i].reshape((n_frames, -1, 2))
        out.append(pd.DataFrame(data.reshape((n_frames, -1)), columns=pdindex, index=range(start_frame, start_frame + n_frames)))
    return out


def save_3d_cheetah_as_2d(position3d_arr, out_dir, scene_fpath, bodyparts, project_func, start_frame, save_as_csv=True,
                          out_fname=None):
    """src/lib/utils.py:237-283, same arguments, files and return value. `project_func` selects the camera model as in the
    reference; only the fisheye model is implemented here. The .h5 copy needs PyTables, like the reference; when it is not
    importable the .csv is still written and the .h5 skipped with a message."""
    assert os.path.dirname(os.path.dirname(scene_fpath)) in out_dir, 'scene_fpath does not belong to the same parent folder as out_dir'
    if getattr(project_func, '__name__', '') != 'project_points_fisheye':
        raise NotImplementedError('only the fisheye camera model is implemented')
    video_fpaths = sorted(glob(os.path.join(out_dir, 'cam[1-9].mp4')))
    if not video_fpaths:
        video_fpaths = sorted(glob(os.path.join(os.path.dirname(out_dir), 'cam[1-9].mp4')))
    if not video_fpaths:
        print('Could not save 3D cheetah to 2D - No videos were found in', out_dir, 'or', os.path.dirname(out_dir))
        return []
    k_arr, d_arr, r_arr, t_arr, cam_res = load_scene(scene_fpath, verbose=False)
    assert len(k_arr) == len(video_fpaths)
    out_fname = os.path.basename(out_dir) if out_fname is None else out_fname
    result_dfs = project_3d_cheetah_to_2d(position3d_arr, k_arr, d_arr, r_arr, t_arr, cam_res, bodyparts, start_frame)
    try:
        import tables  # noqa: F401
        have_h5 = True
    except ImportError:
        have_h5 = False
        print('PyTables is not installed: .h5 output skipped')
    for i, df in enumerate(result_dfs):
        cam_name = os.path.splitext(os.path.basename(video_fpaths[i]))[0]
        fpath = os.path.join(out_dir, cam_name + '_' + out_fname + '.h5')
        if have_h5:
            df.to_hdf(fpath, key=f'{out_fname}_df', format='table', mode='w')
        if save_as_csv:
            df.to_csv(os.path.splitext(fpath)[0] + '.csv')
    fpath = fpath.replace(cam_name, 'cam*')
    if have_h5:
        print('Saved', fpath)
    if save_as_csv:
        print('Saved', os.path.splitext(fpath)[0] + '.csv')
    print()
    return result_dfs


# ========== DeepLabCut tables (src/lib/utils.py:77-151) ==========

def _read_dlc_table(path, dlc_head=False):
    """One DeepLabCut .h5 -> (frames (F,), bodyparts in alphabetical order, values (F, K, 3) = x, y, likelihood).
    Mirrors the per-file part of load_dlc_points_as_df: tables without a likelihood column get 1 where y is present and 0
    elsewhere (utils.py:104-118) and their index (image file names) becomes int(name[-7:-4]); the 'dlc_head' tables of the
    reference's head dataset are renamed, reduced to nose / r_eye / l_eye and get their likelihood from x
    (utils.py:84-103)."""
    df = pd.read_hdf(path)                    # needs PyTables, like the reference
    if dlc_head:
        df = df.rename(columns={'bodypart1': 'r_eye', 'bodypart2': 'l_eye', 'bodypart3': 'nose'}, level=1)
        df = df.drop(columns=['objectA'], level=1)
    parts = list(dict.fromkeys(df.columns.get_level_values(1)))
    if dlc_head:
        parts = [p_ for p_ in parts if p_ in ('nose', 'r_eye', 'l_eye')]       # the reference's reindex keeps these three
    coords = set(df.columns.get_level_values(2))
    scorer = df.columns.get_level_values(0)[0]
    frames = np.asarray(df.index)
    has_lik = 'likelihood' in coords and not dlc_head
    if not has_lik:
        frames = np.array([int(str(s_)[-7:-4]) for s_ in frames])
    # bodyparts come out alphabetical from the reference's unstack; the head tables keep the order of its reindex
    parts_sorted = [p_ for p_ in ('nose', 'r_eye', 'l_eye') if p_ in parts] if dlc_head else sorted(parts)
    vals = np.empty((len(df), len(parts_sorted), 3))
    for j, bp in enumerate(parts_sorted):
        vals[:, j, 0] = df[(scorer, bp, 'x')].to_numpy(dtype=np.float64)
        vals[:, j, 1] = df[(scorer, bp, 'y')].to_numpy(dtype=np.float64)
        if has_lik:
            vals[:, j, 2] = df[(scorer, bp, 'likelihood')].to_numpy(dtype=np.float64)
        else:
            vals[:, j, 2] = (~np.isnan(vals[:, j, 0 if dlc_head else 1])).astype(np.float64)
    return frames, parts_sorted, vals


def _shift_frames(vals, shift):
    """The reference's frame-drift correction (utils.py:125-137): the rows of a camera move down by `shift` frames; vacated
    frames keep their labels with x, y = NaN; the caller's fillna(0) makes their likelihood 0."""
    if shift == 0:
        return vals
    out = np.full_like(vals, np.nan)
    if shift > 0:
        out[shift:] = vals[:len(vals) - shift]
    else:
        out[:shift] = vals[-shift:]
    return out


def load_dlc_tables(dlc_df_fpaths, frame_shifts=None):
    """Per camera (frames, bodyparts, values (F, K, 3)) with the frame shifts applied and NaN likelihoods set to 0."""
    assert frame_shifts is None or len(dlc_df_fpaths) == len(frame_shifts), '`frame_shifts` should be the same size with `dlc_df_fpaths`'
    out = []
    for i, path in enumerate(dlc_df_fpaths):
        frames, parts, vals = _read_dlc_table(path, dlc_head='dlc_head' in dlc_df_fpaths[0])
        if frame_shifts is not None:
            vals = _shift_frames(vals, int(frame_shifts[i]))
        vals[:, :, 2] = np.nan_to_num(vals[:, :, 2], nan=0.0)
        out.append((frames, parts, vals))
    return out


def load_dlc_points_as_df(dlc_df_fpaths, frame_shifts=None, verbose=False):
    """src/lib/utils.py:77-151: the DeepLabCut .h5 files of all cameras as one long table
    [frame, camera, marker, x, y, likelihood], rows camera-major, then frame, then marker (alphabetical, as the
    reference's unstack leaves them). Built from arrays instead of five pandas reshapes per file. Columns are numeric
    (the reference's come out as `object` under pandas >= 2 because it concatenates onto an empty frame)."""
    dfs = []
    for i, (frames, parts, vals) in enumerate(load_dlc_tables(dlc_df_fpaths, frame_shifts)):
        F, K = len(frames), len(parts)
        dfs.append(pd.DataFrame({'frame': np.repeat(frames, K), 'camera': i, 'marker': np.tile(np.array(parts, dtype=object), F),
                                 'x': vals[:, :, 0].ravel(), 'y': vals[:, :, 1].ravel(), 'likelihood': vals[:, :, 2].ravel()}))
    dlc_df = pd.concat(dfs, ignore_index=True) if dfs else pd.DataFrame(columns=['frame', 'camera', 'marker', 'x', 'y', 'likelihood'])
    dlc_df = dlc_df[['frame', 'camera', 'marker', 'x', 'y', 'likelihood']]
    if verbose:
        print(f'DLC points dataframe:\n{dlc_df}\n')
    return dlc_df
