"""Drop-in for residual_error of the reference's lib/metric.py (src/lib/metric.py:36-101): per camera, the reprojection
error table of every marker. The reference projects marker by marker (one OpenCV call each, inside pandas filtering);
here all markers of a camera go through ONE K1 launch and the bookkeeping is vectorised. Same signature, same dict of
DataFrames, same columns and row order (markers in the order given, frames ascending), and — like the reference, whose
np.vstack with the marker names turns every column into strings — string-typed cells.
"""
import numpy as np
import pandas as pd
from . import calib

COLUMNS = ['frame', 'marker', 'camera_distance', 'pixel_residual', 'pck_threshold', 'error_u', 'error_v']


def _nose_to_eye(points_2d_df, cam, frames_3d):
    """Per-frame pixel distance nose <-> eye (left eye, else right eye) of one camera: the PCK threshold
    (src/lib/metric.py:43-61)."""
    cam_df = points_2d_df[points_2d_df['camera'] == cam]
    nose = cam_df[cam_df['marker'] == 'nose'].sort_values(by=['frame'])
    l_eye = cam_df[cam_df['marker'] == 'l_eye'].sort_values(by=['frame'])
    r_eye = cam_df[cam_df['marker'] == 'r_eye'].sort_values(by=['frame'])
    eye = l_eye.combine_first(r_eye).drop_duplicates(subset=['frame'], keep='first')
    valid = np.intersect1d(eye['frame'].to_numpy(), nose['frame'].to_numpy())
    valid = np.intersect1d(valid, np.intersect1d(cam_df['frame'].to_numpy(), frames_3d))
    nose = nose[nose['frame'].isin(valid)]
    eye = eye[eye['frame'].isin(valid)]
    dist = np.linalg.norm(nose[['x', 'y']].to_numpy() - eye[['x', 'y']].to_numpy(), axis=1)
    return pd.DataFrame({'distance': dist}, index=pd.Index(nose['frame'].to_numpy(), name='frame'))


def residual_error(points_2d_df, points_3d_dfs, markers, camera_params):
    """src/lib/metric.py:36-101."""
    k_arr, d_arr, r_arr, t_arr, _, _ = camera_params
    n_cam = len(k_arr)
    if not isinstance(points_3d_dfs, list):
        points_3d_dfs = [points_3d_dfs] * n_cam
    rank = {m: i for i, m in enumerate(markers)}
    error = {}
    for i in range(n_cam):
        p3 = points_3d_dfs[i]
        pck = _nose_to_eye(points_2d_df, i, p3['frame'].to_numpy())
        d2 = points_2d_df[(points_2d_df['camera'] == i) & points_2d_df[['x', 'y']].notnull().all(axis=1)]
        d3 = p3[p3[['x', 'y', 'z']].notnull().all(axis=1)]
        d2 = d2[d2['marker'].isin(rank)][['frame', 'marker', 'x', 'y']]
        d3 = d3[d3['marker'].isin(rank)][['frame', 'marker', 'x', 'y', 'z']]
        m = d2.merge(d3, how='inner', on=['frame', 'marker'], suffixes=('_2d', ''))
        if len(m) == 0:
            error[str(i)] = pd.DataFrame(columns=COLUMNS)
            continue
        order = np.lexsort((m['frame'].to_numpy(), m['marker'].map(rank).to_numpy()))
        m = m.iloc[order]
        pts_2d = m[['x_2d', 'y_2d']].to_numpy(dtype=np.float64)
        pts_3d = m[['x', 'y', 'z']].to_numpy(dtype=np.float64)
        prj = calib.project_points_fisheye(pts_3d, k_arr[i], d_arr[i], r_arr[i], t_arr[i])     # one launch per camera
        cam_pos = np.squeeze(np.asarray(t_arr)[i, :, :])       # the reference uses t, not the camera centre (metric.py:83)
        cam_dist = np.sqrt(np.sum((pts_3d - cam_pos) ** 2, axis=1))
        err = pts_2d - prj
        frames = m['frame'].to_numpy()
        thr = pck.reindex(frames)['distance'].to_numpy()
        table = np.vstack((frames, m['marker'].to_numpy().astype(str), cam_dist, np.sqrt(np.sum(err ** 2, axis=1)), thr, err.T)).T
        error[str(i)] = pd.DataFrame(table, columns=COLUMNS)
    return error


def pck(pix_errors, ratio=0.5):
    """Percentage of correct keypoints over all cameras, as save_error_dists computes it (src/core/metrics.py:36-45): share of
    rows whose pixel residual is <= ratio * nose-to-eye distance; rows without a threshold (NaN) count as incorrect."""
    dist = np.concatenate([np.asarray(df['pixel_residual'], dtype=np.float64) for df in pix_errors.values()] or [np.zeros(0)])
    thr = np.concatenate([np.asarray(df['pck_threshold'], dtype=np.float64) for df in pix_errors.values()] or [np.zeros(0)])
    if len(dist) == 0:
        return float('nan')
    return float(100.0 * (np.sum(dist <= (ratio * thr)) / len(dist)))
