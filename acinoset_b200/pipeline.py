"""Array-level versions of the reference's TRI and SBA pipelines (src/core/tri.py:27-63, src/core/sba.py:27-69) for
callers that hold DeepLabCut output as arrays: threshold (`likelihood > dlc_thresh`, strict), frame window
(`between(start, end)`, inclusive), ring-pair triangulation, optional points-only SBA, and the scatter into the
`positions[(end-start+1), n_markers, 3]` array (NaN where a marker was not reconstructed) that the reference saves.
No pandas join is involved: a (frame, marker) pair IS a row of the dense slot layout, so the join / groupby / scatter of
the reference (where its TRI time goes) disappears; thresholding is fused into the K2 kernel.
"""
import numpy as np
from . import dense
from ._capi import MODE_POINTS_ONLY


def dlc_df_to_dense(points_2d_df, n_cams, markers, start_frame, end_frame):
    """DataFrame[frame, camera, marker, x, y, likelihood] (load_dlc_points_as_df layout, src/lib/utils.py:139-147) ->
    uv (F*K, C, 2), lik (F*K, C) with point index (frame - start_frame) * K + markers.index(marker). Rows outside the
    window or with an unknown marker are dropped; slots without a row get likelihood -1 (never above a threshold)."""
    F, K = int(end_frame) - int(start_frame) + 1, len(markers)
    uv = np.zeros((F * K, n_cams, 2))
    lik = np.full((F * K, n_cams), -1.0)
    if len(points_2d_df) == 0 or F <= 0:
        return uv, lik
    frame = points_2d_df['frame'].to_numpy()
    rank = {m: i for i, m in enumerate(markers)}
    mk = points_2d_df['marker'].map(rank).to_numpy(dtype=np.float64)          # NaN for unknown markers
    cam = points_2d_df['camera'].to_numpy().astype(np.int64)
    ok = (frame >= start_frame) & (frame <= end_frame) & ~np.isnan(mk) & (cam >= 0) & (cam < n_cams)
    p = (frame[ok].astype(np.int64) - int(start_frame)) * K + mk[ok].astype(np.int64)
    c = cam[ok]
    uv[p, c, 0] = points_2d_df['x'].to_numpy(dtype=np.float64)[ok]
    uv[p, c, 1] = points_2d_df['y'].to_numpy(dtype=np.float64)[ok]
    lik[p, c] = points_2d_df['likelihood'].to_numpy(dtype=np.float64)[ok]
    return uv, lik


def dlc_arrays_to_dense(per_camera):
    """per_camera: sequence (one entry per camera) of arrays (n_frames, n_markers, 3) holding x, y, likelihood (the
    DeepLabCut .h5 layout) -> uv (F*K, C, 2), lik (F*K, C)."""
    a = np.stack([np.asarray(x, dtype=np.float64) for x in per_camera], axis=2)        # (F, K, C, 3)
    F, K, C, _ = a.shape
    return np.ascontiguousarray(a[..., :2]).reshape(F * K, C, 2), np.ascontiguousarray(a[..., 2]).reshape(F * K, C)


def load_dlc_points_dense(dlc_df_fpaths, markers, start_frame, end_frame, frame_shifts=None):
    """DeepLabCut .h5 files -> dense slots for the window [start_frame, end_frame] without the long table of
    load_dlc_points_as_df (src/lib/utils.py:77-151) in between: uv (F*K, C, 2), lik (F*K, C) with point
    (frame - start_frame) * K + markers.index(marker); frames or markers a file does not hold get likelihood -1."""
    from . import utils
    tables = utils.load_dlc_tables(dlc_df_fpaths, frame_shifts)
    F, K, C = int(end_frame) - int(start_frame) + 1, len(markers), len(tables)
    uv = np.zeros((F * K, C, 2))
    lik = np.full((F * K, C), -1.0)
    for c, (frames, parts, vals) in enumerate(tables):
        keep = (frames >= start_frame) & (frames <= end_frame)
        fi = (frames[keep] - int(start_frame)).astype(np.int64)
        col = {bp: j for j, bp in enumerate(parts)}
        for k, m in enumerate(markers):
            if m in col:
                v = vals[keep, col[m], :]
                uv[fi * K + k, c, 0] = v[:, 0]
                uv[fi * K + k, c, 1] = v[:, 1]
                lik[fi * K + k, c] = v[:, 2]
    return uv, lik


def tri(uv, lik, n_markers, dlc_thresh, camera_params):
    """core.tri on dense arrays: positions (F, K, 3), NaN where no ring pair saw the marker."""
    k_arr, d_arr, r_arr, t_arr = camera_params[:4]
    xyz, exists, _mask = dense.triangulate_ring(uv, lik, dlc_thresh, k_arr, d_arr, r_arr, t_arr)
    return xyz.reshape(-1, n_markers, 3)


def sba(uv, lik, n_markers, dlc_thresh, camera_params, f_scale=50.0, **solver_opts):
    """core.sba on dense arrays: TRI initialisation + points-only SBA (f_scale = 50, src/lib/sba.py:305).
    Returns (positions (F, K, 3), solver statistics)."""
    k_arr, d_arr, r_arr, t_arr = camera_params[:4]
    st = dense.solve_scene_host(uv, lik, dlc_thresh, k_arr, d_arr, r_arr, t_arr, mode=MODE_POINTS_ONLY, f_scale=float(f_scale),
                                **solver_opts)
    pos = st.pop('points').reshape(-1, n_markers, 3)
    return pos, st
