"""Drop-in for the fisheye camera functions of the reference's lib/calib.py (src/lib/calib.py:120-136).

Same names, argument order, accepted shapes and return shapes; the arithmetic runs in the CUDA
library (K1 / K2). fp32 inputs are up-cast to fp64 (the reference's board path passes fp32 pixels,
src/lib/sba.py:88-89).
"""
from typing import Dict
import numpy as np
from . import _capi


def project_points_fisheye(obj_pts, k, d, r, t, shutter_delay_param: Dict = None):
    """src/lib/calib.py:132-136. obj_pts (N,3) or (3,) -> (N,2) float64 pixels.
    ``shutter_delay_param`` is accepted and ignored, as in the reference."""
    _capi.require_gpu()
    X = _capi.f64(obj_pts).reshape(-1, 3)
    uv = np.empty((X.shape[0], 2))
    cam = _capi.pack_cam(k, d, r, t)
    _capi.check(_capi.lib().acino_project(_capi.ptr(X), X.shape[0], _capi.ptr(cam), _capi.ptr(uv)))
    return uv


def project_points_fisheye_all(obj_pts, k_arr, d_arr, r_arr, t_arr, cam_res=None):
    """project_points_fisheye for every camera of the rig in ONE launch (the per-camera loop of save_3d_cheetah_as_2d,
    src/lib/utils.py:255-263). obj_pts (..., 3) -> (n_cams, N, 2). With cam_res = (width, height) the reference's range rule is
    applied on the device: a pixel outside [0, width] x [0, height] becomes (NaN, NaN) (utils.py:262-263)."""
    _capi.require_gpu()
    X = _capi.f64(obj_pts).reshape(-1, 3)
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    n_cams = cams.shape[0]
    uv = np.empty((n_cams, X.shape[0], 2))
    w, h = (float(cam_res[0]), float(cam_res[1])) if cam_res is not None else (0.0, 0.0)
    _capi.check(_capi.lib().acino_project_all(_capi.ptr(X), X.shape[0], _capi.ptr(cams), n_cams, w, h, _capi.ptr(uv)))
    return uv


def triangulate_points_fisheye(img_pts_1, img_pts_2, k1, d1, r1, t1, k2, d2, r2, t2):
    """src/lib/calib.py:120-129. Pixels (M,1,2) / (M,2) / (2,) per view -> (M,3) float64."""
    _capi.require_gpu()
    p1 = _capi.f64(img_pts_1).reshape(-1, 2)
    p2 = _capi.f64(img_pts_2).reshape(-1, 2)
    if p1.shape[0] != p2.shape[0]:
        raise ValueError('img_pts_1 and img_pts_2 must hold the same number of points')
    xyz = np.empty((p1.shape[0], 3))
    ca = _capi.pack_cam(k1, d1, r1, t1)
    cb = _capi.pack_cam(k2, d2, r2, t2)
    _capi.check(_capi.lib().acino_triangulate_pair(_capi.ptr(p1), _capi.ptr(p2), p1.shape[0], _capi.ptr(ca), _capi.ptr(cb),
                                                   _capi.ptr(xyz)))
    return xyz
