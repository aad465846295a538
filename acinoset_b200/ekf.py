"""The measurement step of the reference's EKF update (src/core/ekf.py:66-96, 245-251) on the device: skeleton forward
kinematics (get_3d_marker_coords, src/lib/misc.py:144-327), fisheye projection into every camera and the forward-
difference Jacobian, for all cameras and all perturbed poses in two kernel launches (acino_ekf_measurement).
The Kalman recursion itself (prediction, gain, smoothing) stays with the caller: it is sequential in time and out of
scope (BASELINE: EKF untouched)."""
import numpy as np
from . import _capi

MODES = {'default': 0, 'head': 1, 'upper_body': 2, 'head_stabilize': 3}
N_POSE = {'default': 29, 'head': 6, 'upper_body': 11, 'head_stabilize': 11}
N_MARKERS = {'default': 21, 'head': 3, 'upper_body': 7, 'head_stabilize': 5}


def measurement(x_pose, k_arr, d_arr, r_arr, t_arr, marker_mode='default', n_states=None, eps=1e-3, with_coords=False):
    """h (n_cams*n_markers*2,) and H (n_cams*n_markers*2, n_states) exactly as the loop at src/core/ekf.py:245-251 fills
    them: h[j*n_markers*2:(j+1)*n_markers*2] = h_function(x, camera j).flatten(), H[..., :n_pose] = numerical_jacobian
    (forward differences, eps = 1e-3), other columns zero. x_pose = states[:vel_idx]."""
    _capi.require_gpu()
    if marker_mode not in MODES:
        raise ValueError(f'unknown marker mode {marker_mode!r}')
    x = _capi.f64(x_pose).reshape(-1)
    n_pose, n_mk = N_POSE[marker_mode], N_MARKERS[marker_mode]
    if x.size != n_pose:
        raise ValueError(f'marker mode {marker_mode!r} has {n_pose} pose parameters, got {x.size}')
    n_states = 3 * n_pose if n_states is None else int(n_states)
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    n_cams = cams.shape[0]
    h = np.empty(n_cams * n_mk * 2)
    H = np.empty((n_cams * n_mk * 2, n_states))
    coords = np.empty((n_mk, 3)) if with_coords else None
    _capi.check(_capi.lib().acino_ekf_measurement(_capi.ptr(x), n_pose, MODES[marker_mode], _capi.ptr(cams), n_cams, float(eps),
                                                  n_states, _capi.ptr(h), _capi.ptr(H), _capi.ptr(coords) if with_coords else None))
    return (h, H, coords) if with_coords else (h, H)


def get_3d_marker_coords(states, mode='default'):
    """src/lib/misc.py:144-327 for numeric states at tau = 0: {'x': pose} -> (n_markers, 3)."""
    rig_k = np.eye(3)[None]
    _, _, coords = measurement(states['x'], rig_k, np.zeros((1, 4)), np.eye(3)[None], np.zeros((1, 3, 1)), mode, with_coords=True)
    return coords


def h_function(x, k, d, r, t, marker_mode='default'):
    """src/core/ekf.py:66-71: pixels (n_markers, 2) of the skeleton in one camera."""
    h, _ = measurement(x, np.asarray(k)[None], np.asarray(d).reshape(1, -1), np.asarray(r)[None], np.asarray(t).reshape(1, 3, 1), marker_mode)
    return h.reshape(-1, 2)
