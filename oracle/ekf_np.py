"""TEST INFRASTRUCTURE ONLY — numpy restatement of the EKF measurement; never imported by the product.

marker_coords       : get_3d_marker_coords({'x': x}, mode) at tau = 0 (/root/reference/src/lib/misc.py:144-327)
measurement         : h and H as the loop of /root/reference/src/core/ekf.py:245-251 builds them from h_function (:66-71)
                      and numerical_jacobian (:81-96)
Pinned by tests/golden/golden_ekf.npz (oracle/make_golden_n3.py, from the reference's own functions).
"""
import numpy as np
from . import fisheye_np as fe


def _rx(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, s], [0, -s, c]])


def _ry(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, -s], [0, 1, 0], [s, 0, c]])


def _rz(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, s, 0], [-s, c, 0], [0, 0, 1]])


def marker_coords(x, mode='default'):
    x = np.asarray(x, dtype=np.float64)
    v = lambda *a: np.array(a, dtype=np.float64)
    head = x[0:3]
    RI0 = _rz(x[5]) @ _rx(x[3]) @ _ry(x[4])
    e, n = (0.03, 0.055) if mode == 'default' else (0.038852231676497324, 0.0571868749393016)
    nose, r_eye, l_eye = head + RI0.T @ v(n, 0, -n), head + RI0.T @ v(0, -e, 0), head + RI0.T @ v(0, e, 0)
    if mode == 'head':
        return np.array([nose, r_eye, l_eye])
    RI1 = _rz(x[9]) @ _rx(x[7]) @ _ry(x[8]) @ RI0
    RI2 = _ry(x[10]) @ RI1
    neck = head + RI1.T @ v(x[6], 0, 0)
    spine = neck + RI2.T @ v(-0.37, 0, 0)
    if mode == 'head_stabilize':
        return np.array([nose, r_eye, l_eye, neck, spine])
    r_sh, l_sh = neck + RI2.T @ v(-0.04, -0.08, -0.10), neck + RI2.T @ v(-0.04, 0.08, -0.10)
    if mode == 'upper_body':
        return np.array([nose, r_eye, l_eye, neck, spine, r_sh, l_sh])
    RI3 = _rz(x[13]) @ _rx(x[11]) @ _ry(x[12]) @ RI2
    RI4 = _rz(x[15]) @ _ry(x[14]) @ RI3
    RI5 = _rz(x[17]) @ _ry(x[16]) @ RI4
    RI6 = _ry(x[18]) @ RI2; RI7 = _ry(x[19]) @ RI6
    RI8 = _ry(x[20]) @ RI2; RI9 = _ry(x[21]) @ RI8
    RI10 = _ry(x[22]) @ RI3; RI11 = _ry(x[23]) @ RI10
    RI12 = _ry(x[24]) @ RI3; RI13 = _ry(x[25]) @ RI12
    tail0 = spine + RI3.T @ v(-0.37, 0, 0)
    tail1 = tail0 + RI4.T @ v(-0.28, 0, 0)
    tail2 = tail1 + RI5.T @ v(-0.36, 0, 0)
    l_fk = l_sh + RI6.T @ v(0, 0, -0.24); l_fa = l_fk + RI7.T @ v(0, 0, -0.28)
    r_fk = r_sh + RI8.T @ v(0, 0, -0.24); r_fa = r_fk + RI9.T @ v(0, 0, -0.28)
    l_hip = tail0 + RI3.T @ v(0.12, 0.08, -0.06); l_bk = l_hip + RI10.T @ v(0, 0, -0.32); l_ba = l_bk + RI11.T @ v(0, 0, -0.25)
    r_hip = tail0 + RI3.T @ v(0.12, -0.08, -0.06); r_bk = r_hip + RI12.T @ v(0, 0, -0.32); r_ba = r_bk + RI13.T @ v(0, 0, -0.25)
    return np.array([nose, r_eye, l_eye, neck, spine, tail0, tail1, tail2, r_sh, r_fk, r_fa, l_sh, l_fk, l_fa,
                     r_hip, r_bk, r_ba, l_hip, l_bk, l_ba, x[26:29]])


def measurement(x, k_arr, d_arr, r_arr, t_arr, mode='default', n_states=None, eps=1e-3):
    x = np.asarray(x, dtype=np.float64)
    n_cams, n_pose = len(k_arr), len(x)
    n_states = 3 * n_pose if n_states is None else n_states

    def hf(xx, j):
        return fe.project_points_fisheye(marker_coords(xx, mode), k_arr[j], d_arr[j], r_arr[j], t_arr[j]).flatten()

    R = len(hf(x, 0))
    h = np.zeros(n_cams * R)
    H = np.zeros((n_cams * R, n_states))
    for j in range(n_cams):
        fx = hf(x, j)
        h[j * R:(j + 1) * R] = fx
        xp = x.copy()
        for i in range(n_pose):
            xp[i] = xp[i] + eps
            H[j * R:(j + 1) * R, i] = (hf(xp, j) - fx) / eps
            xp[i] = x[i]
    return h, H
