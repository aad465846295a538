"""TEST INFRASTRUCTURE ONLY — numpy restatement of the SURVEY.md §8(f) N3 callers; never imported by the product.

manual_points_residuals : residual_arr of adjust_extrinsics_manual_points (/root/reference/src/lib/calib.py:224-256)
redescending_loss       : /root/reference/src/lib/misc.py:329-343, elementwise
ring_critical           : closed form of the threshold loop of initialize_marker_3d (/root/reference/src/lib/app.py:44-57)
Pinned by tests/golden/golden_manual_adjust.npz and golden_init_marker.npz (made by oracle/make_golden_n3.py from the
real reference functions).
"""
import numpy as np
from . import fisheye_np as fe


def redescending_loss(err, a, b, c):
    e = np.abs(err)
    step = lambda s, x: 1 / (1 + np.e ** (-1 * (x - s)))
    piece = lambda s, t, x: step(s, x) - step(t, x)
    cost = (1 - step(a, e)) / 2 * e ** 2
    cost = cost + piece(a, b, e) * (a * e - (a ** 2) / 2)
    cost = cost + piece(b, c, e) * (a * b - (a ** 2) / 2 + (a * (c - b) / 2) * (1 - ((c - e) / (c - b)) ** 2))
    cost = cost + step(c, e) * (a * b - (a ** 2) / 2 + (a * (c - b) / 2))
    return cost


def manual_points_residuals(R_t, img_pts_arr, cam_idxs_to_correct, cam_pairs, k_arr, d_arr, r_arr, t_arr, loss=(3, 10, 20)):
    R = fe.rodrigues_vec_to_mat(np.asarray(R_t[0:3], dtype=np.float64))
    t = np.asarray(R_t[3:6], dtype=np.float64).reshape(3, 1)
    n_cams = len(k_arr)
    t_arr = np.asarray(t_arr, dtype=np.float64).reshape(n_cams, 3, 1)

    def cam(i):
        if i in cam_idxs_to_correct:
            return k_arr[i], d_arr[i], r_arr[i] @ R.T, t_arr[i] - r_arr[i] @ t
        return k_arr[i], d_arr[i], r_arr[i], t_arr[i]

    skew = [fe.triangulate_points_fisheye(img_pts_arr[:, a], img_pts_arr[:, b], *cam(a), *cam(b)) for a, b in cam_pairs]
    out = []
    for i in range(n_cams):
        for pts in skew:
            err = img_pts_arr[:, i] - fe.project_points_fisheye(pts, *cam(i))
            out.append(redescending_loss(err, *loss))
    return np.array(out).flatten()


def ring_critical(lik):
    lik = np.where(np.isnan(lik), -np.inf, np.asarray(lik, dtype=np.float64))
    return np.max(np.minimum(lik, np.roll(lik, -1, axis=1)), axis=1)
