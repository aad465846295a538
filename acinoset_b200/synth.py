"""Synthetic 6-camera fisheye scenes (SURVEY.md §8d generators).

Data generation only: builds the inputs the tests and bench.py feed to the hot path
(dense ``[frame*keypoint][camera]`` observation slots and the equivalent DeepLabCut-style
DataFrame of src/lib/utils.py:77-151). Nothing here is on the product's compute path.
"""
import json
import os
import numpy as np

# src/lib/misc.py:10-17 — the 20 body keypoints of get_markers('default') (lure excluded)
MARKERS = [
    'nose', 'r_eye', 'l_eye', 'neck_base', 'spine', 'tail_base', 'tail1', 'tail2',
    'r_shoulder', 'r_front_knee', 'r_front_ankle', 'l_shoulder', 'l_front_knee', 'l_front_ankle',
    'r_hip', 'r_back_knee', 'r_back_ankle', 'l_hip', 'l_back_knee', 'l_back_ankle',
]

_SCENE = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden', 'scene_6cam.json')


def load_rig(path=None):
    """Scene JSON in the reference's format (src/lib/utils.py:55-74) -> k(n,3,3) d(n,4,1) r(n,3,3) t(n,3,1)."""
    with open(path or _SCENE) as f:
        data = json.load(f)
    cams = data['cameras']
    k = np.array([c['k'] for c in cams], dtype=np.float64)
    d = np.array([c['d'] for c in cams], dtype=np.float64)
    r = np.array([c['r'] for c in cams], dtype=np.float64)
    t = np.array([c['t'] for c in cams], dtype=np.float64)
    return k, d, r, t, tuple(data['camera_resolution'])


def _project_np(X, k, d, r, t):
    # Kannala-Brandt projection, generator-side only (ground truth -> pixels)
    Xc = X @ r.T + t.reshape(3)
    a = Xc[:, 0] / Xc[:, 2]
    b = Xc[:, 1] / Xc[:, 2]
    rr = np.sqrt(a * a + b * b)
    th = np.arctan(rr)
    t2 = th * th
    dd = d.reshape(-1)
    thd = th * (1 + t2 * (dd[0] + t2 * (dd[1] + t2 * (dd[2] + t2 * dd[3]))))
    s = np.where(rr > 1e-8, thd / np.maximum(rr, 1e-300), 1.0)
    return np.stack([k[0, 0] * (a * s + k[0, 1] / k[0, 0] * b * s) + k[0, 2], k[1, 1] * b * s + k[1, 2]], axis=1)


def make_rig_scene(n_frames, n_kpts=20, seed=0, noise_px=1.0, masked_frac=0.5, outlier_frac=0.0,
                   outlier_px=50.0, rig=None, project=None):
    """Ground-truth points uniform in the rig's common view volume x[-1,5] y[5,8] z[0,1] m,
    projected into every camera, N(0, noise_px) pixel noise, likelihood ~ U(0,1) shifted so that
    ``masked_frac`` of the slots fall at or below 0.5 (the bench's dlc_thresh).

    Returns dict(uv (P,C,2) f64, lik (P,C) f64, gt (P,3), k,d,r,t) with P = n_frames*n_kpts and
    point p = frame*n_kpts + keypoint."""
    k, d, r, t, _ = rig if rig is not None else load_rig()
    C = len(k)
    rng = np.random.default_rng(seed)
    P = n_frames * n_kpts
    gt = np.empty((P, 3))
    gt[:, 0] = rng.uniform(-1.0, 5.0, P)
    gt[:, 1] = rng.uniform(5.0, 8.0, P)
    gt[:, 2] = rng.uniform(0.0, 1.0, P)
    uv = np.empty((P, C, 2))
    proj = project or _project_np
    for c in range(C):
        uv[:, c, :] = proj(gt, k[c], d[c], r[c], t[c])
    uv += rng.normal(0.0, noise_px, uv.shape)
    if outlier_frac > 0:
        out = rng.uniform(size=(P, C)) < outlier_frac
        uv[out] += rng.normal(0.0, outlier_px, (int(out.sum()), 2))
    u01 = rng.uniform(size=(P, C))
    # likelihood > 0.5 with probability 1 - masked_frac
    lik = np.where(u01 < masked_frac, 0.5 * u01 / max(masked_frac, 1e-12),
                   0.5 + 0.5 * (u01 - masked_frac) / max(1 - masked_frac, 1e-12) + 1e-9)
    lik = np.minimum(lik, 1.0)
    return dict(uv=uv, lik=lik, gt=gt, k=k, d=d, r=r, t=t, n_frames=n_frames, n_kpts=n_kpts)


def make_track_scene(n_frames, n_kpts=20, seed=0, noise_px=1.0, rig=None):
    """Like make_rig_scene but the keypoints ride a smooth cubic trajectory through the view volume (a running animal)
    instead of being redrawn every frame; likelihood ~ U(0,1). For the trajectory-fit callers (initialize_marker_3d)."""
    k, d, r, t, _ = rig if rig is not None else load_rig()
    C = len(k)
    rng = np.random.default_rng(seed)
    tt = np.linspace(0.0, 1.0, n_frames)
    base = np.stack([-0.5 + 4.5 * tt + 0.6 * tt ** 2 - 0.8 * tt ** 3, 6.0 + 1.2 * tt - 0.9 * tt ** 2, 0.4 + 0.3 * tt ** 2], axis=1)
    off = rng.uniform(-0.25, 0.25, (n_kpts, 3))
    gt = (base[:, None, :] + off[None, :, :]).reshape(-1, 3)
    uv = np.empty((len(gt), C, 2))
    for c in range(C):
        uv[:, c, :] = _project_np(gt, k[c], d[c], r[c], t[c])
    uv += rng.normal(0.0, noise_px, uv.shape)
    lik = rng.uniform(size=(len(gt), C))
    return dict(uv=uv, lik=lik, gt=gt, k=k, d=d, r=r, t=t, n_frames=n_frames, n_kpts=n_kpts)


def make_manual_points_case(n_points=14, seed=0, noise_px=0.3, bad_cams=(3, 4, 5), rig=None):
    """Manually-labelled points seen by every camera, and a rig whose cameras `bad_cams` are off by one common rigid
    transform (what adjust_extrinsics_manual_points corrects, src/lib/calib.py:216-297).
    Returns dict(img_pts (N, C, 2), k, d, r_true, t_true, r_bad, t_bad (C,3,1), bad_cams)."""
    k, d, r, t, _ = rig if rig is not None else load_rig()
    C = len(k)
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-1, 5, n_points), rng.uniform(5, 8, n_points), rng.uniform(0, 1, n_points)], axis=1)
    img = np.empty((n_points, C, 2))
    for c in range(C):
        img[:, c, :] = _project_np(X, k[c], d[c], r[c], t[c])
    img += rng.normal(0.0, noise_px, img.shape)
    Rg = _rot_from_axis_angle(rng.normal(0, 0.01, 3))
    tg = rng.normal(0, 0.02, (3, 1))
    r_bad = np.array(r, dtype=np.float64).copy()
    t_bad = np.array(t, dtype=np.float64).reshape(C, 3, 1).copy()
    for c in bad_cams:                          # inverse of the reference's update r_f = r_i R^T, t_f = t_i - r_i t
        r_bad[c] = r[c] @ Rg
        t_bad[c] = t_bad[c] + r_bad[c] @ tg
    return dict(img_pts=img, k=k, d=d, r_true=np.array(r), t_true=np.array(t).reshape(C, 3, 1), r_bad=r_bad, t_bad=t_bad,
                bad_cams=list(bad_cams), gt=X)


def make_dlc_tables(n_frames, markers, n_cams, seed=0, first_frame=0, nan_frac=0.05):
    """DeepLabCut-style wide tables (what pd.read_hdf returns for a DLC .h5): index = frame number, columns
    (scorer, bodyparts, coords); a few likelihood / coordinate cells are NaN."""
    import pandas as pd
    out = []
    for c in range(n_cams):
        rng = np.random.default_rng(seed * 100 + c)
        cols = pd.MultiIndex.from_product([['DLC_resnet50'], markers, ['x', 'y', 'likelihood']], names=['scorer', 'bodyparts', 'coords'])
        a = rng.uniform(0, 1, (n_frames, len(markers) * 3))
        a[:, 0::3] *= 2704; a[:, 1::3] *= 1520
        a[rng.uniform(size=a.shape) < nan_frac] = np.nan
        out.append(pd.DataFrame(a, columns=cols, index=range(first_frame, first_frame + n_frames)))
    return out


def make_dlc_xy_tables(n_frames, markers, n_cams, seed=0, nan_frac=0.15):
    """Hand-labelled style tables without a likelihood column (index = image file names ending in NNN.png); cells are NaN
    in x only, in y only, or in both. Returns (tables, paths)."""
    import pandas as pd
    out = []
    for c in range(n_cams):
        rng = np.random.default_rng(seed * 100 + c)
        cols = pd.MultiIndex.from_product([['labeller'], markers, ['x', 'y']], names=['scorer', 'bodyparts', 'coords'])
        a = rng.uniform(0, 1, (n_frames, len(markers) * 2))
        a[:, 0::2] *= 2704; a[:, 1::2] *= 1520
        a[rng.uniform(size=a.shape) < nan_frac] = np.nan
        index = [f'labeled-data/cam{c + 1}/img{f:03d}.png' for f in range(4, 4 + n_frames)]
        out.append(pd.DataFrame(a, columns=cols, index=index))
    return out, [f'/data/labels/cam{i + 1}.h5' for i in range(n_cams)]


def make_dlc_head_tables(n_frames, n_cams, seed=0, nan_frac=0.15):
    """The reference's head dataset layout (src/lib/utils.py:84-103): scorer '2019-03-09_lily_run', bodyparts
    bodypart1..3 + objectA, x / y only, paths containing 'dlc_head'. Returns (tables, paths)."""
    import pandas as pd
    out = []
    for c in range(n_cams):
        rng = np.random.default_rng(seed * 100 + c)
        cols = pd.MultiIndex.from_product([['2019-03-09_lily_run'], ['bodypart1', 'bodypart2', 'bodypart3', 'objectA'], ['x', 'y']],
                                          names=['scorer', 'bodyparts', 'coords'])
        a = rng.uniform(0, 1, (n_frames, 8))
        a[:, 0::2] *= 2704; a[:, 1::2] *= 1520
        a[rng.uniform(size=a.shape) < nan_frac] = np.nan
        index = [f'labeled-data/cam{c + 1}/img{f:03d}.png' for f in range(10, 10 + n_frames)]
        out.append(pd.DataFrame(a, columns=cols, index=index))
    return out, [f'/data/dlc_head/cam{i + 1}.h5' for i in range(n_cams)]


def dense_to_df(uv, lik, n_kpts, first_frame=0, markers=None):
    """Dense slots -> DataFrame[frame, camera, marker, x, y, likelihood], rows camera-major then
    frame then marker, like load_dlc_points_as_df (src/lib/utils.py:139-147)."""
    import pandas as pd
    markers = markers or MARKERS[:n_kpts]
    P, C = lik.shape
    F = P // n_kpts
    frame = np.repeat(np.arange(first_frame, first_frame + F), n_kpts)
    marker = np.tile(np.array(markers, dtype=object), F)
    dfs = []
    for c in range(C):
        dfs.append(pd.DataFrame({'frame': frame, 'camera': c, 'marker': marker,
                                 'x': uv[:, c, 0], 'y': uv[:, c, 1], 'likelihood': lik[:, c]}))
    return pd.concat(dfs, ignore_index=True)


def board_object_points(board_shape=(9, 6), square=0.088):
    """Planar checkerboard corner grid, same ordering as create_board_object_pts (src/lib/utils.py:313-316)."""
    g = np.zeros((board_shape[0] * board_shape[1], 3))
    g[:, :2] = np.mgrid[0:board_shape[0], 0:board_shape[1]].T.reshape(-1, 2) * square
    return g


def _rot_from_axis_angle(v):
    th = np.linalg.norm(v)
    if th < 1e-15:
        return np.eye(3)
    kx, ky, kz = v / th
    K = np.array([[0, -kz, ky], [kz, 0, -kx], [-ky, kx, 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def make_board_scene(n_boards, seed=3, noise_px=0.3, board_shape=(9, 6), square=0.088, rig=None,
                     rvec_sigma=8e-4, t_sigma=3e-3, project=None):
    """Synthetic extrinsic-calibration input: ``n_boards`` checkerboard poses, each seen by two
    ring-adjacent cameras (+ a third with probability 0.5), N(0, noise_px) corner noise, stored float32 per camera in
    the reference's load_points layout (img_pts_arr[cam] (n_img, 54, 1, 2) f32, fnames_arr[cam] list of names).
    Also returns perturbed initial extrinsics (truth + N(0, rvec_sigma) rad, N(0, t_sigma) m)."""
    k, d, r, t, res = rig if rig is not None else load_rig()
    C = len(k)
    rng = np.random.default_rng(seed)
    grid = board_object_points(board_shape, square)
    grid = grid - grid.mean(axis=0)
    proj = project or _project_np
    img_pts = [[] for _ in range(C)]
    fnames = [[] for _ in range(C)]
    gt_pts = []
    b = 0
    tries = 0
    while b < n_boards and tries < 200 * n_boards:
        tries += 1
        yaw = rng.uniform(0, 2 * np.pi)
        tilt = rng.uniform(-0.3, 0.3, 2)
        Rb = _rot_from_axis_angle(np.array([0, 0, yaw])) @ _rot_from_axis_angle(np.array([tilt[0], tilt[1], 0.0]))
        centre = np.array([rng.uniform(0, 4), rng.uniform(5.5, 7.5), rng.uniform(0.3, 1.0)])
        Xw = grid @ Rb.T + centre
        a = int(rng.integers(0, C))
        cams = [a, (a + 1) % C]
        if rng.uniform() < 0.5:
            cams.append((a + 2) % C)
        ok = True
        pix = {}
        for c in cams:
            Xc = Xw @ r[c].T + t[c].reshape(3)
            uv = proj(Xw, k[c], d[c], r[c], t[c])
            if (Xc[:, 2] < 0.5).any() or (uv[:, 0] < 0).any() or (uv[:, 0] > res[0]).any() or (uv[:, 1] < 0).any() or (uv[:, 1] > res[1]).any():
                ok = False
                break
            pix[c] = uv + rng.normal(0, noise_px, uv.shape)
        if not ok:
            continue
        name = f'img{b:05d}.jpg'
        for c in sorted(cams):
            img_pts[c].append(pix[c].astype(np.float32).reshape(-1, 1, 2))
            fnames[c].append(name)
        gt_pts.append(Xw)
        b += 1
    if b < n_boards:
        raise RuntimeError('could not place the requested number of boards in view')
    img_pts = [np.array(p, dtype=np.float32).reshape(-1, grid.shape[0], 1, 2) for p in img_pts]
    r0 = np.array([r[c] @ _rot_from_axis_angle(rng.normal(0, rvec_sigma, 3)) for c in range(C)])
    t0 = t + rng.normal(0, t_sigma, t.shape)
    return dict(img_pts_arr=img_pts, fnames_arr=fnames, board_shape=board_shape, square=square, k=k, d=d, r=r, t=t,
                r0=r0, t0=t0, gt=np.array(gt_pts), cam_res=res)
