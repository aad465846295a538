"""TEST INFRASTRUCTURE ONLY — golden vectors for the SURVEY.md §8(f) N3 callers, produced by the REAL reference functions:

  adjust_extrinsics_manual_points (/root/reference/src/lib/calib.py:216-297): run unmodified; scipy's least_squares is
      wrapped for the duration of the call only to record the closure's residual vector at two parameter points;
  initialize_marker_3d (/root/reference/src/lib/app.py:39-84): lib/app.py cannot be imported here (it pulls in the
      OpenGL viewer and video modules), so the function's source text is cut out of the file at generation time and
      executed unmodified in a namespace holding the reference's own utils / calib / points modules.

    python -m oracle.make_golden_n3
"""
import ast
import os
import numpy as np

from . import ref_shim
from .make_golden import GOLD, _versions


def gen_adjust():
    calib, utils, sba = ref_shim.load()
    from acinoset_b200 import synth
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    case = synth.make_manual_points_case(14, seed=3, rig=rig)
    probe = np.array([0.01, -0.02, 0.005, 0.03, -0.01, 0.02])
    rec = {}
    real = calib.least_squares

    def spy(fun, x0, *a, **kw):
        rec['f0'] = np.array(fun(np.asarray(x0, dtype=np.float64)))
        rec['f_probe'] = np.array(fun(probe))
        res = real(fun, x0, *a, **kw)
        rec['x'] = res.x.copy(); rec['cost'] = res.cost; rec['nfev'] = res.nfev; rec['fun'] = res.fun.copy()
        return res

    calib.least_squares = spy
    try:
        r_in = case['r_bad'].copy(); t_in = case['t_bad'].copy()
        with ref_shim.quiet():
            r_out, t_out = calib.adjust_extrinsics_manual_points(calib.calibrate_pair_extrinsics_fisheye, case['img_pts'],
                                                                 case['bad_cams'], case['k'], case['d'], r_in, t_in)
    finally:
        calib.least_squares = real
    np.savez_compressed(os.path.join(GOLD, 'golden_manual_adjust.npz'), probe=probe, f0=rec['f0'], f_probe=rec['f_probe'],
                        x=rec['x'], cost=rec['cost'], nfev=rec['nfev'], fun=rec['fun'], r_out=np.array(r_out), t_out=np.array(t_out),
                        versions=_versions())
    err_before = np.abs(case['r_bad'] - case['r_true']).max()
    err_after = np.abs(np.array(r_out) - case['r_true']).max()
    print(f"manual adjust: {len(rec['f0'])} residuals, cost {rec['cost']:.6e}, nfev {rec['nfev']}, |dR| {err_before:.2e} -> {err_after:.2e}")


def _cut_function(path, name):
    src = open(path).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            return '\n'.join(src.splitlines()[node.lineno - 1:node.end_lineno])
    raise KeyError(name)


def gen_init_marker():
    calib, utils, sba = ref_shim.load()
    from lib import points
    from acinoset_b200 import synth
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    sc = synth.make_track_scene(60, 20, seed=2, rig=rig)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=10)
    ns = dict(np=np, utils=utils, triangulate_points_fisheye=calib.triangulate_points_fisheye, EOM_curve_fit=points.EOM_curve_fit,
              plot_marker_3d=None)
    exec(compile(_cut_function(os.path.join(ref_shim.REF_SRC, 'lib', 'app.py'), 'initialize_marker_3d'), 'reference:lib/app.py', 'exec'), ns)
    out = {}
    import io, contextlib
    for marker in ('nose', 'spine'):
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            fit, fit_deriv = ns['initialize_marker_3d'](df, marker, sc['k'], sc['d'].reshape((-1, 4)), sc['r'], sc['t'])
        line = [l for l in buf.getvalue().splitlines() if l.startswith('Initializing')][0]
        out[f'{marker}_fit'] = fit; out[f'{marker}_fit_deriv'] = fit_deriv; out[f'{marker}_thresh_line'] = line
        print(line)
    np.savez_compressed(os.path.join(GOLD, 'golden_init_marker.npz'), versions=_versions(), **out)




def gen_dlc_load():
    """N1: the reference's load_dlc_points_as_df (/root/reference/src/lib/utils.py:77-151) on synthetic DeepLabCut tables.
    PyTables is not in the image, so pd.read_hdf is pointed at in-memory tables for the duration of the call (harness
    shim); the reshaping, frame shifts and fills are the reference's own code."""
    import pandas as pd
    calib, utils, sba = ref_shim.load()
    from acinoset_b200 import synth
    markers = ['nose', 'r_eye', 'l_eye', 'neck_base', 'spine', 'tail_base']
    tabs = synth.make_dlc_tables(12, markers, 4, seed=7, first_frame=3)
    paths = [f'/data/cam{i + 1}DLC.h5' for i in range(4)]
    store = dict(zip(paths, tabs))
    real = pd.read_hdf
    pd.read_hdf = lambda p, *a, **k: store[p].copy()
    out = {}
    try:
        for name, shifts in (('noshift', None), ('shift', [0, 2, -1, 0])):
            with ref_shim.quiet():
                df = utils.load_dlc_points_as_df(paths, frame_shifts=shifts)
            out[f'{name}_frame'] = df['frame'].to_numpy().astype(np.int64)
            out[f'{name}_camera'] = df['camera'].to_numpy().astype(np.int64)
            out[f'{name}_marker'] = df['marker'].to_numpy().astype(str)
            for col in ('x', 'y', 'likelihood'):
                out[f'{name}_{col}'] = df[col].to_numpy().astype(np.float64)
        # tables without a likelihood column (utils.py:104-118: likelihood from y) and the 'dlc_head' tables
        # (utils.py:84-103: likelihood from x, three markers kept); cells that are NaN in x only / in y only
        for name, tabs2, paths2 in (('nolik',) + synth.make_dlc_xy_tables(9, ['nose', 'r_eye', 'spine'], 3, seed=11),
                                    ('head',) + synth.make_dlc_head_tables(7, 2, seed=12)):
            store.clear()
            store.update(dict(zip(paths2, tabs2)))
            with ref_shim.quiet():
                df = utils.load_dlc_points_as_df(paths2)
            out[f'{name}_frame'] = df['frame'].to_numpy().astype(np.int64)
            out[f'{name}_camera'] = df['camera'].to_numpy().astype(np.int64)
            out[f'{name}_marker'] = df['marker'].to_numpy().astype(str)
            for col in ('x', 'y', 'likelihood'):
                out[f'{name}_{col}'] = df[col].to_numpy().astype(np.float64)
    finally:
        pd.read_hdf = real
    np.savez_compressed(os.path.join(GOLD, 'golden_dlc_load.npz'), versions=_versions(), **out)
    print('dlc load:', len(out['noshift_frame']), 'rows')


def _cut_nested(path, outer, names):
    """Source text of the functions `names` nested inside the top-level function `outer`, dedented."""
    import textwrap
    src = open(path).read()
    lines = src.splitlines()
    out = []
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == outer:
            for sub in node.body:
                if isinstance(sub, ast.FunctionDef) and sub.name in names:
                    out.append(textwrap.dedent('\n'.join(lines[sub.lineno - 1:sub.end_lineno])))
    assert len(out) == len(names)
    return '\n\n'.join(out)


def ekf_pose(mode, seed):
    """A plausible pose for marker mode `mode`: head inside the rig's view volume, moderate joint angles."""
    from acinoset_b200 import ekf as fast
    rng = np.random.default_rng(seed)
    x = rng.uniform(-0.4, 0.4, fast.N_POSE[mode])
    x[0:3] = [2.0 + rng.uniform(-1, 1), 6.5 + rng.uniform(-0.5, 0.5), 0.6]
    if mode != 'head':
        x[6] = -0.28                         # neck length l_1
    if mode == 'default':
        x[26:29] = [3.0, 7.0, 0.1]           # lure
    return x


def gen_ekf():
    """N4: h_function and numerical_jacobian of the reference's EKF (/root/reference/src/core/ekf.py:66-96) — nested
    functions of ekf(), so their source text is cut out and executed unmodified with the reference's lib.misc and
    lib.calib — assembled per camera exactly as the loop at :245-251 does, for float64 poses. (Inside the reference's
    loop the pose is float32 because predict_next_state casts it, so its own H carries ~1e-4 relative noise; a float32
    run is stored too, for the record.)"""
    calib, utils, sba = ref_shim.load()
    from lib import misc
    from acinoset_b200 import synth, ekf as fast
    rig = synth.load_rig(os.path.join(GOLD, 'scene_6cam.json'))
    k, d, r, t, _ = rig
    text = _cut_nested(os.path.join(ref_shim.REF_SRC, 'core', 'ekf.py'), 'ekf', ['h_function', 'numerical_jacobian'])
    out = {}
    for mode in ('default', 'head', 'upper_body', 'head_stabilize'):
        ns = dict(np=np, misc=misc, project_points_fisheye=calib.project_points_fisheye, marker_mode=mode)
        exec(compile(text, 'reference:core/ekf.py', 'exec'), ns)
        x = ekf_pose(mode, seed=5)
        n_mk = len(misc.get_markers(mode=mode))
        for tag, xx in (('f64', x), ('f32', x.astype(np.float32))):
            h = np.zeros(6 * n_mk * 2); H = np.zeros((6 * n_mk * 2, 3 * len(x)))
            for j in range(6):
                h[j * n_mk * 2:(j + 1) * n_mk * 2] = ns['h_function'](xx, k[j], d[j], r[j], t[j]).flatten()
                H[j * n_mk * 2:(j + 1) * n_mk * 2, 0:len(x)] = ns['numerical_jacobian'](ns['h_function'], xx, k[j], d[j], r[j], t[j])
            out[f'{mode}_{tag}_h'] = h; out[f'{mode}_{tag}_H'] = H
        out[f'{mode}_x'] = x
        out[f'{mode}_coords'] = np.array(misc.get_3d_marker_coords({'x': x}, mode=mode), dtype=np.float64)
    np.savez_compressed(os.path.join(GOLD, 'golden_ekf.npz'), versions=_versions(), **out)
    print('ekf:', {m: out[f'{m}_f64_H'].shape for m in ('default', 'head', 'upper_body', 'head_stabilize')})


if __name__ == '__main__':
    if not ref_shim.available():
        raise SystemExit('reference tree not present; fixtures can only be generated in the build container')
    gen_adjust()
    gen_init_marker()
    gen_dlc_load()
    gen_ekf()
