"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/acino_b200.h declares, and the
host-side lowering (DataFrame / observation lists -> dense slots) is exact. No compute calls (no GPU here)."""
import os
import re
import numpy as np
import pandas as pd
import pytest

from acinoset_b200 import _capi, utils, sba, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, 'include', 'acino_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(acino_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    L = _capi.lib()
    names = _header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f'{n} declared in include/acino_b200.h but not exported'
    assert set(names) == set(_capi.SIGNATURES), 'ctypes signature table and header disagree'


def test_no_gpu_means_loud_failure():
    if _capi.lib().acino_device_count() > 0:
        pytest.skip('a GPU is visible')
    with pytest.raises(_capi.AcinoError):
        _capi.require_gpu()
    from acinoset_b200 import calib
    k, d, r, t, _ = synth.load_rig()
    with pytest.raises(_capi.AcinoError):
        calib.project_points_fisheye(np.zeros((1, 3)), k[0], d[0], r[0], t[0])


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'acinoset_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            assert 'oracle' not in open(os.path.join(pkg, fn)).read(), fn


def test_pack_cam_layout(rig):
    k, d, r, t, _ = rig
    c = _capi.pack_cam(k[2], d[2], r[2], t[2])
    assert c.shape == (24,)
    assert c[0] == k[2][0, 0] and c[1] == k[2][1, 1] and c[2] == k[2][0, 2] and c[3] == k[2][1, 2]
    assert (c[5:9] == d[2].ravel()).all() and (c[9:18] == r[2].ravel()).all() and (c[18:21] == t[2].ravel()).all()
    assert (_capi.pack_cam(k[2], d[2].reshape(4), r[2], t[2].reshape(3)) == c).all()
    with pytest.raises(ValueError):
        _capi.pack_cam(k[2], np.zeros(5), r[2], t[2])


def test_df_to_dense_order_and_masking():
    """Point order = sorted (frame, marker name) (groupby at src/lib/utils.py:348); strict > threshold applied by caller."""
    sc = synth.make_rig_scene(4, 20, seed=2)
    df = synth.dense_to_df(sc['uv'], sc['lik'], 20, first_frame=10)
    dft = df.query('likelihood > 0.5')
    uv, present, frames, markers, slot = utils.df_to_dense(dft, 6)
    keys = sorted(set(zip(dft['frame'], dft['marker'])))
    assert list(zip(frames, markers)) == keys
    assert present.sum() == len(dft)
    # every row lands in its slot
    flat = uv.reshape(-1, 2)
    assert (flat[slot, 0] == dft['x'].to_numpy()).all() and (flat[slot, 1] == dft['y'].to_numpy()).all()
    # against the generator's mask
    m = sc['lik'] > 0.5
    assert present.sum() == m.sum()


def test_df_to_dense_rejects_duplicates_and_bad_cameras():
    df = pd.DataFrame({'frame': [1, 1], 'camera': [0, 0], 'marker': ['a', 'a'], 'x': [0., 1.], 'y': [0., 1.]})
    with pytest.raises(ValueError):
        utils.df_to_dense(df, 6)
    df = pd.DataFrame({'frame': [1], 'camera': [7], 'marker': ['a'], 'x': [0.], 'y': [0.]})
    with pytest.raises(ValueError):
        utils.df_to_dense(df, 6)
    out = utils.df_to_dense(df.iloc[:0], 6)
    assert out[0].shape == (0, 6, 2)


def test_lists_to_dense_roundtrip():
    rng = np.random.default_rng(0)
    P, C = 50, 6
    pairs = [(p, c) for p in range(P) for c in range(C) if rng.uniform() < 0.5]
    rng.shuffle(pairs)
    pidx = np.array([p for p, _ in pairs]); cidx = np.array([c for _, c in pairs])
    obs = rng.normal(size=(len(pairs), 2))
    uv, mask, slot = sba._lists_to_dense(obs, pidx, cidx, P, C)
    assert (uv.reshape(-1, 2)[slot] == obs).all()
    want = np.zeros(P, dtype=np.uint8)
    for p, c in pairs:
        want[p] |= 1 << c
    assert (mask == want).all()
    with pytest.raises(NotImplementedError):
        sba._lists_to_dense(np.zeros((2, 2)), np.array([0, 0]), np.array([1, 1]), P, C)
    with pytest.raises(ValueError):
        sba._lists_to_dense(np.zeros((1, 2)), np.array([P]), np.array([0]), P, C)


def test_rotation_matrix_to_rvec_matches_opencv():
    import cv2
    rng = np.random.default_rng(1)
    for _ in range(50):
        v = rng.normal(0, 1.2, 3)
        R = cv2.Rodrigues(v)[0]
        assert np.abs(sba.rotation_matrix_to_rvec(R) - cv2.Rodrigues(R)[0].ravel()).max() < 1e-9
    assert np.abs(sba.rotation_matrix_to_rvec(np.eye(3))).max() == 0.0


def test_install_rebinds_only_names_the_reference_defines():
    """INTEGRATION.md drop-in: every attribute app.install() sets exists in the reference module it patches (checked
    against the reference sources when the tree is present, i.e. in the build container)."""
    import ast
    import os
    import types
    import pytest
    ref = '/root/reference/src/lib'
    if not os.path.isdir(ref):
        pytest.skip('reference tree not present')
    import acinoset_b200.app as fast
    mods = {n: types.ModuleType(n) for n in ('calib', 'utils', 'sba', 'app', 'metric')}
    fast.install(mods['calib'], mods['utils'], mods['sba'], mods['app'], mods['metric'])
    def names_of(mod_name, follow_star=True):
        tree = ast.parse(open(os.path.join(ref, mod_name + '.py')).read())
        defined = set()
        for node in tree.body:
            if isinstance(node, ast.FunctionDef):
                defined.add(node.name)
            elif isinstance(node, ast.ImportFrom):
                for a in node.names:
                    if a.name == '*' and follow_star and node.level == 1:       # `from .utils import *` (src/lib/sba.py:6)
                        defined |= names_of(node.module, follow_star=False)
                    else:
                        defined.add(a.asname or a.name)
        return defined

    for name, m in mods.items():
        defined = names_of(name)
        patched = {k for k in vars(m) if not k.startswith('__')}
        assert patched and patched <= defined, (name, patched - defined)


def test_df_to_dense_key_paths_agree():
    """utils._point_keys takes a short cut for integer frames in a compact range and factorises the marker column as the
    pandas column it is; object-dtype markers, widely scattered frame numbers and float frames go the general way. All
    must give the sorted (frame, marker-name) point order of groupby(['frame', 'marker']) (src/lib/utils.py:348)."""
    rng = np.random.default_rng(5)
    n = 4000
    frames = rng.integers(10, 60, n)
    markers = rng.choice(np.array(['nose', 'l_eye', 'tail1', 'spine', 'r_hip'], dtype=object), n)
    cams = rng.integers(0, 6, n)
    df = pd.DataFrame({'frame': frames, 'camera': cams, 'marker': markers, 'x': rng.normal(size=n), 'y': rng.normal(size=n)})
    df = df.drop_duplicates(['frame', 'camera', 'marker']).reset_index(drop=True)
    want_order = sorted(set(zip(df['frame'].tolist(), df['marker'].tolist())))
    variants = {
        'compact int frames': df,
        'object markers': df.assign(marker=df['marker'].astype(object)),
        'scattered frames': df.assign(frame=df['frame'] * 1_000_003),
        'float frames': df.assign(frame=df['frame'].astype(np.float64)),
    }
    ref = None
    for name, d in variants.items():
        uv, present, f, m, slot, codes = utils.df_to_dense(d, 6, return_keys=True)
        scale = 1_000_003 if name == 'scattered frames' else 1
        got_order = list(zip((np.asarray(f) // scale).astype(int).tolist(), [str(x) for x in m]))
        assert got_order == [(int(a), str(b)) for a, b in want_order], name
        if ref is None:
            ref = (uv, present, slot, codes)
        else:
            assert np.array_equal(uv, ref[0]) and np.array_equal(present, ref[1]), name
            assert np.array_equal(slot, ref[2]) and np.array_equal(codes, ref[3]), name
        # every row lands in the slot of its own (frame, marker, camera)
        rows = np.arange(len(d))
        assert np.array_equal(uv.reshape(-1, 2)[slot], d[['x', 'y']].to_numpy())
        assert np.array_equal(slot // 6, codes) and np.array_equal(slot % 6, d['camera'].to_numpy())


def test_struct_mirrors_follow_the_header():
    """The ctypes mirrors of acino_sba_opts / acino_sba_stats list the header's fields in the header's order with the
    header's types (a field added on one side only would shift everything behind it silently)."""
    import re
    import ctypes as ct
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'include', 'acino_b200.h')).read()
    ctype = {'int32_t': ct.c_int32, 'int64_t': ct.c_int64, 'double': ct.c_double}

    def fields(struct_name):
        end = hdr.index('} ' + struct_name + ';')
        body = hdr[hdr.rindex('typedef struct {', 0, end) + len('typedef struct {'):end]
        body = re.sub(r'/\*.*?\*/', '', body, flags=re.S)
        out = []
        for decl in body.split(';'):
            decl = decl.strip()
            if not decl:
                continue
            ty, names = decl.split(None, 1)
            out += [(n.strip(), ctype[ty]) for n in names.split(',')]
        return out

    for struct_name, mirror in (('acino_sba_opts', _capi.SbaOpts), ('acino_sba_stats', _capi.SbaStats)):
        want = fields(struct_name)
        got = [(n, t) for n, t in mirror._fields_]
        assert got == want, (struct_name, got, want)
