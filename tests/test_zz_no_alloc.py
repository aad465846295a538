"""GPU test of acino_residual_cost_noalloc (include/acino_b200.h), the allocation-free entry point added:
residual vector + robust cost with the scratch in a file-scope device array of the library.

Collected last on purpose (the file name sorts after every other test file): a
failure here cannot hide another test. The entry point is called through the C ABI twice in a row on one stream - the
second call with other points and a shorter list, so that anything the shared scratch array kept from the first call
would show - and compared with the existing entry point acino_residuals_list (1e-9 relative on the residuals, the
tolerance of the other residual tests; 1e-12 relative on the cost computed from the existing entry point's residuals).
"""
import ctypes as ct
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

from acinoset_b200 import _capi, dense, synth

MAX_OBS = 4194304          # the stated limit of the entry point


def _problem(seed, n_points, rig):
    k, d, r, t, _ = rig
    C = len(k)
    rng = np.random.default_rng(seed)
    X = np.stack([rng.uniform(-1, 5, n_points), rng.uniform(5, 8, n_points), rng.uniform(0, 1, n_points)], 1)
    sel = rng.uniform(size=(n_points, C)) < 0.6
    pidx, cidx = np.nonzero(sel)
    pidx, cidx = pidx.astype(np.int32), cidx.astype(np.int32)
    cams = _capi.pack_cams(k, d, r, t)
    proj = np.empty((len(pidx), 2))
    zero = np.zeros((len(pidx), 2))
    _capi.check(_capi.lib().acino_residuals_list(_capi.ptr(X), n_points, _capi.ptr(pidx), _capi.ptr(cidx), _capi.ptr(zero), len(pidx),
                                                 _capi.ptr(cams), C, _capi.ptr(proj)))
    obs = proj + rng.normal(0, 1.0, proj.shape)
    obs[rng.uniform(size=len(obs)) < 0.05] += rng.normal(0, 50.0, 2)          # a few gross outliers: the robust part of the cost
    return X, pidx, cidx, obs, cams, C


def _existing(X, pidx, cidx, obs, cams, C):
    f = np.empty(2 * len(pidx))
    _capi.check(_capi.lib().acino_residuals_list(_capi.ptr(X), len(X), _capi.ptr(pidx), _capi.ptr(cidx), _capi.ptr(obs), len(pidx),
                                                 _capi.ptr(cams), C, _capi.ptr(f)))
    return f


def test_residual_cost_noalloc_matches_existing_entry_point_twice_in_a_row(rig):
    _capi.require_gpu()
    L = _capi.lib()
    X, pidx, cidx, obs, cams, C = _problem(41, 9000, rig)
    n = len(pidx)
    assert n > 30000          # > 100 blocks: the fold by the last block is exercised
    f_scale = 50.0
    X_d, p_d, c_d, o_d = (dense.DeviceArray.from_host(a) for a in (X, pidx, cidx, obs))
    f_d = dense.DeviceArray((n, 2), np.float64)
    cost_d = dense.DeviceArray((1,), np.float64)
    # second problem: moved points, a shorter list (fewer blocks than the first call left partial sums for)
    X2 = X + np.random.default_rng(1).normal(0, 0.01, X.shape)
    n2 = n // 3 + 7
    X2_d = dense.DeviceArray.from_host(X2)
    f2_d = dense.DeviceArray((n2, 2), np.float64)
    cost2_d = dense.DeviceArray((1,), np.float64)
    f_d.upload(np.full((n, 2), -7.0)); f2_d.upload(np.full((n2, 2), -7.0))
    cost_d.upload(np.array([-1.0])); cost2_d.upload(np.array([-1.0]))

    # two calls in a row on one stream (the default one), no synchronisation in between
    _capi.check(L.acino_residual_cost_noalloc(X_d.ptr, len(X), p_d.ptr, c_d.ptr, o_d.ptr, n, _capi.ptr(cams), C, f_scale,
                                              f_d.ptr, cost_d.ptr, None))
    _capi.check(L.acino_residual_cost_noalloc(X2_d.ptr, len(X), p_d.ptr, c_d.ptr, o_d.ptr, n2, _capi.ptr(cams), C, f_scale,
                                              f2_d.ptr, cost2_d.ptr, None))
    _capi.check(L.acino_sync())

    for Xh, m, fd, cd in ((X, n, f_d, cost_d), (X2, n2, f2_d, cost2_d)):
        want = _existing(Xh, pidx[:m], cidx[:m], obs[:m], cams, C)
        got = fd.download().ravel()
        assert (np.abs(got - want) / np.maximum(1.0, np.abs(want))).max() < 1e-9
        cost_want = 0.5 * f_scale ** 2 * np.sum(np.log1p((want / f_scale) ** 2))
        cost_got = float(cd.download()[0])
        assert abs(cost_got - cost_want) <= 1e-12 * cost_want, (cost_got, cost_want)

    # a third call, repeating the first: same bits (the sum is deterministic and nothing of call two is left behind)
    first = (f_d.download().copy(), float(cost_d.download()[0]))
    _capi.check(L.acino_residual_cost_noalloc(X_d.ptr, len(X), p_d.ptr, c_d.ptr, o_d.ptr, n, _capi.ptr(cams), C, f_scale,
                                              f_d.ptr, cost_d.ptr, None))
    _capi.check(L.acino_sync())
    assert np.array_equal(f_d.download(), first[0]) and float(cost_d.download()[0]) == first[1]

    # cost only (no residual output), and the empty list
    _capi.check(L.acino_residual_cost_noalloc(X_d.ptr, len(X), p_d.ptr, c_d.ptr, o_d.ptr, n, _capi.ptr(cams), C, f_scale,
                                              None, cost2_d.ptr, None))
    _capi.check(L.acino_sync())
    assert float(cost2_d.download()[0]) == first[1]
    _capi.check(L.acino_residual_cost_noalloc(None, 0, None, None, None, 0, _capi.ptr(cams), C, f_scale, None, cost2_d.ptr, None))
    _capi.check(L.acino_sync())
    assert float(cost2_d.download()[0]) == 0.0
    for a in (X_d, p_d, c_d, o_d, f_d, cost_d, X2_d, f2_d, cost2_d):
        a.free()


def test_residual_cost_noalloc_rejects_sizes_beyond_its_scratch(rig):
    _capi.require_gpu()
    L = _capi.lib()
    k, d, r, t, _ = rig
    cams = _capi.pack_cams(k, d, r, t)
    cost_d = dense.DeviceArray((1,), np.float64)
    dummy = dense.DeviceArray((8,), np.float64)
    rc = L.acino_residual_cost_noalloc(dummy.ptr, 1, dummy.ptr, dummy.ptr, dummy.ptr, MAX_OBS + 1, _capi.ptr(cams), len(cams), 1.0,
                                       None, cost_d.ptr, None)
    assert rc == -1          # ACINO_EINVAL
    msg = L.acino_last_error().decode()
    assert 'exceeds' in msg and str(MAX_OBS) in msg
    rc = L.acino_residual_cost_noalloc(dummy.ptr, 1, dummy.ptr, dummy.ptr, dummy.ptr, 1, _capi.ptr(cams), 9, 1.0, None, cost_d.ptr, None)
    assert rc == -1 and 'cameras' in L.acino_last_error().decode()
    cost_d.free(); dummy.free()
