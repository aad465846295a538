"""ctypes binding of include/acino_b200.h (libacino_b200.so, built in-tree by build.sh).

There is no CPU fallback: a missing library or a box without a CUDA device raises.
"""
import ctypes as ct
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'libacino_b200.so')
CAM_STRIDE = 24
MAX_CAMS = 8
MODE_POINTS_ONLY = 0
MODE_POINTS_AND_EXTRINSICS = 1


class AcinoError(RuntimeError):
    pass


class SbaOpts(ct.Structure):
    _fields_ = [('mode', ct.c_int32), ('max_iter', ct.c_int32), ('inner_iters', ct.c_int32), ('force_iters', ct.c_int32),
                ('f_scale', ct.c_double), ('gtol', ct.c_double), ('xtol', ct.c_double), ('lambda0', ct.c_double),
                ('ftol', ct.c_double), ('hess_floor', ct.c_double), ('point_refine_iters', ct.c_double), ('point_refine_skip', ct.c_double), ('polish_tol', ct.c_double),
                ('lambda_dec', ct.c_double)]


class SbaStats(ct.Structure):
    _fields_ = [('n_points', ct.c_int64), ('n_obs', ct.c_int64), ('n_unconverged', ct.c_int64),
                ('iterations', ct.c_int32), ('launches', ct.c_int32),
                ('cost_initial', ct.c_double), ('cost_final', ct.c_double), ('solve_ms', ct.c_double),
                ('status', ct.c_double), ('n_accept', ct.c_double), ('lambda_final', ct.c_double), ('gmax', ct.c_double),
                ('init_ms', ct.c_double), ('polish_ms', ct.c_double), ('syncs', ct.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if k != 'reserved'}


_P = ct.c_void_p
_I64 = ct.c_int64
_D = ct.c_double
_I = ct.c_int

# name -> (restype, argtypes); every symbol include/acino_b200.h declares
SIGNATURES = {
    'acino_last_error': (ct.c_char_p, []),
    'acino_version': (_I, []),
    'acino_device_count': (_I, []),
    'acino_set_device': (_I, [_I]),
    'acino_dev_alloc': (_I, [ct.POINTER(_P), _I64]),
    'acino_dev_free': (_I, [_P]),
    'acino_h2d': (_I, [_P, _P, _I64]),
    'acino_d2h': (_I, [_P, _P, _I64]),
    'acino_sync': (_I, []),
    'acino_stream_create': (_I, [ct.POINTER(_P)]),
    'acino_stream_destroy': (_I, [_P]),
    'acino_stream_sync': (_I, [_P]),
    'acino_h2d_async': (_I, [_P, _P, _I64, _P]),
    'acino_d2h_async': (_I, [_P, _P, _I64, _P]),
    'acino_project': (_I, [_P, _I64, _P, _P]),
    'acino_project_dev': (_I, [_P, _I64, _P, _P, _P]),
    'acino_project_all': (_I, [_P, _I64, _P, _I, _D, _D, _P]),
    'acino_residuals_list': (_I, [_P, _I64, _P, _P, _P, _I64, _P, _I, _P]),
    'acino_ekf_measurement': (_I, [_P, _I, _I, _P, _I, _D, _I, _P, _P, _P]),
    'acino_manual_residuals': (_I, [_P, _I, _P, _I, _P, _I, _P, _I, _P, _I, _D, _D, _D, _P]),
    'acino_ring_critical': (_I, [_P, _I64, _I, _P]),
    'acino_residual_cost_noalloc': (_I, [_P, _I64, _P, _P, _P, _I64, _P, _I, _D, _P, _P, _P]),
    'acino_jacobian_list': (_I, [_P, _I64, _P, _P, _I64, _P, _I, _P]),
    'acino_rodrigues': (_I, [_P, _I, _P]),
    'acino_triangulate_pair': (_I, [_P, _P, _I64, _P, _P, _P]),
    'acino_triangulate_ring': (_I, [_P, _P, _D, _I64, _I, _P, _P, _P, _P]),
    'acino_triangulate_ring_dev': (_I, [_P, _P, _D, _I64, _I, _P, _P, _P, _P, _P]),
    'acino_sba_default_opts': (None, [ct.POINTER(SbaOpts), _I]),
    'acino_sba_create': (_P, [_I, _I64]),
    'acino_sba_destroy': (None, [_P]),
    'acino_sba_set_cameras': (_I, [_P, _P]),
    'acino_sba_set_observations': (_I, [_P, _P, _P]),
    'acino_sba_set_observations_dev': (_I, [_P, _P, _P]),
    'acino_sba_set_points': (_I, [_P, _P]),
    'acino_sba_set_points_dev': (_I, [_P, _P]),
    'acino_sba_get_points': (_I, [_P, _P]),
    'acino_sba_get_points_dev': (_I, [_P, _P]),
    'acino_sba_get_cameras': (_I, [_P, _P]),
    'acino_sba_get_rvec_t': (_I, [_P, _P, _P]),
    'acino_sba_set_rvec_t': (_I, [_P, _P, _P]),
    'acino_sba_residuals': (_I, [_P, _P]),
    'acino_sba_solve': (_I, [_P, ct.POINTER(SbaOpts), ct.POINTER(SbaStats)]),
    'acino_sba_iterate': (_I, [_P, ct.POINTER(SbaOpts)]),
    'acino_sba_reset_state': (_I, [_P]),
    'acino_sba_stream': (_P, [_P]),
    'acino_sba_profile': (_I, [_P, _I]),
    'acino_sba_profile_read': (_I, [_P, ct.POINTER(_D), ct.POINTER(_D), ct.POINTER(_I)]),
    'acino_sba_profile_read_kernels': (_I, [_P, _P, ct.POINTER(_I)]),
    'acino_sba_lm_state': (_I, [_P, _P]),
    'acino_sba_history': (_I, [_P, _P, _I, ct.POINTER(_I)]),
    'acino_fp64_peak': (_I, [ct.POINTER(_D)]),
    'acino_nccl_unique_id': (_I, [_P]),
    'acino_sba_comm_init': (_I, [_P, _P, _I, _I]),
    'acino_sba_comm_mode': (_I, [_P]),
}

_lib = None


def lib():
    """Loads the shared library (once). Raises AcinoError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise AcinoError(f'{LIB_PATH} is missing: run ./build.sh (or __graft_entry__.build()); there is no CPU fallback')
        l = ct.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        raise AcinoError(f'acino_b200 error {rc}: {lib().acino_last_error().decode()}')


def require_gpu():
    n = lib().acino_device_count()
    if n <= 0:
        raise AcinoError('no CUDA device visible: the acinoset_b200 hot path is CUDA-only (sm_100a), there is no CPU fallback')
    return n


def ptr(a):
    return a.ctypes.data_as(ct.c_void_p) if a is not None else None


def f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.reshape(shape) if shape is not None else a


def pack_cam(k, d, r, t):
    """One cam24 block from reference-shaped k (3,3), d (4,)/(4,1), r (3,3), t (3,)/(3,1)."""
    k = np.asarray(k, dtype=np.float64).reshape(3, 3)
    d = np.asarray(d, dtype=np.float64).reshape(-1)
    if d.size != 4:
        raise ValueError(f'fisheye model needs 4 distortion coefficients, got {d.size}')
    c = np.zeros(CAM_STRIDE)
    c[0], c[1], c[2], c[3], c[4] = k[0, 0], k[1, 1], k[0, 2], k[1, 2], k[0, 1] / k[0, 0]
    c[5:9] = d
    c[9:18] = np.asarray(r, dtype=np.float64).reshape(9)
    c[18:21] = np.asarray(t, dtype=np.float64).reshape(3)
    return c


def pack_cams(k_arr, d_arr, r_arr, t_arr):
    n = len(k_arr)
    d_arr = np.asarray(d_arr, dtype=np.float64).reshape(n, -1)
    return np.stack([pack_cam(k_arr[i], d_arr[i], r_arr[i], t_arr[i]) for i in range(n)])
