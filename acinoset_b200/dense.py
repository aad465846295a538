"""Dense-slot entry points: the array layout the CUDA kernels consume.

Point p = (frame, keypoint); slot (p, c) = camera c's observation of it. ``uv`` is (P, C, 2) float64,
``lik`` (P, C) float64 likelihoods, ``obs_mask`` (P,) uint8 bitmask. These are what the DataFrame / list
front-ends in utils.py and sba.py lower to, and what bench.py drives directly at full scale.
"""
import ctypes as ct
import numpy as np
from . import _capi
from ._capi import SbaOpts, SbaStats, MODE_POINTS_ONLY, MODE_POINTS_AND_EXTRINSICS


def triangulate_ring(uv, lik, thresh, k_arr, d_arr, r_arr, t_arr):
    """Ring-pair DLT of src/lib/utils.py:319-349 fused with `likelihood > thresh` (src/core/tri.py:41).
    Returns xyz (P,3) (NaN where absent), exists (P,) bool, obs_mask (P,) uint8."""
    _capi.require_gpu()
    uv = _capi.f64(uv)
    P, C = uv.shape[0], uv.shape[1]
    lik = _capi.f64(lik).reshape(P, C)
    cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
    xyz = np.empty((P, 3))
    mask = np.empty(P, dtype=np.uint8)
    exists = np.empty(P, dtype=np.uint8)
    _capi.check(_capi.lib().acino_triangulate_ring(_capi.ptr(uv), _capi.ptr(lik), float(thresh), P, C, _capi.ptr(cams),
                                                   _capi.ptr(xyz), _capi.ptr(mask), _capi.ptr(exists)))
    return xyz, exists.astype(bool), mask


class DeviceArray:
    """A caller-owned HBM buffer (acino_dev_alloc) with numpy upload / download."""

    def __init__(self, shape, dtype):
        self.shape = tuple(int(s) for s in np.atleast_1d(shape))
        self.dtype = np.dtype(dtype)
        self.nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = ct.c_void_p()
        _capi.check(_capi.lib().acino_dev_alloc(ct.byref(p), self.nbytes))
        self.ptr = p

    @classmethod
    def from_host(cls, a):
        a = np.ascontiguousarray(a)
        d = cls(a.shape, a.dtype)
        d.upload(a)
        return d

    def upload(self, a):
        a = np.ascontiguousarray(a, dtype=self.dtype)
        assert a.nbytes == self.nbytes
        _capi.check(_capi.lib().acino_h2d(self.ptr, _capi.ptr(a), self.nbytes))

    def download(self, out=None):
        if out is None:
            out = np.empty(self.shape, dtype=self.dtype)
        _capi.check(_capi.lib().acino_d2h(_capi.ptr(out), self.ptr, self.nbytes))
        return out

    def free(self):
        if self.ptr is not None and self.ptr.value:
            _capi.lib().acino_dev_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def default_opts(mode=MODE_POINTS_ONLY, **kw):
    o = SbaOpts()
    _capi.lib().acino_sba_default_opts(ct.byref(o), mode)
    for k, v in kw.items():
        if not hasattr(o, k):
            raise TypeError(f'unknown solver option {k}')
        setattr(o, k, v)
    return o


class SbaSolver:
    """One solver handle = one GPU = one stream (include/acino_b200.h: acino_sba_*)."""

    def __init__(self, n_cams, n_points):
        _capi.require_gpu()
        self.C, self.P = int(n_cams), int(n_points)
        self._h = _capi.lib().acino_sba_create(self.C, self.P)
        if not self._h:
            raise _capi.AcinoError('acino_sba_create: ' + _capi.lib().acino_last_error().decode())
        self._keep = []

    def close(self):
        if self._h:
            _capi.lib().acino_sba_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_cameras(self, k_arr, d_arr, r_arr, t_arr):
        cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
        assert cams.shape[0] == self.C
        _capi.check(_capi.lib().acino_sba_set_cameras(self._h, _capi.ptr(cams)))

    def set_observations(self, uv, obs_mask):
        uv = _capi.f64(uv).reshape(self.P, self.C, 2)
        m = np.ascontiguousarray(obs_mask, dtype=np.uint8).reshape(self.P)
        _capi.check(_capi.lib().acino_sba_set_observations(self._h, _capi.ptr(uv), _capi.ptr(m)))

    def set_observations_dev(self, uv_dev: DeviceArray, mask_dev: DeviceArray):
        self._keep = [uv_dev, mask_dev]
        _capi.check(_capi.lib().acino_sba_set_observations_dev(self._h, uv_dev.ptr, mask_dev.ptr))

    def set_points(self, xyz):
        x = _capi.f64(xyz).reshape(self.P, 3)
        _capi.check(_capi.lib().acino_sba_set_points(self._h, _capi.ptr(x)))

    def set_points_dev(self, xyz_dev: DeviceArray):
        _capi.check(_capi.lib().acino_sba_set_points_dev(self._h, xyz_dev.ptr))

    def set_rvec_t(self, rvecs, tvecs):
        rv = _capi.f64(rvecs).reshape(self.C, 3)
        tv = _capi.f64(tvecs).reshape(self.C, 3)
        _capi.check(_capi.lib().acino_sba_set_rvec_t(self._h, _capi.ptr(rv), _capi.ptr(tv)))

    def get_rvec_t(self):
        rv = np.empty((self.C, 3)); tv = np.empty((self.C, 3))
        _capi.check(_capi.lib().acino_sba_get_rvec_t(self._h, _capi.ptr(rv), _capi.ptr(tv)))
        return rv, tv

    def get_points(self):
        x = np.empty((self.P, 3))
        _capi.check(_capi.lib().acino_sba_get_points(self._h, _capi.ptr(x)))
        return x

    def get_cameras(self):
        c = np.empty((self.C, _capi.CAM_STRIDE))
        _capi.check(_capi.lib().acino_sba_get_cameras(self._h, _capi.ptr(c)))
        return c

    def residuals(self):
        f = np.empty((self.P, self.C, 2))
        _capi.check(_capi.lib().acino_sba_residuals(self._h, _capi.ptr(f)))
        return f

    def reset_state(self):
        _capi.check(_capi.lib().acino_sba_reset_state(self._h))

    def solve(self, opts=None, **kw):
        o = opts or default_opts(**kw)
        st = SbaStats()
        _capi.check(_capi.lib().acino_sba_solve(self._h, ct.byref(o), ct.byref(st)))
        return st.as_dict()

    def iterate(self, opts):
        _capi.check(_capi.lib().acino_sba_iterate(self._h, ct.byref(opts)))

    def profile(self, enable=True):
        _capi.check(_capi.lib().acino_sba_profile(self._h, int(bool(enable))))

    def profile_read(self):
        """Average device ms of the build (K4) and back-substitution (K6) kernels over the profiled iterations."""
        b, k, n = ct.c_double(), ct.c_double(), ct.c_int()
        _capi.check(_capi.lib().acino_sba_profile_read(self._h, ct.byref(b), ct.byref(k), ct.byref(n)))
        ms4 = np.zeros(4)
        _capi.check(_capi.lib().acino_sba_profile_read_kernels(self._h, _capi.ptr(ms4), ct.byref(n)))
        return dict(build_ms=b.value, backsub_ms=k.value, iterations=n.value, linearise_ms=float(ms4[0]),
                    accumulate_ms=float(ms4[1]), solve_ms=float(ms4[2]))

    def history(self):
        """Iteration log of the latest points+extrinsics solve: (n, 6) = iteration, cost, cost reduction, max |step|,
        max |gradient|, accepted; row 0 is the starting point."""
        n = ct.c_int()
        _capi.check(_capi.lib().acino_sba_history(self._h, None, 0, ct.byref(n)))
        out = np.zeros((n.value, 6))
        if n.value:
            _capi.check(_capi.lib().acino_sba_history(self._h, _capi.ptr(out), n.value, ct.byref(n)))
        return out

    def lm_state(self):
        out = np.empty(16)
        _capi.check(_capi.lib().acino_sba_lm_state(self._h, _capi.ptr(out)))
        return dict(cost=out[0], lam=out[1], gmax=out[2], iterations=int(out[3]), accepted=int(out[4]), status=int(out[5]),
                    cost_trial=out[6], pred=out[7], rho=out[8], dxmax=out[9], xmax=out[10], nu=out[11], last_rel_dec=out[12],
                    rejected=int(out[13]), chol_fail=int(out[14]), cost_first=out[15])

    @property
    def stream(self):
        return _capi.lib().acino_sba_stream(self._h)

    def comm_init(self, unique_id: bytes, rank: int, nranks: int):
        buf = ct.create_string_buffer(unique_id, 128)
        _capi.check(_capi.lib().acino_sba_comm_init(self._h, buf, rank, nranks))

    def comm_mode(self) -> str:
        """How the per-iteration all-reduces travel: 'single', 'nccl' or 'peer' (NVLink peer-memory exchange)."""
        m = _capi.lib().acino_sba_comm_mode(self._h)
        if m < 0:
            _capi.check(m)
        return ('single', 'nccl', 'peer')[m]


def nccl_unique_id() -> bytes:
    buf = ct.create_string_buffer(128)
    _capi.check(_capi.lib().acino_nccl_unique_id(buf))
    return buf.raw


class ScenePipeline:
    """core.tri + core.sba on host buffers with the device allocations (and the NCCL communicator) kept across calls:
    upload observations and likelihoods, threshold + ring triangulation (K2), SBA solve, download the points
    (src/core/tri.py:41-47, src/core/sba.py:41-43). With ``comm=(unique_id, rank, world)`` every rank holds a frame shard
    and ``run`` is collective."""

    def __init__(self, n_points, n_cams, k_arr, d_arr, r_arr, t_arr, mode=MODE_POINTS_ONLY, comm=None):
        _capi.require_gpu()
        self.P, self.C, self.mode = int(n_points), int(n_cams), mode
        self.cam_args = (k_arr, d_arr, r_arr, t_arr)
        self.cams = _capi.pack_cams(k_arr, d_arr, r_arr, t_arr)
        self.uv_d = DeviceArray((self.P, self.C, 2), np.float64)
        self.lik_d = DeviceArray((self.P, self.C), np.float64)
        self.xyz_d = DeviceArray((self.P, 3), np.float64)
        self.mask_d = DeviceArray((self.P,), np.uint8)
        self.ex_d = DeviceArray((self.P,), np.uint8)
        self.solver = SbaSolver(self.C, self.P)
        self.solver.set_observations_dev(self.uv_d, self.mask_d)
        if comm is not None and comm[2] > 1:
            self.solver.comm_init(*comm)

    def run(self, uv, lik, thresh, out_points=None, rvec_t0=None, breakdown=False, **solver_opts):
        """``breakdown=True`` synchronises between the phases and adds ``st['breakdown_ms']`` = h2d, tri, init (the
        once-per-solve set-up of the points+extrinsics solver), lm (joint iterations + per-point refinement), polish,
        d2h - host clock; the extra synchronisations make such a run a little slower than a plain one."""
        import time
        L = _capi.lib()
        t0 = time.perf_counter()
        self.uv_d.upload(uv)
        self.lik_d.upload(lik)
        t1 = time.perf_counter()
        _capi.check(L.acino_triangulate_ring_dev(self.uv_d.ptr, self.lik_d.ptr, float(thresh), self.P, self.C,
                                                 _capi.ptr(self.cams), self.xyz_d.ptr, self.mask_d.ptr, self.ex_d.ptr,
                                                 ct.c_void_p(self.solver.stream)))
        if breakdown:
            _capi.check(L.acino_sync())
        t2 = time.perf_counter()
        self.solver.set_cameras(*self.cam_args)
        if self.mode == MODE_POINTS_AND_EXTRINSICS and rvec_t0 is not None and rvec_t0[0] is not None:
            self.solver.set_rvec_t(rvec_t0[0], rvec_t0[1])
        self.solver.set_points_dev(self.xyz_d)
        st = self.solver.solve(default_opts(self.mode, **solver_opts))
        t3 = time.perf_counter()
        if out_points is None:
            out_points = np.empty((self.P, 3))
        _capi.check(L.acino_sba_get_points(self.solver._h, _capi.ptr(out_points)))
        t4 = time.perf_counter()
        st['points'] = out_points
        if breakdown:
            ms = lambda a, b: 1e3 * (b - a)
            st['breakdown_ms'] = dict(h2d=ms(t0, t1), tri=ms(t1, t2), init=st.get('init_ms', 0.0), polish=st.get('polish_ms', 0.0),
                                      lm=ms(t2, t3) - st.get('init_ms', 0.0) - st.get('polish_ms', 0.0), d2h=ms(t3, t4),
                                      total=ms(t0, t4))
        return st

    def run_many(self, scenes, thresh, out_points=None, rvec_t0=None, **solver_opts):
        """The same as ``run`` for a sequence of scenes ``[(uv, lik), ...]`` of this pipeline's shape, with the transfers
        taken off the critical path: the upload of scene i + 1 (one of two input buffers, a copy stream of the
        pipeline's: ``acino_h2d_async``) and the download of the points of scene i - 1 (a device copy of them,
        ``acino_sba_get_points_dev``, on a second copy stream) run beside the triangulation and the solve of scene i.
        The host arrays must be page-locked for that (e.g. ``torch.Tensor.pin_memory``); pageable ones work, without
        the overlap. Returns the list of solver statistics, all results on the host; ``st['t_solved']`` is the host
        clock (``time.perf_counter``) when the scene's solve had ended and its download was queued."""
        import time
        L = _capi.lib()
        scenes = list(scenes)
        n = len(scenes)
        if n == 0:
            return []
        if getattr(self, '_copy_stream', None) is None:
            sp, sq = ct.c_void_p(), ct.c_void_p()
            _capi.check(L.acino_stream_create(ct.byref(sp)))
            _capi.check(L.acino_stream_create(ct.byref(sq)))
            self._copy_stream, self._down_stream = sp, sq
            self._uv2 = [self.uv_d, DeviceArray((self.P, self.C, 2), np.float64)]
            self._lik2 = [self.lik_d, DeviceArray((self.P, self.C), np.float64)]
            self._out2 = [DeviceArray((self.P, 3), np.float64), DeviceArray((self.P, 3), np.float64)]
        cs, ds = self._copy_stream, self._down_stream
        keep = []

        def upload(i):
            uv = _capi.f64(scenes[i][0]).reshape(self.P, self.C, 2)
            lik = _capi.f64(scenes[i][1]).reshape(self.P, self.C)
            keep.append((uv, lik))          # alive until the copy has run
            _capi.check(L.acino_h2d_async(self._uv2[i & 1].ptr, _capi.ptr(uv), uv.nbytes, cs))
            _capi.check(L.acino_h2d_async(self._lik2[i & 1].ptr, _capi.ptr(lik), lik.nbytes, cs))

        upload(0)
        res = []
        for i in range(n):
            _capi.check(L.acino_stream_sync(cs))        # scene i is on the device (its copy ran during solve i - 1)
            if i + 1 < n:
                upload(i + 1)                           # runs beside everything below
            uv_d, lik_d = self._uv2[i & 1], self._lik2[i & 1]
            _capi.check(L.acino_triangulate_ring_dev(uv_d.ptr, lik_d.ptr, float(thresh), self.P, self.C, _capi.ptr(self.cams),
                                                     self.xyz_d.ptr, self.mask_d.ptr, self.ex_d.ptr, ct.c_void_p(self.solver.stream)))
            self.solver.set_observations_dev(uv_d, self.mask_d)
            self.solver.set_cameras(*self.cam_args)
            if self.mode == MODE_POINTS_AND_EXTRINSICS and rvec_t0 is not None and rvec_t0[0] is not None:
                self.solver.set_rvec_t(rvec_t0[0], rvec_t0[1])
            self.solver.set_points_dev(self.xyz_d)
            st = self.solver.solve(default_opts(self.mode, **solver_opts))
            out = out_points[i] if out_points is not None else np.empty((self.P, 3))
            stage = self._out2[i & 1]
            if i >= 2:
                _capi.check(L.acino_stream_sync(ds))    # the download of scene i - 2 has left this staging buffer
            _capi.check(L.acino_sba_get_points_dev(self.solver._h, stage.ptr))
            _capi.check(L.acino_d2h_async(_capi.ptr(out), stage.ptr, stage.nbytes, ds))     # runs beside scene i + 1
            st['points'] = out
            st['t_solved'] = time.perf_counter()
            res.append(st)
        _capi.check(L.acino_stream_sync(ds))
        self.solver.set_observations_dev(self.uv_d, self.mask_d)      # back to the buffer ``run`` uploads into
        return res

    def close(self):
        self.solver.close()
        if getattr(self, '_copy_stream', None) is not None:
            _capi.lib().acino_stream_destroy(self._copy_stream)
            _capi.lib().acino_stream_destroy(self._down_stream)
            self._uv2[1].free(); self._lik2[1].free(); self._out2[0].free(); self._out2[1].free()
            self._copy_stream = None
        for a in (self.uv_d, self.lik_d, self.xyz_d, self.mask_d, self.ex_d):
            a.free()


def solve_scene_host(uv, lik, thresh, k_arr, d_arr, r_arr, t_arr, mode=MODE_POINTS_ONLY, out_points=None, rvec_t0=None,
                     **solver_opts):
    """One-shot ScenePipeline: returns the solver statistics; absent points are NaN rows of ``st['points']``."""
    uv = _capi.f64(uv)
    pipe = ScenePipeline(uv.shape[0], uv.shape[1], k_arr, d_arr, r_arr, t_arr, mode)
    try:
        return pipe.run(uv, lik, thresh, out_points, rvec_t0, **solver_opts)
    finally:
        pipe.close()


def fp64_peak_tflops():
    """Measured fp64 FMA throughput of the current device (acino_fp64_peak)."""
    _capi.require_gpu()
    v = ct.c_double()
    _capi.check(_capi.lib().acino_fp64_peak(ct.byref(v)))
    return v.value


def frame_shard(n_frames, rank, world):
    """Contiguous, balanced frame range [start, stop) of ``rank`` (observations shard by frame)."""
    if not (0 <= rank < world):
        raise ValueError('rank outside the world')
    base, extra = divmod(int(n_frames), int(world))
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)
