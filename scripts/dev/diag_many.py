import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch, ctypes as ct
from acinoset_b200 import synth, dense, sba, _capi
sc = synth.make_rig_scene(119490, 20, seed=4, masked_frac=0.5)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
P, C = sc['lik'].shape
uv_h = torch.from_numpy(sc['uv']).pin_memory(); lik_h = torch.from_numpy(sc['lik']).pin_memory()
out_h = torch.empty((P, 3), dtype=torch.float64).pin_memory()
E = _capi.MODE_POINTS_AND_EXTRINSICS
pipe = dense.ScenePipeline(P, C, k, d, r, t, E)
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
L = _capi.lib()
for rep in range(2):
    res = pipe.run_many([(uv_h.numpy(), lik_h.numpy())] * 5, 0.5, [out_h.numpy()] * 5, (rv, tv), ftol=1e-10)
    print([round(1e3 * (res[i]['t_solved'] - res[i - 1]['t_solved']), 2) for i in range(1, 5)], [round(x['solve_ms'], 2) for x in res], 'init', [round(x['init_ms'], 2) for x in res], 'polish', [round(x['polish_ms'], 2) for x in res])
# phases by hand
sp = pipe._copy_stream
def up(slot):
    _capi.check(L.acino_h2d_async(pipe._uv2[slot].ptr, uv_h.data_ptr(), uv_h.numel() * 8, sp))
    _capi.check(L.acino_h2d_async(pipe._lik2[slot].ptr, lik_h.data_ptr(), lik_h.numel() * 8, sp))
# copy alone
_capi.check(L.acino_sync()); t0 = time.perf_counter(); up(1); _capi.check(L.acino_stream_sync(sp)); print('copy alone ms', 1e3 * (time.perf_counter() - t0))
# copy beside a solve: how long until the copy stream is idle, and the solve
for rep in range(2):
    _capi.check(L.acino_sync())
    t0 = time.perf_counter(); up(1)
    st = pipe.run(uv_h.numpy(), lik_h.numpy(), 0.5, out_h.numpy(), (rv, tv), ftol=1e-10)
    t1 = time.perf_counter(); _capi.check(L.acino_stream_sync(sp)); t2 = time.perf_counter()
    print('run (with its own sync upload) beside an async copy: run ms', 1e3 * (t1 - t0), 'copy idle after', 1e3 * (t2 - t0), 'solve_ms', st['solve_ms'])
