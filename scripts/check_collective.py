"""Sharded vs single-GPU points+extrinsics solve (run under torchrun with >= 2 GPUs):
every rank solves its frame shard collectively (NCCL allreduce of the reduced camera system per iteration); rank 0 then
solves the whole scene alone. The two runs minimise the same sum, so cameras, cost and points must agree to rounding.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/check_collective.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from acinoset_b200 import synth, dense, sba, _capi

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
_capi.check(_capi.lib().acino_set_device(local))
dist.init_process_group('nccl')
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
K = 20
sc = synth.make_rig_scene(frames, K, seed=4)
k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
C = len(k)
rng = np.random.default_rng(100)
rv = np.array([sba.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
ids = [dense.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(ids, src=0)
f0, f1 = dense.frame_shard(frames, rank, world)
uv, lik = sc['uv'][f0 * K:f1 * K], sc['lik'][f0 * K:f1 * K]
pipe = dense.ScenePipeline(len(uv), C, k, d, r, t, mode=_capi.MODE_POINTS_AND_EXTRINSICS, comm=(ids[0], rank, world))
opts = dict(max_iter=15, ftol=0.0, gtol=0.0, xtol=0.0)     # a fixed number of iterations: the same path on both sides
st = pipe.run(uv, lik, 0.5, rvec_t0=(rv, tv), **opts)
cams_sharded = pipe.solver.get_cameras()
pts_sharded = st['points']
allc = [None] * world
dist.all_gather_object(allc, cams_sharded)
ok = True
if rank == 0:
    for c in allc[1:]:
        ok &= bool(np.array_equal(c, allc[0]))             # every rank took identical decisions
    print('cameras identical on all ranks:', ok)
    print('collective:', pipe.solver.comm_mode())
    full = dense.solve_scene_host(sc['uv'], sc['lik'], 0.5, k, d, r, t, mode=_capi.MODE_POINTS_AND_EXTRINSICS, rvec_t0=(rv, tv), **opts)
    s1 = dense.SbaSolver  # noqa
    print(f"sharded: cost {st['cost_final']:.12e} it {st['iterations']} | single: cost {full['cost_final']:.12e} it {full['iterations']}")
    rel = abs(st['cost_final'] - full['cost_final']) / full['cost_final']
    dp = np.nanmax(np.abs(pts_sharded - full['points'][f0 * K:f1 * K]))
    print(f'relative cost difference {rel:.3e}, max point difference on rank 0 shard {dp:.3e} m')
    ok &= rel < 1e-9 and dp < 1e-7
    print('COLLECTIVE_CHECK', 'OK' if ok else 'FAILED')
pipe.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
