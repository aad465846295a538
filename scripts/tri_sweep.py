"""Config C3 (BASELINE.json configs[2]): ring-pair DLT triangulation sweep, 10^3 .. 10^7 frames x 20 keypoints x 6 cameras
on one GPU (frames shard with no communication, so N GPUs are N independent replicas of this). Inputs resident in HBM;
scenes above 10^5 frames replicate a 10^5-frame synthetic block on the device (the kernel's work per point does not depend
on which frame it is). Prints one JSON line per size. Needs a GPU."""
import sys, os, json, ctypes as ct
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from acinoset_b200 import synth, dense, _capi

rank, world, local = (int(os.environ.get(k_, d_)) for k_, d_ in (('RANK', '0'), ('WORLD_SIZE', '1'), ('LOCAL_RANK', '0')))
torch.cuda.set_device(local)
L = _capi.lib()
_capi.check(L.acino_set_device(local))
if world > 1:
    import torch.distributed as dist
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
K, C = 20, 6
sizes = [int(float(a)) for a in sys.argv[1:]] or [1000, 10000, 100000, 1000000, 10000000]
block_frames = 100000
sc = synth.make_rig_scene(block_frames, K, seed=4)
cams = _capi.pack_cams(sc['k'], sc['d'], sc['r'], sc['t'])
uv_b = torch.from_numpy(sc['uv']).cuda(); lik_b = torch.from_numpy(sc['lik']).cuda()
peak = 6538.9
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))['hbm_gbs']
except Exception:
    pass
for F_total in sizes:
    F = F_total // world             # frames shard across the ranks, no communication
    if F == 0:
        continue
    P = F * K
    reps = -(-F // block_frames)
    uv = uv_b.repeat(reps, 1, 1)[:P].contiguous() if reps > 1 else uv_b[:P].contiguous()
    lik = lik_b.repeat(reps, 1)[:P].contiguous() if reps > 1 else lik_b[:P].contiguous()
    xyz = torch.empty((P, 3), dtype=torch.float64, device='cuda')
    mask = torch.empty(P, dtype=torch.uint8, device='cuda'); ex = torch.empty(P, dtype=torch.uint8, device='cuda')
    s = torch.cuda.current_stream()
    def run():
        _capi.check(L.acino_triangulate_ring_dev(ct.c_void_p(uv.data_ptr()), ct.c_void_p(lik.data_ptr()), 0.5, P, C, _capi.ptr(cams),
                                                 ct.c_void_p(xyz.data_ptr()), ct.c_void_p(mask.data_ptr()), ct.c_void_p(ex.data_ptr()),
                                                 ct.c_void_p(s.cuda_stream)))
    for _ in range(3):
        run()
    n = max(3, min(50, int(2e8 / P)))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    if world > 1:
        tt = torch.tensor([ms], device='cuda', dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    P = P * world                    # whole-job figures: all ranks' points over the slowest rank's time
    byts = P * 169
    if rank == 0:
      print(json.dumps(dict(config='C3 TRI sweep', n_gpus=world, frames=F * world, points=P, existing=int(ex.sum().item()), ms=ms, points_per_s=P / ms * 1e3,
                          algorithmic_gbs=byts / ms / 1e6, frac_hbm=byts / ms / 1e6 / (peak * world), launches_timed=n,
                          note='acino_triangulate_ring_dev incl. its stream synchronisation; inputs resident in HBM; max over ranks')), flush=True)
    del uv, lik, xyz, mask, ex
    torch.cuda.empty_cache()
if world > 1:
    dist.destroy_process_group()
