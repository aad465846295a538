#!/usr/bin/env python
"""bench.py — SBA LM iterations/s (and solve ms, TRI points/s) on the synthetic 6-camera fisheye scene.

Contract: `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0. For N>1 it is
launched by torch.distributed.run (one rank per GPU); the frames shard across ranks (weak scaling: every rank
holds one C4-sized shard of 119,490 frames x 20 keypoints x 6 cameras).

A "step" is one Levenberg-Marquardt iteration over the rank's whole shard (residuals + analytic Jacobian +
normal equations + solve + trial evaluation + accept/reject), launched through the C ABI
(acino_sba_iterate) with inputs resident in HBM. `value` = N * K / t, i.e. LM iterations per second in units of
the 119,490-frame scene. `e2e` = the same quantity through the host-buffer API (pinned host observations ->
HBM -> TRI -> full solve -> points back on the host). `--impl reference` times the reference's CPU algorithm
(oracle/ref_port.py: per-observation cv2.fisheye.projectPoints loop inside scipy least_squares, exactly the
call at src/lib/sba.py:189) on a bounded sample and scales it linearly to the same workload.
"""
import argparse
import ctypes as ct
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

C4_FRAMES = 119490
N_KPTS = 20
N_CAMS = 6
THRESH = 0.5
METRIC = 'SBA LM iters/sec, 6-cam fisheye, 20 kpts x 119490 frames'
UNIT = 'LM iterations/s (119,490-frame scene equivalents)'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--impl', default='native', choices=['native', 'reference'])
    ap.add_argument('--mode', default='auto', choices=['auto', 'points', 'extrinsics'])
    ap.add_argument('--frames', type=int, default=C4_FRAMES, help='frames per GPU shard')
    ap.add_argument('--cpu-frames', type=int, default=60, help='frames of the CPU-baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--masked-frac', type=float, default=0.5, help='fraction of slots at or below dlc_thresh')
    ap.add_argument('--outlier-frac', type=float, default=0.0, help='fraction of observations with N(0,50 px) outlier noise (C5)')
    ap.add_argument('--e2e-solves', type=int, default=3)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index=0):
        self.f = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits', '-lms', '50',
                                       '-i', str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for line in self.f.read().splitlines():
            parts = [s.strip() for s in line.split(',')]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(nm)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


def dist_env():
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    return rank, world, local


MASKED_FRAC, OUTLIER_FRAC = 0.5, 0.0


def make_shard(frames, seed):
    from acinoset_b200 import synth
    sc = synth.make_rig_scene(frames, N_KPTS, seed=seed, noise_px=1.0, masked_frac=MASKED_FRAC, outlier_frac=OUTLIER_FRAC)
    return sc


# algorithmic bytes per point and LM iteration (DESIGN.md "Algorithmic bytes"): the kernel reads the point's
# observation mask (1 B), its kept observations (16 B each; dense slots, 50 % kept -> 3 of 6 on average, counted
# exactly from the mask), the point (24 B), damping (8 B), status (1 B); writes point (24 B), damping (8 B),
# cost (8 B), status (1 B).
def lm_points_bytes(n_points_active, n_obs):
    return n_points_active * (1 + 24 + 8 + 1 + 24 + 8 + 8 + 1) + n_obs * 16


def tri_bytes(P, C):
    return P * (C * 16 + C * 8 + 24 + 1 + 1)


def native_arm(args):
    import torch
    import torch.distributed as dist
    from acinoset_b200 import _capi, dense
    rank, world, local = dist_env()
    if args.gpus != world and world > 1:
        raise SystemExit(f'--gpus {args.gpus} but WORLD_SIZE={world}')
    _capi.require_gpu()
    _capi.check(_capi.lib().acino_set_device(local))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    mode = args.mode
    if mode == 'auto':
        mode = 'extrinsics' if HAVE_SCHUR() else 'points'
    sc = make_shard(args.frames, seed=4 + rank)
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    P, C = sc['lik'].shape
    cams = _capi.pack_cams(k, d, r, t)
    L = _capi.lib()

    # ---- resident inputs
    uv_d = dense.DeviceArray.from_host(sc['uv'])
    lik_d = dense.DeviceArray.from_host(sc['lik'])
    xyz_d = dense.DeviceArray((P, 3), np.float64)
    mask_d = dense.DeviceArray((P,), np.uint8)
    ex_d = dense.DeviceArray((P,), np.uint8)
    solver = dense.SbaSolver(C, P)
    stream = torch.cuda.ExternalStream(solver.stream)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def tri():
        _capi.check(L.acino_triangulate_ring_dev(uv_d.ptr, lik_d.ptr, THRESH, P, C, _capi.ptr(cams), xyz_d.ptr, mask_d.ptr,
                                                 ex_d.ptr, ct.c_void_p(solver.stream)))
    # TRI (K2): timed alone, 3 warm-ups + 5 timed launches
    for _ in range(3):
        tri()
    tri_ms = []
    for _ in range(5):
        e0, e1 = ev(), ev()
        e0.record(stream); tri(); e1.record(stream); e1.synchronize()
        tri_ms.append(e0.elapsed_time(e1))
    tri_ms = float(np.median(tri_ms))
    mask = mask_d.download()
    n_obs = int(np.unpackbits(mask).sum())
    n_pts = int((mask != 0).sum())

    solver.set_cameras(k, d, r, t)
    solver.set_observations_dev(uv_d, mask_d)
    xyz0 = np.nan_to_num(xyz_d.download())
    if mode == 'extrinsics':
        rng = np.random.default_rng(100)          # same perturbed extrinsics on every rank
        from acinoset_b200 import sba as sba_mod
        rv = np.array([sba_mod.rotation_matrix_to_rvec(m) for m in r]) + rng.normal(0, 8e-4, (C, 3))
        tv = t.reshape(C, 3) + rng.normal(0, 3e-3, (C, 3))
        solver.set_rvec_t(rv, tv)
    solver.set_points(xyz0)
    if world > 1:
        uid = [dense.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        solver.comm_init(uid[0], rank, world)
    comm_mode = solver.comm_mode()

    m = _capi.MODE_POINTS_AND_EXTRINSICS if mode == 'extrinsics' else _capi.MODE_POINTS_ONLY
    # full solve from the TRI initialisation (solve ms, iterations to converge); at N > 1 the ranks enter together and the
    # first collective (NCCL connection set-up) is spent before the timed solve
    if world > 1:
        solver.iterate(dense.default_opts(m, force_iters=1))
        solver.reset_state()
        torch.cuda.synchronize()
        dist.barrier()
    st_solve = solver.solve(dense.default_opts(m, **(dict(ftol=1e-10) if mode == 'extrinsics' else {})))
    # fixed-work LM iterations for the throughput number: every point does a full iteration each step
    solver.reset_state()
    # the timed step is the pure LM iteration (K4-K7); the per-point refinement between camera steps that full solves use
    # (point_refine_iters) is left out of the fixed-work step and reported through `solve` / `e2e`
    opts = dense.default_opts(m, inner_iters=1, force_iters=1, **(dict(point_refine_iters=0) if mode == 'extrinsics' else {}))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local) if rank == 0 else None
    for _ in range(max(args.warmup, 3)):
        solver.iterate(opts)
    if mode == 'extrinsics':
        solver.profile(True)          # CUDA events around K4 / K6 on the solver's stream, inside the timed region
    barrier()
    e0, e1 = ev(), ev()
    e0.record(stream)
    for _ in range(args.steps):
        solver.iterate(opts)
    e1.record(stream)
    e1.synchronize()
    barrier()
    ms = e0.elapsed_time(e1)
    kern = None
    if mode == 'extrinsics':
        pr = solver.profile_read()
        solver.profile(False)
        # K4 (build) reads the mask of every slot-group, and for existing points the point and its kept observations
        build_bytes = P * 1 + n_pts * 24 + n_obs * 16
        back_bytes = build_bytes + n_pts * 24
        kern = dict(name=f'schur_build_kernel<{C}>', ms_per_launch=pr['build_ms'], bytes_per_launch=build_bytes,
                    launches_per_step=5 if world == 1 else (7 if comm_mode == 'peer' else 8), collective=comm_mode, backsub_ms=pr['backsub_ms'], backsub_bytes=back_bytes,
                    profiled_iterations=pr['iterations'])
    fp64_peak = dense.fp64_peak_tflops()
    if world > 1:
        tms = torch.tensor([ms], device='cuda', dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    clk = clocks.stop() if clocks else None
    ms_per_step = ms / args.steps
    value = world * args.steps / (ms * 1e-3) * (args.frames / C4_FRAMES)
    return dict(locals())


def HAVE_SCHUR():
    """True when the library's points+extrinsics solver is built (acino_version() >= 200)."""
    from acinoset_b200 import _capi
    return _capi.lib().acino_version() >= 200


def e2e_arm(args, ctx):
    """Full solves through the host-buffer API (dense.ScenePipeline): pinned host observations -> HBM -> TRI -> solve ->
    host points. Every solve uploads its inputs and downloads its result inside the timed region; at N > 1 every rank
    feeds its own frame shard and the solve is collective (NCCL allreduce of the reduced camera system)."""
    import torch
    import torch.distributed as dist
    from acinoset_b200 import dense
    sc, P, C, world, rank = ctx['sc'], ctx['P'], ctx['C'], ctx['world'], ctx['rank']
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    uv_h = torch.from_numpy(sc['uv']).pin_memory()
    lik_h = torch.from_numpy(sc['lik']).pin_memory()
    out_h = torch.empty((P, 3), dtype=torch.float64).pin_memory()
    comm = None
    if world > 1 and ctx['mode'] == 'extrinsics':
        uid = [dense.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        comm = (uid[0], rank, world)
    pipe = dense.ScenePipeline(P, C, k, d, r, t, ctx['m'], comm)
    rv0 = (ctx.get('rv'), ctx.get('tv')) if ctx['mode'] == 'extrinsics' else None
    # the reference's own stopping literal for this mode (src/lib/sba.py:170-173 ftol=1e-10); points-only keeps its defaults
    opts = dict(ftol=1e-10) if ctx['mode'] == 'extrinsics' else {}
    iters, times, last = 0, [], None
    for i in range(args.e2e_solves + 1):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        last = pipe.run(uv_h.numpy(), lik_h.numpy(), THRESH, out_h.numpy(), rv0, **opts)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if i > 0:                     # first solve is the warm-up
            times.append(dt)
            iters += last['iterations']
    pipe.close()
    total = float(np.sum(times))
    if world > 1:
        tt = torch.tensor([total], device='cuda', dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total = float(tt.item())
    h2d = uv_h.numel() * 8 + lik_h.numel() * 8
    d2h = out_h.numel() * 8
    return dict(value=world * iters / total * (args.frames / C4_FRAMES), unit=UNIT,
                step='one full solve per GPU shard: H2D observations+likelihoods, TRI, LM until the stopping rule, D2H points',
                solve_ms=1e3 * total / len(times), lm_iters_per_solve=iters / len(times), exit_status=last.get('status'),
                cost_initial=last['cost_initial'], cost_final=last['cost_final'],
                h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h))


def cpu_baseline(frames, mode, full_obs):
    """The reference's CPU algorithm (oracle/ref_port.py) on a bounded sample; scaled linearly in observations to the
    bench workload (the per-observation Python/OpenCV loop is exactly linear, SURVEY.md §6, and the number of finite-
    difference column groups per Jacobian does not depend on the scene size)."""
    from oracle import ref_port
    t0 = time.perf_counter()
    if mode == 'extrinsics':
        frames = max(2, min(frames, 40))      # ~10 s of host work at 25 function evaluations
        sc = make_shard(frames, seed=4)
        rng = np.random.default_rng(100)
        from acinoset_b200 import synth
        r0 = np.array([sc['r'][c] @ synth._rot_from_axis_angle(rng.normal(0, 8e-4, 3)) for c in range(N_CAMS)])
        t0_ = sc['t'] + rng.normal(0, 3e-3, sc['t'].shape)
        out = ref_port.sba_extrinsics_scene(sc['uv'], sc['lik'], THRESH, sc['k'], sc['d'], r0, t0_)
        what = ('points+extrinsics SBA (reference algorithm with the layout-consistent sparsity pattern, SURVEY.md F6; '
                'bounded at 25 nfev)')
    else:
        sc = make_shard(frames, seed=4)
        out = ref_port.sba_points_scene(sc['uv'], sc['lik'], THRESH, sc['k'], sc['d'], sc['r'], sc['t'], f_scale=50.0)
        what = 'points-only SBA'
    dt = time.perf_counter() - t0
    its_per_s_sample = max(out['iterations'], 1) / out['sba_seconds']
    scale = out['n_obs'] / float(full_obs)
    return dict(value=its_per_s_sample * scale, unit=UNIT, cores=1, kind='port',
                sample=(f'{frames} frames x 20 kpts x 6 cams ({out["n_obs"]} observations, {out["n_points"]} points): '
                        f'{what} {out["sba_seconds"]:.2f} s, {out["iterations"]} TRF iterations, '
                        f'{out["nfev"]} nfev, {out["cost_calls"]} cost-function sweeps; iterations/s scaled by observations '
                        f'({out["n_obs"]}/{int(full_obs)}) to the bench workload'),
                host_cpus=os.cpu_count(), sample_seconds=dt, sample_iters_per_s=its_per_s_sample,
                us_per_obs_per_sweep=1e6 * out['sba_seconds'] / (out['cost_calls'] * out['n_obs']))


def main():
    global MASKED_FRAC, OUTLIER_FRAC
    args = parse_args()
    MASKED_FRAC, OUTLIER_FRAC = args.masked_frac, args.outlier_frac
    rank, world, local = dist_env()
    if args.impl == 'reference':
        if rank != 0:
            return
        full_obs = args.frames * N_KPTS * N_CAMS * 0.42     # kept observations of existing points at thresh 0.5 (measured ratio)
        vals = []
        cb = None
        mode = 'points' if args.mode == 'points' else 'extrinsics'
        for i in range(max(1, min(args.steps, 2))):
            cb = cpu_baseline(args.cpu_frames, mode, full_obs)
            vals.append(cb['value'])
        v = float(np.median(vals))
        cb['value'] = v
        line = dict(impl='reference', metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=len(vals), warmup=0,
                    ms_per_step=1e3 / v, higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f64',
                    data='synthetic', config=dict(workload=f'{"C4" if args.frames == C4_FRAMES else "custom"} {"points+extrinsics (Schur)" if mode == "extrinsics" else "points-only"} SBA: '
                                                           f'{args.frames} frames x 20 keypoints x 6 cameras per GPU, dlc_thresh 0.5, 1 px noise',
                                                  masked_frac=args.masked_frac, outlier_frac=args.outlier_frac, mode=mode,
                                                  parallelism=f'frame-shard x{args.gpus}',
                                                  note='reference CPU algorithm on rank 0, single-threaded as in the reference (its per-observation '
                                                       'Python/OpenCV loop has no parallel section); bounded sample scaled to the workload'),
                    cpu_baseline=cb, e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
        print(json.dumps(line))
        return

    ctx = native_arm(args)
    from acinoset_b200 import _capi
    peak, peak_src = peaks()
    mode, P, C = ctx['mode'], ctx['P'], ctx['C']
    e2e = e2e_arm(args, ctx)
    if rank != 0:
        import torch.distributed as dist
        dist.destroy_process_group()
        return
    n_obs, n_pts = ctx['n_obs'], ctx['n_pts']
    if mode == 'points':
        bytes_step = lm_points_bytes(n_pts, n_obs) + (P - n_pts) * 2
        kern_ms = ctx['ms_per_step']
        kern_name = 'lm_points_kernel<6>'
        launches = args.steps
    else:
        bytes_step = ctx['kern']['bytes_per_launch']
        kern_ms = ctx['kern']['ms_per_launch']
        kern_name = ctx['kern']['name']
        launches = ctx['kern']['launches_per_step'] * args.steps
    achieved = bytes_step / (kern_ms * 1e-3) / 1e9
    roofline = dict(bound='hbm', achieved=achieved, peak=peak, unit='GB/s', frac=achieved / peak, traffic=None,
                    kernel=kern_name, algorithmic_bytes_per_launch=int(bytes_step), kernel_ms=kern_ms, peak_source=peak_src,
                    fp64_peak_tflops_measured=ctx['fp64_peak'],
                    note='fp64-pipe bound, not HBM bound: see profiles/ for the ncu pipe utilisation')
    if mode == 'extrinsics':
        k6 = ctx['kern']
        roofline['backsub_kernel'] = dict(kernel=f'schur_backsub_kernel<{C}>', kernel_ms=k6['backsub_ms'],
                                          algorithmic_bytes_per_launch=int(k6['backsub_bytes']),
                                          achieved=k6['backsub_bytes'] / (k6['backsub_ms'] * 1e-3) / 1e9,
                                          frac=k6['backsub_bytes'] / (k6['backsub_ms'] * 1e-3) / 1e9 / peak)
        roofline['step_share_of_kernels'] = (k6['ms_per_launch'] + k6['backsub_ms']) / ctx['ms_per_step']
    # DRAM traffic and executed fp64 flops per launch come from the committed ncu capture of the same workload
    # (profiles/r01_ncu_counts.json); they are attached only when this run IS that workload
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'profiles', 'r01_ncu_counts.json')) as f:
            ncu = json.load(f)
        w = ncu['workload']
        if (args.frames, args.masked_frac, args.outlier_frac) == (w['frames'], w['masked_frac'], w['outlier_frac']) and mode == 'extrinsics':
            def fp64_block(name, ms):
                k = ncu['kernels'][name]
                tf = k['fp64_flops'] / (ms * 1e-3) / 1e12
                return dict(flops_per_launch=k['fp64_flops'], achieved_tflops=tf, peak_tflops=ctx['fp64_peak'], frac=tf / ctx['fp64_peak'],
                            pipe_active_pct_ncu=k['fp64_pipe_pct'], source='executed DFMA x2 + DMUL + DADD, ncu SASS page (profiles/r01_ncu_counts.json)')
            roofline['traffic'] = ncu['kernels'][kern_name]['dram_bytes']
            roofline['traffic_source'] = 'ncu dram__bytes_read.sum + dram__bytes_write.sum per launch (profiles/r01_ncu_counts.json)'
            roofline['fp64'] = fp64_block(kern_name, kern_ms)
            roofline['traffic_gbs'] = roofline['traffic'] / (kern_ms * 1e-3) / 1e9
            roofline['traffic_frac_of_hbm_peak'] = roofline['traffic_gbs'] / peak
            bk = roofline['backsub_kernel']
            bk['traffic'] = ncu['kernels'][f'schur_backsub_kernel<{C}>']['dram_bytes']
            bk['traffic_gbs'] = bk['traffic'] / (ctx['kern']['backsub_ms'] * 1e-3) / 1e9
            bk['traffic_frac_of_hbm_peak'] = bk['traffic_gbs'] / peak       # K6 is the HBM-bound kernel of the path
            roofline['backsub_kernel']['fp64'] = fp64_block(f'schur_backsub_kernel<{C}>', ctx['kern']['backsub_ms'])
    except (OSError, KeyError, ValueError):
        pass
    tri_gbs = tri_bytes(P, C) / (ctx['tri_ms'] * 1e-3) / 1e9
    cb = None
    if not args.no_cpu_baseline and world == 1:
        cb = cpu_baseline(args.cpu_frames, mode, n_obs)
    st = ctx['st_solve']
    line = dict(metric=METRIC, value=ctx['value'], unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                ms_per_step=ctx['ms_per_step'], higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f64',
                data='synthetic',
                config=dict(workload=f'{"C4" if args.frames == C4_FRAMES else "custom"} {"points+extrinsics (Schur)" if mode == "extrinsics" else "points-only"} SBA: '
                                     f'{args.frames} frames x 20 keypoints x 6 cameras per GPU, dlc_thresh 0.5, 1 px noise',
                            masked_frac=args.masked_frac, outlier_frac=args.outlier_frac, mode=mode, points_per_gpu=n_pts, observations_per_gpu=n_obs,
                            l2='inputs larger than L2 (229 MB of observations per GPU vs 126 MB L2)', parallelism=f'frame-shard x{world}',
                            collective={'single': 'none (one GPU)', 'nccl': 'NCCL all-reduce, 2 grouped launches per iteration',
                                        'peer': 'NVLink peer-memory exchange fused into K5 / K7 (csrc/acino_sba_xchg.cuh)'}[(ctx['kern'] or {}).get('collective', 'single')]),
                clocks=ctx['clk'], e2e=e2e, gpu_launches=launches, roofline=roofline, cpu_baseline=cb,
                solve=dict(ms=st['solve_ms'], iterations=st['iterations'], launches=st['launches'], cost_initial=st['cost_initial'],
                           cost_final=st['cost_final'], n_unconverged=st['n_unconverged']),
                tri=dict(ms=ctx['tri_ms'], points_per_s=P / (ctx['tri_ms'] * 1e-3), gbs=tri_gbs, frac_hbm=tri_gbs / peak,
                         algorithmic_bytes=tri_bytes(P, C)))
    print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
