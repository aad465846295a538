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
    ap.add_argument('--e2e-solves', type=int, default=5)
    ap.add_argument('--min-time', type=float, default=0.5, help='the timed region repeats its K steps until it lasts this many seconds')
    ap.add_argument('--strong-frames', type=int, default=C4_FRAMES,
                    help='frames of the ONE scene that is sharded over all ranks for the strong-scaling block (0: skip)')
    ap.add_argument('--no-c1', action='store_true', help='skip the run-to-completion C1 comparison')
    return ap.parse_args()


def kernel_source_sha():
    """Hash of the CUDA sources: profiles/r02_ncu_counts.json carries the hash of the sources its ncu capture was taken
    with, and its per-launch DRAM bytes / executed flops are attached to the bench line only when they match."""
    import hashlib
    h = hashlib.sha1()
    d = os.path.join(ROOT, 'acinoset_b200', 'csrc')
    for name in sorted(os.listdir(d)):
        with open(os.path.join(d, name), 'rb') as f:
            h.update(name.encode()); h.update(f.read())
    return h.hexdigest()[:16]


def c1_scene():
    """Config C1: 102 frames x 20 keypoints x 6 cameras, seed 0, dlc_thresh 0.5 - the scene of
    tests/golden/golden_c1.npz, small enough for the reference's CPU algorithm to run to completion."""
    from acinoset_b200 import synth
    return synth.make_rig_scene(102, N_KPTS, seed=0)


def c1_native():
    """C1 to completion through the reference-signature path (app.sba_points_fisheye: DataFrame in, DataFrame out)."""
    import contextlib, io
    from acinoset_b200 import app, synth
    sc = c1_scene()
    df = synth.dense_to_df(sc['uv'], sc['lik'], N_KPTS)
    df = df[df['likelihood'] > THRESH]
    scene = os.path.join(ROOT, 'tests', 'golden', 'scene_6cam.json')
    out = None
    secs = []
    for _ in range(3):                     # first call pays library load / context creation
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            out_df, res = app.sba_points_fisheye(scene, df)
        secs.append(time.perf_counter() - t0)
    from acinoset_b200 import sba as sba_mod
    st = sba_mod.bundle_adjust_points_only.last_stats if hasattr(sba_mod.bundle_adjust_points_only, 'last_stats') else {}
    return dict(seconds=float(min(secs)), seconds_first_call=float(secs[0]), iterations=int(st.get('iterations', -1)),
                rmse=float(np.sqrt(np.mean(res['after'] ** 2))), rmse_before=float(np.sqrt(np.mean(res['before'] ** 2))),
                n_points=int(len(out_df)), n_obs=int(len(res['after']) // 2),
                path='app.sba_points_fisheye (DataFrame in -> TRI -> points-only SBA -> DataFrame out), host buffers')


def c1_reference():
    """C1 to completion with the reference's CPU algorithm (oracle/ref_port.py: pandas ring triangulation + per-observation
    OpenCV projection inside scipy least_squares, src/lib/sba.py:181-195), one core as in the reference."""
    from oracle import ref_port
    sc = c1_scene()
    t0 = time.perf_counter()
    out = ref_port.sba_points_scene(sc['uv'], sc['lik'], THRESH, sc['k'], sc['d'], sc['r'], sc['t'], f_scale=50.0)
    dt = time.perf_counter() - t0
    return dict(seconds=float(dt), tri_seconds=float(out['tri_seconds']), sba_seconds=float(out['sba_seconds']),
                iterations=int(out['iterations']), nfev=int(out['nfev']),
                rmse=float(np.sqrt(np.mean(out['residuals']['after'] ** 2))),
                rmse_before=float(np.sqrt(np.mean(out['residuals']['before'] ** 2))),
                n_points=int(out['n_points']), n_obs=int(out['n_obs']), cores=1,
                path='oracle/ref_port.sba_points_scene (the reference algorithm: pandas TRI + cv2 per observation in scipy TRF)')


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
    m = _capi.MODE_POINTS_AND_EXTRINSICS if mode == 'extrinsics' else _capi.MODE_POINTS_ONLY
    t1_ms = None
    if world > 1:
        # the same fixed-work iteration on this rank ALONE (no communicator yet): the single-GPU time the scaling
        # efficiencies of this run are quoted against
        o1 = dense.default_opts(m, inner_iters=1, force_iters=1, **(dict(point_refine_iters=0) if mode == 'extrinsics' else {}))
        for _ in range(5):
            solver.iterate(o1)
        e0, e1 = ev(), ev()
        e0.record(stream)
        for _ in range(50):
            solver.iterate(o1)
        e1.record(stream); e1.synchronize()
        t1_ms = e0.elapsed_time(e1) / 50
        solver.reset_state()
    if world > 1:
        uid = [dense.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        solver.comm_init(uid[0], rank, world)
    comm_mode = solver.comm_mode()

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
    barrier()
    # EXACTLY K steps per timed block; the block repeats until the timed region lasts --min-time seconds, so that the
    # clock sampler (and the driver's) sees the GPU under this load. ms_per_step is the mean over all timed steps.
    ms, repeats = 0.0, 0
    t_wall0 = time.perf_counter()
    while True:
        e0, e1 = ev(), ev()
        e0.record(stream)
        for _ in range(args.steps):
            solver.iterate(opts)
        e1.record(stream)
        e1.synchronize()
        barrier()
        ms += e0.elapsed_time(e1)
        repeats += 1
        go = torch.tensor([1.0 if (time.perf_counter() - t_wall0) < args.min_time else 0.0], device='cuda')
        if world > 1:
            dist.all_reduce(go, op=dist.ReduceOp.MIN)      # every rank takes the same number of blocks
        if go.item() == 0.0 or repeats >= 10000:
            break
    timed_steps = args.steps * repeats
    if mode == 'extrinsics':
        # per-kernel device times: a profiled pass of the same step right after the timed region (CUDA events between the
        # kernels on the solver's stream; with the events in place the iteration is launched kernel by kernel, the timed
        # region above replays it as a CUDA graph)
        solver.profile(True)
        e0, e1 = ev(), ev()
        e0.record(stream)
        for _ in range(max(args.steps, 50)):
            solver.iterate(opts)
        e1.record(stream)
        e1.synchronize()
        ms_per_step_profiled = e0.elapsed_time(e1) / max(args.steps, 50)
        barrier()
    kern = None
    if mode == 'extrinsics':
        pr = solver.profile_read()
        solver.profile(False)
        # the Schur build (K4a linearisation + K4b accumulation) reads the mask of every slot group and, for existing
        # points, the point and its kept observations; the back-substitution (K6) the same and writes the trial points
        build_bytes = P * 1 + n_pts * 24 + n_obs * 16
        back_bytes = build_bytes + n_pts * 24
        kern = dict(name=f'schur_accum_kernel<{C}>', ms_per_launch=pr['build_ms'], bytes_per_launch=build_bytes,
                    launches_per_step=6 if world == 1 or comm_mode == 'peer' else 9, collective=comm_mode,
                    backsub_ms=pr['backsub_ms'], backsub_bytes=back_bytes, linearise_ms=pr['linearise_ms'],
                    accumulate_ms=pr['accumulate_ms'], solve_ms=pr['solve_ms'], profiled_iterations=pr['iterations'],
                    ms_per_step_profiled=ms_per_step_profiled)
    fp64_peak = dense.fp64_peak_tflops()
    if world > 1:
        tms = torch.tensor([ms], device='cuda', dtype=torch.float64)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms = float(tms.item())
    clk = clocks.stop() if clocks else None
    ms_per_step = ms / timed_steps
    value = world * timed_steps / (ms * 1e-3) * (args.frames / C4_FRAMES)
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
    # Throughput: a sequence of solves with the upload of solve i + 1 and the download of solve i - 1 overlapped
    # with the LM of solve i (ScenePipeline.run_many: two input buffers, two copy streams); the first solve is the warm-up and the clock runs from
    # its completion to the completion of the last, so every timed solve has its upload, triangulation, LM and
    # download inside the region.
    n_seq = max(1, args.e2e_solves) + 1      # at least one timed solve after the warm-up
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    out2_h = torch.empty((P, 3), dtype=torch.float64).pin_memory()      # consecutive solves download into different arrays
    outs = [(out_h if i % 2 == 0 else out2_h).numpy() for i in range(n_seq)]
    res = pipe.run_many([(uv_h.numpy(), lik_h.numpy())] * n_seq, THRESH, outs, rv0, **opts)
    torch.cuda.synchronize()
    t_end = time.perf_counter()
    last = res[-1]
    # from the end of the warm-up solve (its download queued) to the last result on the host: n_seq - 1 whole solves
    times = [(t_end - res[0]['t_solved']) / (n_seq - 1)] * (n_seq - 1)
    iters = sum(r['iterations'] for r in res[1:])
    # Latency: one solve on its own, nothing overlapped (what the breakdown below splits up)
    lat = []
    for i in range(2):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pipe.run(uv_h.numpy(), lik_h.numpy(), THRESH, out_h.numpy(), rv0, **opts)
        torch.cuda.synchronize()
        lat.append(time.perf_counter() - t0)
    latency = min(lat)
    if world > 1:
        lt = torch.tensor([latency], device='cuda', dtype=torch.float64)
        dist.all_reduce(lt, op=dist.ReduceOp.MAX)
        latency = float(lt.item())
    if world > 1:
        dist.barrier()
    bd = pipe.run(uv_h.numpy(), lik_h.numpy(), THRESH, out_h.numpy(), rv0, breakdown=True, **opts).get('breakdown_ms')
    pipe.close()
    total = float(np.sum(times))
    if world > 1:
        tt = torch.tensor([total], device='cuda', dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        total = float(tt.item())
    h2d = uv_h.numel() * 8 + lik_h.numel() * 8
    d2h = out_h.numel() * 8
    return dict(value=world * iters / total * (args.frames / C4_FRAMES), unit=UNIT,
                step='one full solve per GPU shard: H2D observations+likelihoods, TRI, LM until the stopping rule, D2H points; '
                     'a sequence of solves through dense.ScenePipeline.run_many, which uploads solve i+1 and downloads solve i-1 (copy '
                     'streams, second buffers) while solve i runs - every timed solve has all four phases inside the region',
                pipelined=True, solves_timed=len(times),
                solve_ms=1e3 * total / len(times), latency_ms=1e3 * latency,
                latency_note='one solve on its own through ScenePipeline.run, nothing overlapped (breakdown_ms splits this one up)',
                lm_iters_per_solve=iters / len(times), exit_status=last.get('status'),
                cost_initial=last['cost_initial'], cost_final=last['cost_final'],
                launches_per_solve=last.get('launches'), host_syncs_per_solve=last.get('syncs'),
                breakdown_ms=bd, breakdown_note='one extra solve of this rank with a synchronisation between the phases (host clock); '
                                                'lm = joint LM iterations + per-point refinement, polish = points to convergence with the cameras fixed',
                h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h))


def strong_arm(args, ctx):
    """Strong scaling as BASELINE.json states configs 4 and 5: ONE scene of --strong-frames frames (seed 4; 119,490 frames
    = config C4) sharded by frame over all ranks, the reduced camera system exchanged once per LM iteration. Measured in
    the same run as the weak numbers: the fixed-work LM iteration (max over ranks, device time), a full collective solve
    from the TRI start on resident inputs, and the same solve end to end from pinned host buffers."""
    import torch
    import torch.distributed as dist
    from acinoset_b200 import _capi, dense
    rank, world, C, mode, m = ctx['rank'], ctx['world'], ctx['C'], ctx['mode'], ctx['m']
    F = args.strong_frames
    if F >= 500000 and F % 8 == 0 and world in (1, 2, 4, 8):
        # a large scene (config C5) is defined as eight chunks of F/8 frames with seeds 500..507, so that a rank generates
        # only its own frames (the scene is the same at every GPU count)
        per = 8 // world
        parts = [make_shard(F // 8, seed=500 + c) for c in range(rank * per, (rank + 1) * per)]
        uv = np.concatenate([p_['uv'] for p_ in parts]); lik = np.concatenate([p_['lik'] for p_ in parts])
        sc = parts[0]
        a, b = rank * per * (F // 8), (rank + 1) * per * (F // 8)
        del parts
    else:
        sc = ctx['sc'] if (rank == 0 and F == args.frames) else make_shard(F, seed=4)
        a, b = dense.frame_shard(F, rank, world)
        uv = np.ascontiguousarray(sc['uv'].reshape(F, N_KPTS, C, 2)[a:b].reshape(-1, C, 2))
        lik = np.ascontiguousarray(sc['lik'].reshape(F, N_KPTS, C)[a:b].reshape(-1, C))
    k, d, r, t = sc['k'], sc['d'], sc['r'], sc['t']
    Ps = uv.shape[0]
    L = _capi.lib()
    cams = _capi.pack_cams(k, d, r, t)
    uv_d, lik_d = dense.DeviceArray.from_host(uv), dense.DeviceArray.from_host(lik)
    xyz_d = dense.DeviceArray((Ps, 3), np.float64)
    mask_d, ex_d = dense.DeviceArray((Ps,), np.uint8), dense.DeviceArray((Ps,), np.uint8)
    solver = dense.SbaSolver(C, Ps)
    stream = torch.cuda.ExternalStream(solver.stream)
    ev = lambda: torch.cuda.Event(enable_timing=True)
    _capi.check(L.acino_triangulate_ring_dev(uv_d.ptr, lik_d.ptr, THRESH, Ps, C, _capi.ptr(cams), xyz_d.ptr, mask_d.ptr, ex_d.ptr,
                                             ct.c_void_p(solver.stream)))
    solver.set_cameras(k, d, r, t)
    solver.set_observations_dev(uv_d, mask_d)
    if mode == 'extrinsics':
        solver.set_rvec_t(ctx['rv'], ctx['tv'])
    solver.set_points(np.nan_to_num(xyz_d.download()))
    if world > 1:
        uid = [dense.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        solver.comm_init(uid[0], rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def rank_max(x):
        if world == 1:
            return float(x)
        tt = torch.tensor([x], device='cuda', dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    fixed = dense.default_opts(m, inner_iters=1, force_iters=1, **(dict(point_refine_iters=0) if mode == 'extrinsics' else {}))
    for _ in range(5):
        solver.iterate(fixed)
    solver.reset_state()
    barrier()
    st = solver.solve(dense.default_opts(m, **(dict(ftol=1e-10) if mode == 'extrinsics' else {})))
    solve_ms = rank_max(st['solve_ms'])
    solver.reset_state()
    for _ in range(5):
        solver.iterate(fixed)
    barrier()
    ms, steps, t0w = 0.0, 0, time.perf_counter()
    while True:
        e0, e1 = ev(), ev()
        e0.record(stream)
        for _ in range(50):
            solver.iterate(fixed)
        e1.record(stream)
        e1.synchronize()
        barrier()
        ms += e0.elapsed_time(e1)
        steps += 50
        go = torch.tensor([1.0 if (time.perf_counter() - t0w) < 0.3 else 0.0], device='cuda')
        if world > 1:
            dist.all_reduce(go, op=dist.ReduceOp.MIN)
        if go.item() == 0.0:
            break
    ms_per_step = rank_max(ms / steps)
    kern = None
    if mode == 'extrinsics':
        solver.profile(True)          # a profiled pass (kernel by kernel, events between the kernels) for the per-kernel times
        for _ in range(50):
            solver.iterate(fixed)
        barrier()
        pr = solver.profile_read()
        solver.profile(False)
        kern = dict(linearise_ms=pr['linearise_ms'], accumulate_ms=pr['accumulate_ms'], fold_exchange_solve_ms=pr['solve_ms'],
                    backsub_ms=pr['backsub_ms'])
    solver.close()
    for arr in (uv_d, lik_d, xyz_d, mask_d, ex_d):
        arr.free()
    # end to end: pinned host shard -> HBM -> TRI -> collective solve -> points back
    uv_h, lik_h = torch.from_numpy(uv).pin_memory(), torch.from_numpy(lik).pin_memory()
    out_h = torch.empty((Ps, 3), dtype=torch.float64).pin_memory()
    comm = None
    if world > 1 and mode == 'extrinsics':
        uid = [dense.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        comm = (uid[0], rank, world)
    pipe = dense.ScenePipeline(Ps, C, k, d, r, t, m, comm)
    rv0 = (ctx.get('rv'), ctx.get('tv')) if mode == 'extrinsics' else None
    opts = dict(ftol=1e-10) if mode == 'extrinsics' else {}
    times, last = [], None
    for i in range(3):
        barrier()
        t0 = time.perf_counter()
        last = pipe.run(uv_h.numpy(), lik_h.numpy(), THRESH, out_h.numpy(), rv0, **opts)
        torch.cuda.synchronize()
        if i > 0:
            times.append(time.perf_counter() - t0)
    pipe.close()
    e2e_ms = rank_max(1e3 * float(np.mean(times)))
    # the single-GPU time of the same amount of work: this run's own at N = 1, else the rank-alone iteration over a scene
    # of the same size measured before the communicator was made (native_arm)
    t1 = ms_per_step if world == 1 else (ctx.get('t1_ms') if args.frames == F else None)
    out = dict(frames_total=F, frames_per_rank=b - a, scaling='strong', ms_per_step=ms_per_step,
               value=1e3 / ms_per_step * (F / C4_FRAMES), unit=UNIT, timed_steps=steps,
               solve_ms=solve_ms, solve_iterations=st['iterations'], solve_launches=st['launches'], solve_status=st.get('status'),
               cost_initial=st['cost_initial'], cost_final=st['cost_final'],
               e2e_solve_ms=e2e_ms, e2e_iterations=last['iterations'], kernels_ms_rank0=kern,
               single_gpu_ms_per_step=t1,
               efficiency_vs_n1=(t1 / (world * ms_per_step)) if t1 else None)
    if kern and t1:
        work = kern['linearise_ms'] + kern['accumulate_ms'] + kern['backsub_ms']
        out['fixed_cost_us'] = 1e3 * (ms_per_step - work)
        out['fixed_cost_note'] = ('ms_per_step minus the three kernels whose time shrinks with the shard (K4a, K4b, K6) on rank 0: fold, '
                                  'exchange pushes and waits, the 36x36 solve (K5), the decision (K7) and launch gaps')
    return out


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
                cores_note=('the reference algorithm is one Python loop over the observations (cv2.fisheye.projectPoints per '
                            'observation) inside one scipy.optimize.least_squares call, src/lib/sba.py:142-178: it uses one '
                            'thread however many the host has, and the joint points+extrinsics problem cannot be split into '
                            'independent processes'),
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
        c1 = None
        if not args.no_c1:
            try:
                c1 = c1_reference()
            except Exception as e:
                c1 = dict(error=f'{type(e).__name__}: {e}')
        line = dict(impl='reference', metric=METRIC, value=v, unit=UNIT, n_gpus=args.gpus, steps=len(vals), warmup=0,
                    ms_per_step=1e3 / v, higher_is_better=True, scaling='weak', vs_baseline=None, dtype='f64',
                    data='synthetic', config=dict(workload=f'{"C4" if args.frames == C4_FRAMES else "custom"} {"points+extrinsics (Schur)" if mode == "extrinsics" else "points-only"} SBA: '
                                                           f'{args.frames} frames x 20 keypoints x 6 cameras per GPU, dlc_thresh 0.5, 1 px noise',
                                                  masked_frac=args.masked_frac, outlier_frac=args.outlier_frac, mode=mode,
                                                  parallelism=f'frame-shard x{args.gpus}',
                                                  note='reference CPU algorithm on rank 0, single-threaded as in the reference (its per-observation '
                                                       'Python/OpenCV loop has no parallel section); bounded sample scaled to the workload'),
                    cpu_baseline=cb, e2e=dict(value=v, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0,
                    c1=c1, c1_note='config C1 (102 frames) run TO COMPLETION, not extrapolated: compare `c1.seconds` with the native '
                                   "arm's `c1.seconds`; `value` above is the C4 rate scaled from a bounded sample, labelled as such")
        print(json.dumps(line))
        return

    ctx = native_arm(args)
    from acinoset_b200 import _capi
    peak, peak_src = peaks()
    mode, P, C = ctx['mode'], ctx['P'], ctx['C']
    e2e = e2e_arm(args, ctx)
    strong = strong_arm(args, ctx) if args.strong_frames > 0 else None
    if rank != 0:
        import torch.distributed as dist
        dist.destroy_process_group()
        return
    n_obs, n_pts = ctx['n_obs'], ctx['n_pts']
    timed_steps = ctx['timed_steps']
    if mode == 'points':
        bytes_step = lm_points_bytes(n_pts, n_obs) + (P - n_pts) * 2
        kern_ms = ctx['ms_per_step']
        kern_name = f'lm_points_kernel<{C}>'
        launches = timed_steps
    else:
        bytes_step = ctx['kern']['bytes_per_launch']
        kern_ms = ctx['kern']['ms_per_launch']
        kern_name = ctx['kern']['name']
        launches = ctx['kern']['launches_per_step'] * timed_steps
    achieved = bytes_step / (kern_ms * 1e-3) / 1e9
    roofline = dict(bound='hbm', achieved=achieved, peak=peak, unit='GB/s', frac=achieved / peak, traffic=None,
                    kernel=kern_name, algorithmic_bytes_per_launch=int(bytes_step), kernel_ms=kern_ms, peak_source=peak_src,
                    fp64_peak_tflops_measured=ctx['fp64_peak'])
    if mode == 'extrinsics':
        k6 = ctx['kern']
        roofline['kernel'] = f'schur_linearise_kernel<{C}, false> + schur_accum_kernel<{C}> (the Schur build, K4a + K4b; K4b dominates)'
        roofline['note'] = ('the Schur build is bound by the fp64 pipe, not by HBM: `achieved` / `frac` are the ALGORITHMIC bytes of the '
                            'build (mask + point + observations) over K4a + K4b; see `fp64` for the binding roof and `kernels` '
                            'for each kernel against its own traffic')
        roofline['step_share_of_kernels'] = (k6['ms_per_launch'] + k6['backsub_ms']) / k6['ms_per_step_profiled']
        roofline['kernel_times_from'] = (f"{k6['profiled_iterations']} profiled iterations right after the timed region (CUDA events between the kernels "
                                         f"on the solver's stream, kernel-by-kernel launches: {k6['ms_per_step_profiled']:.4f} ms per step; the timed "
                                         f"region replays the iteration as a CUDA graph: {ctx['ms_per_step']:.4f} ms per step)")
        kernels = {
            f'schur_linearise_kernel<{C}, false>': dict(kernel_ms=k6['linearise_ms']),
            f'schur_accum_kernel<{C}>': dict(kernel_ms=k6['accumulate_ms']),
            f'schur_backsub_kernel<{C}>': dict(kernel_ms=k6['backsub_ms'], algorithmic_bytes_per_launch=int(k6['backsub_bytes']),
                                               achieved=k6['backsub_bytes'] / (k6['backsub_ms'] * 1e-3) / 1e9,
                                               frac=k6['backsub_bytes'] / (k6['backsub_ms'] * 1e-3) / 1e9 / peak),
            'fold + exchange + schur_solve_kernel': dict(kernel_ms=k6['solve_ms']),
        }
        roofline['kernels'] = kernels
        # DRAM traffic and executed fp64 flops per launch come from a committed ncu capture of the same workload; they are
        # attached only when the capture was taken with THESE kernel sources (hash of csrc/) and this workload
        sha = kernel_source_sha()
        roofline['kernel_source_sha'] = sha
        try:
            with open(os.path.join(ROOT, 'profiles', 'r02_ncu_counts.json')) as f:
                ncu = json.load(f)
            w = ncu['workload']
            same = (args.frames, args.masked_frac, args.outlier_frac) == (w['frames'], w['masked_frac'], w['outlier_frac'])
            if not same or ncu.get('kernel_source_sha') != sha:
                roofline['ncu_counts'] = (f"not attached: profiles/r02_ncu_counts.json was captured for sources {ncu.get('kernel_source_sha')} / "
                                          f"workload {w}, this run is {sha} / frames {args.frames}")
            else:
                src = 'profiles/r02_ncu_counts.json (ncu --set full of scripts/profile_target.py, same sources and workload)'
                flops = dram = 0
                for name, blk in kernels.items():
                    kk = ncu['kernels'].get(name)
                    if not kk:
                        continue
                    ms_k = blk['kernel_ms']
                    blk['traffic'] = kk['dram_bytes']
                    blk['traffic_gbs'] = kk['dram_bytes'] / (ms_k * 1e-3) / 1e9
                    blk['traffic_frac_of_hbm_peak'] = blk['traffic_gbs'] / peak
                    tf = kk['fp64_flops'] / (ms_k * 1e-3) / 1e12
                    blk['fp64'] = dict(flops_per_launch=kk['fp64_flops'], achieved_tflops=tf, peak_tflops=ctx['fp64_peak'],
                                       frac=tf / ctx['fp64_peak'], pipe_active_pct_ncu=kk['fp64_pipe_pct'])
                    if 'linearise' in name or 'accum' in name:
                        flops += kk['fp64_flops']; dram += kk['dram_bytes']
                roofline['traffic'] = dram
                roofline['traffic_source'] = 'ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, K4a + K4b; ' + src
                roofline['traffic_gbs'] = dram / (kern_ms * 1e-3) / 1e9
                roofline['traffic_frac_of_hbm_peak'] = roofline['traffic_gbs'] / peak
                tf = flops / (kern_ms * 1e-3) / 1e12
                roofline['fp64'] = dict(flops_per_launch=flops, achieved_tflops=tf, peak_tflops=ctx['fp64_peak'], frac=tf / ctx['fp64_peak'],
                                        source='executed DFMA x2 + DMUL + DADD of K4a + K4b, ncu SASS page; ' + src)
        except (OSError, KeyError, ValueError) as e:
            roofline['ncu_counts'] = f'not attached: {type(e).__name__}'
    tri_gbs = tri_bytes(P, C) / (ctx['tri_ms'] * 1e-3) / 1e9
    cb = None
    if not args.no_cpu_baseline and world == 1:
        cb = cpu_baseline(args.cpu_frames, mode, n_obs)
    st = ctx['st_solve']
    c1 = None
    if not args.no_c1 and world == 1:
        try:
            c1 = c1_native()
        except Exception as e:          # the headline must not depend on the comparison block
            c1 = dict(error=f'{type(e).__name__}: {e}')
    line = dict(metric=METRIC, value=ctx['value'], unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
                timed_steps=timed_steps, timed_blocks=ctx['repeats'], timed_region_s=ctx['ms'] * 1e-3,
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
                           cost_final=st['cost_final'], n_unconverged=st['n_unconverged'], init_ms=st.get('init_ms'),
                           polish_ms=st.get('polish_ms'), host_syncs=st.get('syncs'),
                           iterations_per_s=1e3 * st['iterations'] / st['solve_ms'] if st['solve_ms'] else None,
                           note='full solve from the TRI start on resident inputs (refinement and polishing on): the '
                                'solve-representative rate; `value` is the fixed-work LM iteration'),
                strong=strong, c1=c1, single_rank_ms_per_step=ctx.get('t1_ms'),
                tri=dict(ms=ctx['tri_ms'], points_per_s=P / (ctx['tri_ms'] * 1e-3), gbs=tri_gbs, frac_hbm=tri_gbs / peak,
                         algorithmic_bytes=tri_bytes(P, C)))
    print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
