"""World-size-2 CPU test (gloo) of the frame-sharded reduction the multi-GPU solver relies on: cameras are shared,
points are disjoint per shard, so the reduced camera system of the whole scene is the SUM over ranks of the shard systems
(what acino_sba_* all-reduces with NCCL), and the cost / predicted reduction are sums of shard scalars. The per-shard
systems come from the numpy LM oracle; no GPU is involved."""
import os
import socket
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from acinoset_b200 import dense, synth


def test_frame_shard_partition():
    for n, w in [(119490, 8), (10, 3), (7, 8), (0, 2)]:
        spans = [dense.frame_shard(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        dense.frame_shard(10, 2, 2)


def _problem():
    from oracle import fisheye_np as fe
    k, d, r, t, _ = synth.load_rig()
    rng = np.random.default_rng(5)
    n_frames, n_k = 6, 4
    P = n_frames * n_k
    Xg = np.stack([rng.uniform(-1, 5, P), rng.uniform(5, 8, P), rng.uniform(0, 1, P)], 1)
    pidx, cidx, obs = [], [], []
    for p in range(P):
        for c in range(6):
            if rng.uniform() < 0.6:
                pidx.append(p); cidx.append(c)
                obs.append(fe.project_points_fisheye(Xg[p], k[c], d[c], r[c], t[c])[0] + rng.normal(0, 0.5, 2))
    return k, d, r, t.reshape(6, 3), Xg + rng.normal(0, 0.01, Xg.shape), np.array(obs), np.array(pidx), np.array(cidx), n_frames, n_k


def _reduced(prob, R, T, X, lam):
    H, g, cost = prob.linearise(R, T, X)
    S, rhs, _, D = prob.reduced_system(H, g, lam)
    nc = 6 * prob.C
    # the damping term lam*diag(U) is added after the sum (K5 does it on the all-reduced diag(U)): remove it per shard
    S = S - lam * np.diag(D[:nc])
    return S, rhs, D[:nc], cost


def _worker(rank, world, port, out):
    from oracle import sba_lm_np as lm_np
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    k, d, r, t, X, obs, pidx, cidx, n_frames, n_k = _problem()
    f0, f1 = dense.frame_shard(n_frames, rank, world)
    sel = (pidx >= f0 * n_k) & (pidx < f1 * n_k)
    prob = lm_np.Problem(obs[sel], pidx[sel] - f0 * n_k, cidx[sel], k, d, (f1 - f0) * n_k, f_scale=1.0)
    S, rhs, D, cost = _reduced(prob, r, t, X[f0 * n_k:f1 * n_k], 1e-3)
    buf = torch.from_numpy(np.concatenate([S.ravel(), rhs, D, [cost]]))
    dist.all_reduce(buf, op=dist.ReduceOp.SUM)
    if rank == 0:
        np.save(out, buf.numpy())
    dist.destroy_process_group()


def test_shard_sum_equals_whole_scene(tmp_path):
    from oracle import sba_lm_np as lm_np
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / 'reduced.npy')
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    k, d, r, t, X, obs, pidx, cidx, n_frames, n_k = _problem()
    prob = lm_np.Problem(obs, pidx, cidx, k, d, len(X), f_scale=1.0)
    S, rhs, D, cost = _reduced(prob, r, t, X, 1e-3)
    want = np.concatenate([S.ravel(), rhs, D, [cost]])
    assert np.allclose(got, want, rtol=1e-10, atol=1e-8 * np.abs(want).max())
