"""Multi-GPU parity (needs >= 2 GPUs, skipped otherwise): the frame-sharded collective points+extrinsics solve takes the
same iterates as the single-GPU solve of the whole scene (scripts/check_collective.py under torchrun, NCCL).
The CPU-side counterpart (gloo, world 2) is tests/test_multirank_cpu.py."""
import os
import subprocess
import sys
import pytest

pytestmark = pytest.mark.gpu


def test_sharded_solve_equals_single_gpu_solve():
    from acinoset_b200 import _capi
    n = _capi.lib().acino_device_count()
    if n < 2:
        pytest.skip('needs two GPUs')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29541', os.path.join(root, 'scripts', 'check_collective.py'), '2000']
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert 'COLLECTIVE_CHECK OK' in out.stdout
