"""The product API's multi-GPU path on real GPUs: RadarOperator under torch.distributed.run with one process
per GPU (NCCL), against the single-GPU result, bit for bit.  Needs >= 2 GPUs (skipped otherwise; the driver's
single-GPU box skips it, `gpurun --gpus 2 -- python -m pytest tests/test_gpu_sharded.py -m gpu` runs it; the log of
that run is kept under profiles/)."""
import os
import socket
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


@pytest.mark.parametrize('world', [2])
def test_sharded_scan_equals_single_gpu(world):
    if _n_gpus() < world:
        pytest.skip('needs {} GPUs'.format(world))
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world),
           '--master-addr', '127.0.0.1', '--master-port', str(port),
           os.path.join(ROOT, 'tests', 'helpers', 'sharded_worker.py')]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, universal_newlines=True, timeout=600)
    assert r.returncode == 0, r.stdout[-4000:]
    for k in range(world):
        assert 'SHARDED_OK rank={}'.format(k) in r.stdout, r.stdout[-4000:]
