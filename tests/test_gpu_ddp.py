"""-m gpu, needs >= 2 GPUs (`gpurun --gpus 2`): the overlapped NCCL data-parallel step on hardware, ragged captions, against
the single-process step on the concatenated batch (tests/ddp_worker.py).  Skipped on a 1-GPU box; the same check runs inside
`bench.py --gpus N` (N > 1) as `ddp_check`, which is what the driver's scaling run records."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_overlapped_ddp_step_equals_single_process_on_ragged_captions():
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2', '--master-addr', '127.0.0.1',
           '--master-port', str(port), os.path.join(ROOT, 'tests', 'ddp_worker.py')]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    line = [l for l in r.stdout.splitlines() if l.startswith('DDP_CHECK')]
    assert r.returncode == 0, r.stdout[-4000:]
    assert line, r.stdout[-2000:]
    print(line[-1])
