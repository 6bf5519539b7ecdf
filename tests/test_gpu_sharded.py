"""View-sharded runs through the drop-in entry points (NCCL, one GPU per rank): every rank returns what the unsharded run
returns (config.gather_views on) or its own view range of it (off); counting noise is independent of the sharding.
Needs two GPUs -- skipped on a one-GPU box (the host-side sharding logic is covered with gloo in test_host_logic.py)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs (NCCL does not run two ranks on one device)')
def test_sharded_dropin_calls_equal_the_unsharded_run():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    ws = 2
    procs = []
    for r in range(ws):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(ws), MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, 'dropin_worker.py')], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=240)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f'dropin worker {r}/{ws} ok' in out, out[-3000:]
