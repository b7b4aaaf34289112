"""Row-sharded search on real GPUs: NCCL all-gather + merge kernel, against the oracle on the full
row set.  Runs with world_size 1 always (still the CUDA path of sharded.py) and 2 when the box has
at least two GPUs."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("exchange", ["p2p", "nccl"])
@pytest.mark.parametrize("world", [1, 2])
def test_sharded_search_nccl(world, exchange):
    """exchange: the one-shot all-gather over peer memory (default where the GPUs allow it) and NCCL's."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "_sharded_gpu_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, B200KNN_EXCHANGE=exchange))
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "SHARDED-GPU-OK world=%d" % world in out.stdout
    if exchange == "nccl":
        assert "exchange=nccl" in out.stdout
