"""Multi-GPU: the material-temperature update's all-reduce over NCCL (one process per GPU under torchrun).  Needs two
GPUs on the box (`gpurun --gpus 2`); skipped otherwise."""
import os
import re
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["nccl", "torch"])
def test_two_ranks_nccl_match_the_oracle(exchange):
    """exchange = nccl: the library's own communicator and ncclAllReduce (rt_set_nccl, the default) ; torch: the host's
    torch.distributed.all_reduce called back through rt_set_allreduce"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "nccl_temperature_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, RT_CUDA_ALLREDUCE=exchange))
    # the two ranks write to one pipe: their report lines may arrive glued together, so records are counted, not lines
    records = re.findall(r"NCCL_TEMPERATURE rank=(\d)/2 groups=\[\d+,\d+\) ok=(True|False)", r.stdout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert sorted(records) == [("0", "True"), ("1", "True")], r.stdout[-2000:]
