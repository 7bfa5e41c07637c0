"""CPU test of the N>1 host path: two ranks over gloo (127.0.0.1), energy groups sharded, results gathered.

The ranks run the CPU oracle on the whole problem and keep only the rows of the groups they own (groups never couple,
SURVEY F6) -- what is under test is the sharding arithmetic and the torch.distributed plumbing bench.py and the
Solver use at N > 1: group_range, gather_group_rows, reduce_scalar, reduce_vector_sum."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, G, N, ret):
    import importlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    for pth in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if pth not in sys.path:
            sys.path.insert(0, pth)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rt = importlib.import_module("radiative-transfer_b200")
        sh = importlib.import_module("radiative-transfer_b200.sharding")
        import oracle_py
        from parity_util import make_oracle, synthetic_parameters
        p = synthetic_parameters(rt, M=4, G=G, N=N, dt=1e-3, steps=1)
        o = make_oracle(oracle_py, rt, p)
        o.solve()
        phi = np.ascontiguousarray(o.moments()[0])  # [G, N]
        gt = o.group_terms()
        o.close()
        g0, gn = sh.group_range(G, world, rank)
        full = sh.gather_group_rows(phi[g0:g0 + gn].copy(), G, dist)
        sums = sh.reduce_vector_sum(np.array([gt["B"][g0:g0 + gn].sum(), gt["dBdT"][g0:g0 + gn].sum(), float(gn)]), dist)
        tmax = sh.reduce_scalar(10.0 + rank, "max", dist)
        tsum = sh.reduce_scalar(float(gn), "sum", dist)
        ok = (np.array_equal(full, phi) and tmax == 10.0 + world - 1 and tsum == G and sums[2] == G
              and abs(sums[0] - gt["B"].sum()) <= 1e-12 * abs(gt["B"].sum()))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("G", [6, 7])
def test_two_ranks_gloo(G):
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, port, G, 257, ret)) for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(timeout=240)
    for pr in procs:
        if pr.is_alive():
            pr.terminate()
            pytest.fail("gloo worker hung")
        assert pr.exitcode == 0
    assert dict(ret) == {0: True, 1: True}
