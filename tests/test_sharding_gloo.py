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


def _worker_temperature(rank, world, port, G, N, ret):
    """the exchange step of the material-temperature extension (rt_params::cv): every rank forms the group sums A, E, D
    over the groups it owns, the all-reduce (sharding.allreduce_tensor_, the function the device hook calls) makes
    them whole, and the update computed from them equals the oracle's"""
    import importlib
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    for pth in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if pth not in sys.path:
            sys.path.insert(0, pth)
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rt = importlib.import_module("radiative-transfer_b200")
        sh = importlib.import_module("radiative-transfer_b200.sharding")
        import oracle_py
        from parity_util import host_setup, make_oracle, synthetic_parameters
        p = synthetic_parameters(rt, M=4, G=G, N=N, dt=1e-3, steps=1, bc_left=1, bc_right=1)
        p.T_cell = 1.0 + 0.2 * np.sin(np.arange(N) / 11.0)
        p.cv = 0.8
        mu, wt, e, kap = host_setup(rt, p)
        o = make_oracle(oracle_py, rt, p)
        o.run_stages(4)
        phi = np.ascontiguousarray(o.moments()[0])  # [G, N], the state the update saw
        T1, A, E, D = o.temperature(with_sums=True)
        o.close()
        g0, gn = sh.group_range(G, world, rank)
        W = float(np.sum(wt))
        part = np.zeros((3, N))
        # B_g(T(c)) and dB_g/dT per cell through the oracle's Planck
        Bc = np.zeros((G, N)); dBc = np.zeros((G, N))
        for c in range(N):
            Bc[:, c], dBc[:, c] = oracle_py.planck(float(p.T_cell[c]), e)
        sig = p.rho * kap[g0:g0 + gn, None]
        part[0] = (sig * phi[g0:g0 + gn]).sum(axis=0)
        part[1] = (W * (sig * Bc[g0:g0 + gn])).sum(axis=0)
        part[2] = (W * (sig * dBc[g0:g0 + gn])).sum(axis=0)
        t = torch.from_numpy(part.copy())
        sh.allreduce_tensor_(t, dist)
        tot = t.numpy()
        Tn = p.T_cell + p.dt * (tot[0] - tot[1]) / (p.rho * p.cv + p.dt * tot[2])
        ok = (np.allclose(tot[0], A, rtol=1e-12) and np.allclose(tot[1], E, rtol=1e-12) and np.allclose(tot[2], D, rtol=1e-12)
              and np.allclose(Tn, T1, rtol=1e-12))
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_temperature_group_sums_over_two_ranks_gloo():
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker_temperature, args=(r, world, port, 7, 40, ret)) for r in range(world)]
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
