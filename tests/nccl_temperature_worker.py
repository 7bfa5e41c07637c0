"""Worker for tests/test_gpu_temperature_nccl.py, launched under torchrun (one rank per GPU, NCCL):
every rank owns a contiguous block of energy groups, the material-temperature update (rt_params::cv) all-reduces its
three group sums through the library's all-reduce hook = torch.distributed NCCL on the library's own device buffer.
Checked against the CPU oracle of the WHOLE problem on every rank, each time step.  Prints one line per rank."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for pth in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if pth not in sys.path:
        sys.path.insert(0, pth)


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rt = importlib.import_module("radiative-transfer_b200")
    sh = importlib.import_module("radiative-transfer_b200.sharding")
    import oracle_py
    from parity_util import CELL_FLOOR, make_oracle, state_ok, synthetic_parameters
    G, N, steps = 12, 2000, 3
    p = synthetic_parameters(rt, M=8, G=G, N=N, dt=1e-3, V=1.0, steps=steps, bc_left=2, bc_right=1)
    x = (np.arange(N) + 0.5) / N
    p.T_cell = 0.8 + 0.4 * np.sin(2 * np.pi * x) ** 2
    p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
    p.cv = 0.7
    o = make_oracle(oracle_py, rt, p, num_threads=4)
    g0, gn = sh.group_range(G, world, rank)
    worst = 0.0
    ok = True
    with rt.Solver(p, device=local, group_begin=g0, group_count=gn, dist=dist) as s:
        for k in range(steps):
            o.run_stages(4)
            s.run_stages(4)
            T_ref, A, E, D = o.temperature(with_sums=True)
            T, a, e, d = s.temperature(with_sums=True)
            errs = [np.max(np.abs(T - T_ref) / T_ref), np.max(np.abs(a - A) / A), np.max(np.abs(e - E) / E), np.max(np.abs(d - D) / D)]
            worst = max(worst, *errs)
            B = np.maximum(o.group_terms()["B"], 1e-300)
            good, w, _ = state_ok(s.psi_mat_ref, o.psi()[:, g0:g0 + gn, :], B[g0:g0 + gn], floor=CELL_FLOOR)
            ok = ok and good and max(errs) <= 1e-9
        moved = float(np.max(np.abs(s.temperature() - p.T_cell)))
        tm = s.timing()
    o.close()
    # every rank must hold the same temperature field bit for bit (the all-reduce result is identical on all ranks)
    t = torch.from_numpy(T.copy()).cuda()
    lo, hi = t.clone(), t.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same = bool(torch.equal(lo, hi))
    print(f"NCCL_TEMPERATURE rank={rank}/{world} groups=[{g0},{g0 + gn}) ok={ok and same and moved > 1e-4} worst_rel={worst:.3e} "
          f"moved={moved:.3e} identical_T={same} updates={tm.tupdate_count}", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if (ok and same and moved > 1e-4) else 1)


if __name__ == "__main__":
    main()
