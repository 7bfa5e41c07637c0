"""GPU parity of the material-temperature update (extension; rt_params::cv) through the C ABI.

PARITY UNPINNED: the reference has no temperature update, so the checker is the oracle's own restatement
(oracle/rt_oracle.cpp rto_temperature_update; its properties: tests/test_oracle_temperature.py).  Tolerance: 1e-9
relative on T(c), the group sums and the state at each time step, like the pinned path."""
import ctypes as C
import os

import numpy as np
import pytest

from parity_util import CELL_FLOOR, make_oracle, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu

RTOL = 1e-9


def _fields(p, seed=5):
    rng = np.random.Generator(np.random.PCG64(seed))
    x = (np.arange(p.N) + 0.5) / p.N
    p.T_cell = 0.8 + 0.4 * np.sin(2 * np.pi * x) ** 2
    p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
    p.kappa_cell = np.asarray(p.group_kappa)[:, None] * 10.0 ** rng.uniform(-0.5, 0.5, size=(p.G, p.N))
    return p


def _oracle_steps(oracle, rt, p, steps, per_step, threads=4):
    o = make_oracle(oracle, rt, p, num_threads=threads)
    out = []
    for _ in range(steps):
        o.run_stages(per_step)
        T, A, E, D = o.temperature(with_sums=True)
        out.append(dict(T=T, A=A, E=E, D=D, psi=o.psi(), phi=o.moments()[0]))
    B = np.maximum(o.group_terms()["B"], 1e-300)
    o.close()
    return out, B


def _check_step(s, ref, B, groups=slice(None)):
    T, A, E, D = s.temperature(with_sums=True)
    assert np.max(np.abs(T - ref["T"]) / ref["T"]) <= RTOL
    for got, want in ((A, ref["A"]), (E, ref["E"]), (D, ref["D"])):
        assert np.max(np.abs(got - want) / np.abs(want)) <= RTOL
    ok, w, _ = state_ok(s.psi_mat_ref, ref["psi"][:, groups, :], B[groups], floor=CELL_FLOOR)
    assert ok, f"psi err/tol {w:.3e}"
    ok, w, _ = state_ok(s.moments()[0], ref["phi"][groups, :], B[groups], group_axis=0, floor=CELL_FLOOR)
    assert ok, f"phi err/tol {w:.3e}"


@pytest.mark.parametrize("ts,bcl,bcr,V", [(3, 2, 1, 1.0), (3, 1, 1, 0.0), (3, 0, 0, 1.0), (1, 1, 1, 1.0), (2, 2, 1, 1.0)])
def test_temperature_update_vs_oracle_each_step(rt, oracle, ts, bcl, bcr, V):
    p = synthetic_parameters(rt, M=8, G=6, N=1500, dt=1e-3, V=V, steps=3, bc_left=bcl, bc_right=bcr, ts_method=ts)
    _fields(p)
    p.cv = 0.7
    per = 4 if ts == 3 else 1
    ref, B = _oracle_steps(oracle, rt, p, 3, per)
    assert np.max(np.abs(ref[-1]["T"] - p.T_cell)) > 1e-4  # the update does something
    with rt.Solver(p) as s:
        for k in range(3):
            s.run_stages(per)
            _check_step(s, ref[k], B)
        tm = s.timing()
        fused = ts == 3 and os.environ.get("RT_CUDA_VAR_FUSED", "1") != "0" and os.environ.get("RT_CUDA_FUSED", "1") != "0" \
            and os.environ.get("RT_CUDA_VAR_THREADS", "64") != "0" and os.environ.get("RT_CUDA_VAR_CPT", "4") == "4"
        assert tm.tupdate_count == 3 and tm.step_launches == (3 if fused else 0)  # BDF2 steps are fused launches


def test_scalar_inputs_with_cv_switch_to_the_cell_dependent_path(rt, oracle):
    """no per-cell arrays given: cv > 0 alone builds them from the scalars"""
    p = synthetic_parameters(rt, M=4, G=5, N=700, dt=1e-3, V=1.0, steps=2, bc_left=2, bc_right=2)
    p.cv = 2.0
    ref, B = _oracle_steps(oracle, rt, p, 2, 4)
    with rt.Solver(p) as s:
        s.solve()
        _check_step(s, ref[-1], B)
        assert np.min(s.temperature()) < 1.0 - 1e-6  # the leaky right wall has cooled the slab


def test_equilibrium_stays_and_reset_restores_the_initial_temperature(rt, oracle):
    p = synthetic_parameters(rt, M=4, G=4, N=600, dt=1e-3, V=0.0, steps=2, bc_left=1, bc_right=1)
    p.cv = 0.5
    with rt.Solver(p) as s:
        s.solve()
        assert np.max(np.abs(s.temperature() - 1.0)) < 1e-10
    _fields(p)
    ref, B = _oracle_steps(oracle, rt, p, 2, 4)
    with rt.Solver(p) as s:
        s.solve()
        T_end = s.temperature()
        assert np.max(np.abs(T_end - p.T_cell)) > 1e-4
        s.reset()
        assert np.array_equal(s.temperature(), p.T_cell)
        s.solve()  # the second run repeats the first one exactly
        assert np.array_equal(s.temperature(), T_end)
        _check_step(s, ref[-1], B)


@pytest.mark.parametrize("N", [1, 5, 127, 513])
def test_ragged_sizes(rt, oracle, N):
    p = synthetic_parameters(rt, M=4, G=3, N=N, dt=1e-3, V=1.0, steps=2, bc_left=2, bc_right=1)
    _fields(p)
    p.cv = 1.0
    ref, B = _oracle_steps(oracle, rt, p, 2, 4, threads=2)
    with rt.Solver(p) as s:
        s.solve()
        _check_step(s, ref[-1], B)


def test_many_groups_exercise_the_chunked_group_sum(rt, oracle):
    """more groups than one chunk of the group sum (32): chunk partials are added in ascending order"""
    p = synthetic_parameters(rt, M=4, G=75, N=300, dt=1e-3, V=1.0, steps=2, bc_left=1, bc_right=1)
    _fields(p)
    p.cv = 1.0
    ref, B = _oracle_steps(oracle, rt, p, 2, 4)
    with rt.Solver(p) as s:
        s.solve()
        _check_step(s, ref[-1], B)


def test_group_shards_coupled_through_the_allreduce_hook(rt, oracle):
    """two handles on one GPU, each owning a group shard, driven from two host threads; their all-reduce hooks meet at a
    barrier, add the two partial-sum buffers and write the total back into both -- what NCCL does across GPUs
    (tests/test_gpu_temperature_nccl.py runs the real thing under torchrun)"""
    import threading
    from importlib import import_module

    import torch
    sharding = import_module(rt.__name__ + ".sharding")
    capi = rt.capi
    p = synthetic_parameters(rt, M=4, G=9, N=800, dt=1e-3, V=1.0, steps=3, bc_left=2, bc_right=1)
    _fields(p)
    p.cv = 0.9
    ref, B = _oracle_steps(oracle, rt, p, 3, 4)
    barrier = threading.Barrier(2, timeout=120)
    parked = {}

    def make_hook(me, other):
        def hook(buf, count, stream, user):
            try:
                torch.cuda.synchronize()
                parked[me] = torch.as_tensor(sharding._DevicePointer(int(buf), int(count)), device="cuda:0")
                barrier.wait()
                total = parked[me] + parked[other]
                torch.cuda.synchronize()
                barrier.wait()
                parked[me].copy_(total)
                torch.cuda.synchronize()
                return 0
            except Exception as exc:
                print("hook:", repr(exc))
                return 1
        return capi.ALLREDUCE_FN(hook)

    ha, hb = make_hook("a", "b"), make_hook("b", "a")
    with rt.Solver(p, group_begin=0, group_count=4) as a, rt.Solver(p, group_begin=4, group_count=5) as b:
        capi.check(capi.lib().rt_set_allreduce(a._h, ha, None))
        capi.check(capi.lib().rt_set_allreduce(b._h, hb, None))
        for k in range(3):
            errs = []

            def drive(solver):
                try:
                    solver.run_stages(4)
                except Exception as exc:
                    errs.append(exc)
                    barrier.abort()

            ta, tb = threading.Thread(target=drive, args=(a,)), threading.Thread(target=drive, args=(b,))
            ta.start(); tb.start()
            ta.join(180); tb.join(180)
            assert not errs, errs
            _check_step(a, ref[k], B, slice(0, 4))
            _check_step(b, ref[k], B, slice(4, 9))


def test_grey_single_group_with_temperature_update(rt, oracle):
    """G = 1 (the shape of prm/single_group.prm): the last-group remainder IS the group, the edge stencils take their
    G = 1 branch (correction.cpp:206-211 overwrite), the group sum has one term"""
    p = synthetic_parameters(rt, M=2, G=1, N=400, dt=1e-3, V=1.0, steps=3, bc_left=2, bc_right=2)
    p.T_cell = 0.9 + 0.2 * np.sin(np.arange(p.N) / 23.0) ** 2
    p.cv = 0.6
    ref, B = _oracle_steps(oracle, rt, p, 3, 4, threads=1)
    with rt.Solver(p) as s:
        for k in range(3):
            s.run_stages(4)
            _check_step(s, ref[k], B)


def test_a_runaway_temperature_is_reported_not_carried_on(rt):
    """The explicitly coupled update is only conditionally stable: on a stiff slab (512 narrow groups with opacities spread
    over three decades, c dt = 30 cells) T leaves the physical range within a few steps -- in the oracle's restatement
    exactly as on the device (checked on the CPU, tools/gpu_debug_tc.py).  The solve must say so instead of returning NaN."""
    N = 256
    p = synthetic_parameters(rt, M=8, G=512, N=N, dt=1e-3, V=1.0, steps=10, bc_left=1, bc_right=1)
    x = (np.arange(N) + 0.5) / N
    p.T_cell = 1.0 + 0.3 * np.sin(2 * np.pi * x) ** 2
    p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
    p.kappa_cell = np.asarray(p.group_kappa)[:, None] * (1.0 + 0.5 * np.sin(7 * np.pi * x))[None, :]
    p.cv = 1.0
    with rt.Solver(p, fetch_psi=False) as s:
        with pytest.raises(rt.capi.RtError, match="left the physical range"):
            s.solve()
        # the same slab at a step inside the stability range runs through and stays physical
    p.dt, p.max_timesteps = 1e-5, 4
    with rt.Solver(p, fetch_psi=False) as s:
        s.solve()
        T = s.temperature()
        assert np.isfinite(T).all() and T.min() > 0.9 and T.max() < 1.4
