"""GPU parity of the per-step history (rt_set_step_history): radiation energy and material temperature at EVERY BDF2 step.

north_star: "1e-9 relative on scalar intensity, radiation energy and material temperature at each BDF2 step".  The scalar
intensity per step is covered by the snapshot tests; here the other two, as delivered by the library for a whole run
(no per-step round trip): energy[k][c] = (1/c) sum_g phi_g(c) after step k against the same sum of the oracle's phi
(reference order of accumulation over m, solver.cpp:191-204; the sum over groups is the test's), T[k][c] against the
oracle's update (PARITY UNPINNED for T: the reference has no temperature update)."""
import numpy as np
import pytest

from parity_util import RTOL, load_case, load_golden, make_oracle, synthetic_parameters

pytestmark = pytest.mark.gpu
C_LIGHT = 299.792458  # Constants::SPEED_OF_LIGHT (cm/sh), include/Constants.h


def _oracle_energy(oracle, rt, p, steps, per_step, with_T=False):
    o = make_oracle(oracle, rt, p, num_threads=4)
    E, T = [], []
    for _ in range(steps):
        o.run_stages(per_step)
        E.append(o.moments()[0].sum(axis=0) / C_LIGHT)
        if with_T:
            T.append(o.temperature().copy())
    o.close()
    return np.array(E), (np.array(T) if with_T else None)


@pytest.mark.parametrize("ts,bcl,bcr,V", [(3, 2, 1, 1.0), (3, 1, 1, 0.0), (1, 0, 0, 1.0), (2, 2, 1, 1.0)])
def test_energy_of_every_step_vs_oracle(rt, oracle, ts, bcl, bcr, V):
    steps, per_step = 4, (4 if ts == 3 else 1)
    p = synthetic_parameters(rt, M=8, G=6, N=3000, dt=1e-3, V=V, steps=steps, bc_left=bcl, bc_right=bcr, ts_method=ts)
    E_ref, _ = _oracle_energy(oracle, rt, p, steps, per_step)
    with rt.Solver(p) as s:
        E, _ = s.set_step_history(steps)
        s.solve()
        assert np.max(np.abs(E - E_ref) / E_ref) <= RTOL
        assert s.plan().step_kernel == (4 if ts == 3 else s.plan().step_kernel)  # BDF2: phi comes out of the step kernel itself
        whole = E.copy()
    # the same rows from a run in pieces (rows are delivered by the call that ran the steps), from group batches (batches
    # add their share to the same row) and from shards (the multi handle adds the shards' shares)
    with rt.Solver(p) as s:
        E, _ = s.set_step_history(steps)
        s.run_stages(per_step)
        assert np.array_equal(E[0], whole[0]) and not E[1:].any()
        s.run_stages(3 * per_step)
        assert np.array_equal(E, whole)
    with rt.Solver(p, groups_per_batch=4, fetch_psi=False) as s:
        assert s.plan().n_batches == 2
        E, _ = s.set_step_history(steps)
        s.solve()
        assert np.max(np.abs(E - whole) / whole) <= 1e-14
    with rt.Solver(p, devices=[0, 0, 0]) as s:
        E, _ = s.set_step_history(steps)
        s.solve()
        assert np.max(np.abs(E - whole) / whole) <= 1e-14


@pytest.mark.parametrize("name", ["multi_group_equilibrium", "llnl_slab_test"])
def test_energy_on_the_reference_cases(rt, name):
    """the reference's own .prm cases: the history's last row against the golden phi of the UNMODIFIED reference"""
    p = load_case(rt, name)
    g = load_golden(name)
    steps = p.max_timesteps
    with rt.Solver(p) as s:
        E, _ = s.set_step_history(steps)
        s.solve()
        ref = g["phi"].sum(axis=0) / C_LIGHT
        assert np.max(np.abs(E[-1] - ref) / ref) <= RTOL
        ks = sorted(int(k[len("psi_steps"):]) for k in g if k.startswith("psi_steps"))
        assert ks
        for k in ks:  # and at the intermediate steps the goldens hold: sum_m w_m psi of the snapshot
            phi_k = np.tensordot(g["wt"], g[f"psi_steps{k}"], axes=(0, 0))
            ref_k = phi_k.sum(axis=0) / C_LIGHT
            assert np.max(np.abs(E[k - 1] - ref_k) / ref_k) <= RTOL, k


def test_temperature_and_energy_of_every_step_with_the_update(rt, oracle):
    N, steps = 2000, 4
    p = synthetic_parameters(rt, M=8, G=12, N=N, dt=1e-3, V=1.0, steps=steps, bc_left=2, bc_right=1)
    x = (np.arange(N) + 0.5) / N
    p.T_cell = 0.8 + 0.4 * np.sin(2 * np.pi * x) ** 2
    p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
    p.cv = 0.7
    E_ref, T_ref = _oracle_energy(oracle, rt, p, steps, 4, with_T=True)
    with rt.Solver(p) as s:
        E, T = s.set_step_history(steps, temperature=True)
        s.solve()
        assert np.max(np.abs(T - T_ref) / T_ref) <= RTOL
        assert np.max(np.abs(E - E_ref) / E_ref) <= RTOL
        assert np.max(np.abs(T[-1] - T[0])) > 1e-5  # the temperature did move


def test_history_contract(rt):
    p = synthetic_parameters(rt, M=4, G=3, N=500, dt=1e-3, steps=3)
    with rt.Solver(p) as s:
        with pytest.raises(rt.capi.RtError, match="cv > 0"):
            s.set_step_history(3, temperature=True)
        E, _ = s.set_step_history(2)  # capacity below the number of steps: the later steps are not recorded
        s.solve()
        assert E.all()
        full = None
    with rt.Solver(p) as s:
        full, _ = s.set_step_history(3)
        s.solve()
    assert np.array_equal(full[:2], E)
