"""CPU: the oracle's material-temperature update.

PARITY UNPINNED: the reference has no temperature update (T is a scalar that never changes), so there is
nothing of the reference's to compare with.  The restatement in oracle/rt_oracle.cpp (rto_temperature_update) is the
definition the device path is checked against; here its own properties are checked: it is a no-op in equilibrium, it
reduces to the cell-dependent solve when the heat capacity is huge, it moves the material towards the radiation and
it accounts for the energy the radiation loses."""
import numpy as np
import pytest

from parity_util import make_oracle, synthetic_parameters

C_LIGHT = 299.792458


def _run(oracle, rt, p, steps, per_step=4):
    o = make_oracle(oracle, rt, p, num_threads=2)
    Ts, phis = [o.temperature()], [o.moments()[0]]
    for _ in range(steps):
        o.run_stages(per_step)
        Ts.append(o.temperature())
        phis.append(o.moments()[0])
    psi = o.psi()
    o.close()
    return np.array(Ts), np.array(phis), psi


def test_equilibrium_is_a_fixed_point(rt, oracle):
    """V = 0, psi = B_g(T) everywhere, source boundaries at B_g: absorption equals emission, T does not move"""
    p = synthetic_parameters(rt, M=4, G=6, N=64, dt=1e-3, V=0.0, steps=3, bc_left=1, bc_right=1)
    p.cv = 0.5
    Ts, _, _ = _run(oracle, rt, p, 3)
    assert np.max(np.abs(Ts - 1.0)) < 1e-11


def test_huge_heat_capacity_freezes_the_temperature(rt, oracle):
    """cv -> infinity: the update vanishes and the solve equals the cell-dependent one with a fixed T(c)"""
    p = synthetic_parameters(rt, M=4, G=4, N=128, dt=1e-3, V=1.0, steps=2)
    p.T_cell = 1.0 + 0.2 * np.sin(np.arange(p.N) / 9.0)
    _, _, frozen = _run(oracle, rt, p, 2)
    p.cv = 1e30
    Ts, _, psi = _run(oracle, rt, p, 2)
    assert np.max(np.abs(Ts[-1] - p.T_cell)) < 1e-14
    assert np.max(np.abs(psi - frozen) / np.abs(frozen)) < 1e-12


@pytest.mark.parametrize("ts", [3, 1, 2])
def test_leaky_wall_cools_the_slab_from_the_right(rt, oracle, ts):
    """reflective left wall, no inflow at the right one: radiation leaks out on the right, the material re-emits and
    cools -- monotonically in time, most at the right wall, with every time-stepping method"""
    p = synthetic_parameters(rt, M=4, G=6, N=96, dt=2e-4, V=0.0, steps=4, bc_left=2, bc_right=2, ts_method=ts)
    p.cv = 5.0
    Ts, _, _ = _run(oracle, rt, p, 4, 4 if ts == 3 else 1)
    assert np.all(Ts <= 1.0 + 1e-12)
    assert np.all(np.diff(Ts, axis=0)[:, -8:] < 0)
    assert np.argmin(Ts[-1]) >= p.N - 2
    assert abs(Ts[-1][0] - 1.0) < 1e-6  # c t = 0.24 << slab width 0.96: the left end has not noticed yet


def test_energy_exchange_matches_the_group_sums(rt, oracle):
    """the temperature change of a backward-Euler update is dt (A - E) / (rho cv + dt D) with the returned group sums"""
    p = synthetic_parameters(rt, M=4, G=5, N=80, dt=5e-4, V=1.0, steps=1, bc_left=1, bc_right=1)
    p.cv = 2.0
    p.T_cell = 1.0 + 0.3 * np.cos(np.arange(p.N) / 5.0)
    p.rho_cell = 1.0 + 0.5 * np.sin(np.arange(p.N) / 7.0)
    o = make_oracle(oracle, rt, p, num_threads=2)
    T0 = o.temperature()
    o.run_stages(4)
    T1, A, E, D = o.temperature(with_sums=True)
    phi = o.moments()[0]
    o.close()
    assert np.all(D > 0) and np.all(E > 0) and np.all(A > 0)
    want = T0 + p.dt * (A - E) / (p.rho_cell * p.cv + p.dt * D)
    assert np.allclose(T1, want, rtol=1e-14, atol=0)
    assert np.max(np.abs(T1 - T0)) > 1e-6
    # A is the absorbed power: sum_g rho kappa_g phi_g with the opacities the solve used
    kap = np.asarray(p.group_kappa)
    assert np.allclose(A, p.rho_cell * (kap[:, None] * phi).sum(axis=0), rtol=1e-12)


def test_bdf2_update_uses_the_previous_temperature(rt, oracle):
    """second and later BDF2 steps: T' = [rho cv (2T - T_prev/2) + dt (A - E + D T)] / (3/2 rho cv + dt D)"""
    p = synthetic_parameters(rt, M=4, G=4, N=48, dt=5e-4, V=0.0, steps=3, bc_left=1, bc_right=1)
    p.cv = 1.5
    p.T_cell = 1.0 + 0.25 * np.sin(np.arange(p.N) / 4.0)
    o = make_oracle(oracle, rt, p, num_threads=2)
    T0 = o.temperature()
    o.run_stages(4)
    T1 = o.temperature()
    o.run_stages(4)
    T2, A, E, D = o.temperature(with_sums=True)
    o.close()
    rcv = p.rho * p.cv
    want = (rcv * (2 * T1 - 0.5 * T0) + p.dt * (A - E + D * T1)) / (1.5 * rcv + p.dt * D)
    assert np.allclose(T2, want, rtol=1e-14, atol=0)


def test_relaxation_towards_radiation_temperature(rt, oracle):
    """cold material in a hot radiation field (source boundaries at T = 1): the material heats up step by step,
    monotonically, and never overshoots the radiation temperature"""
    p = synthetic_parameters(rt, M=4, G=6, N=64, dt=2e-4, V=0.0, steps=8, bc_left=1, bc_right=1)
    p.cv = 20.0
    o = make_oracle(oracle, rt, p, num_threads=2)
    o.close()
    # material at 0.8 but radiation initialised at B_g(0.8) too; the boundaries feed B_g(1.0)
    p.T_cell = np.full(p.N, 0.8)
    p.use_mg_equilib = False
    from parity_util import host_setup
    mu, wt, e, kap = host_setup(rt, p)
    B1, _ = oracle.planck(1.0, e)
    p.psi_source = np.repeat(B1[None, :], p.M, axis=0)
    Ts, _, _ = _run(oracle, rt, p, 8)
    mean = Ts.mean(axis=1)
    assert np.all(np.diff(mean) > 0)
    assert np.all(Ts < 1.0 + 1e-12) and np.all(Ts >= 0.8 - 1e-12)
