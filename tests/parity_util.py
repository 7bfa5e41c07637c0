"""Helpers shared by the parity tests (TEST INFRASTRUCTURE)."""
import os

import numpy as np

from conftest import GOLDEN, PRM_DIR, host_setup

#: north_star tolerance: 1e-9 relative on scalar intensity / radiation energy at each BDF2 step
RTOL = 1e-9

#: absolute floor of the state comparison for CELL-DEPENDENT temperature fields, as a fraction of the group's largest entry.
#: The reference-pinned uniform cases use 1e-6 (tests/parity_util.py).  With T(c) a high-energy group's intensity varies
#: by orders of magnitude from cell to cell, and the last group is the reference's own cancellation remainder
#: a c T^4 - sum of the others (Planck.cpp:58-76): its cold-cell entries carry a relative uncertainty of eps * sum(B) / B_last(T_cold)
#: that the uniform-T allowance does not describe.  These extensions are parity-unpinned (no reference counterpart).
CELL_FLOOR = 1e-3


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


def load_case(rt, name):
    cwd = os.getcwd()
    os.chdir(GOLDEN)  # table files are looked up as ../prm/<name> like the reference does
    try:
        return rt.prm.load(os.path.join(PRM_DIR, name + ".prm"))
    finally:
        os.chdir(cwd)


def rel_err(a, b):
    """max elementwise |a-b|/|b| ; entries where both are exactly equal (incl. 0 == 0) count as 0"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = np.abs(a - b)
    den = np.abs(b)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(d == 0.0, 0.0, d / den)
    return float(np.max(r)) if r.size else 0.0


def state_err(a, b, group_axis=1, floor=1e-6):
    """max |a-b| / (|b| + floor*gmax), gmax = max |b| over the entry's energy group.

    <= RTOL means: 1e-9 relative elementwise, with an absolute floor of 1e-15 x the group's scale for entries that
    are more than a million times smaller than the group's largest one (zero crossings of the oscillating DG solution,
    where an elementwise relative error is meaningless for reference and reformulation alike).  Round 1 used a floor of
    1e-3; the errors observed on B200 under this stricter measure are <= 0.08 of the tolerance on every golden and
    synthetic case (tools/gpu_tolerance_audit.py), so a regression of one order of magnitude now fails."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if b.size == 0:
        return 0.0
    axes = tuple(k for k in range(b.ndim) if k != group_axis)
    gmax = np.max(np.abs(b), axis=axes, keepdims=True)
    d = np.abs(a - b)
    den = np.abs(b) + floor * gmax
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(d == 0.0, 0.0, d / den)
    return float(np.max(r))


def group_rtol(B):
    """Per-group relative tolerance: RTOL everywhere, except that the LAST group's Planck integral is defined by the
    reference as a remainder, B_{G-1} = a c T^4 - sum_{g<G-1} B_g (Planck.cpp:58-76).  That cancellation makes the
    reference's own value uncertain by ~eps * sum(B) / B_{G-1} relative (every B_g carries a few ulp of libm rounding),
    and everything proportional to B_{G-1} (psi, phi of that group) inherits it.  The allowance is 32 eps * sum(B)/B_last
    (llnl_slab_test: sum B / B_last = 1.5e7)."""
    B = np.asarray(B, dtype=np.float64)
    r = np.full(B.size, RTOL)
    if B.size > 1 and B[-1] > 0:
        r[-1] = RTOL + 32 * np.finfo(np.float64).eps * float(np.sum(np.abs(B))) / float(B[-1])
    return r


def state_ok(a, b, B, group_axis=1, floor=1e-6):
    """(ok, worst ratio err/tol, per-group errors) for a state array with an energy-group axis; see state_err, group_rtol"""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    axes = tuple(k for k in range(b.ndim) if k != group_axis)
    gmax = np.max(np.abs(b), axis=axes, keepdims=True)
    d = np.abs(a - b)
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(d == 0.0, 0.0, d / (np.abs(b) + floor * gmax))
    per_group = np.max(r, axis=axes)
    ratio = per_group / group_rtol(B)
    return bool(np.all(ratio <= 1.0)), float(np.max(ratio)), per_group


def flux_err(F, F_ref, phi_ref):
    """max |F - F_ref| against max |F_ref| -- the flux's OWN scale -- wherever the flux is a meaningful quantity
    (max |F| > 1e-6 max phi); in equilibrium cases F is a cancellation residue of O(eps phi) and the phi scale is the
    only one there is (GrayTest's criterion, tests/test_gray.cpp:89, is of that kind).  <= RTOL required."""
    mF, mphi = float(np.max(np.abs(F_ref))), float(np.max(np.abs(phi_ref)))
    scale = mF if mF > 1e-6 * mphi else mphi
    return scaled_err(F, F_ref, scale)


def scaled_err(a, b, scale):
    """max |a-b| / scale -- for quantities that are differences of large terms (flux, balance)"""
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b)))) / float(scale) if np.size(a) else 0.0


def make_oracle(oracle, rt, p, num_threads=1):
    mu, wt, e, kap = host_setup(rt, p)
    return oracle.oracle_from_parameters(p, mu, wt, e, kap, num_threads=num_threads)


def synthetic_parameters(rt, *, M, G, N, dt, V=1.0, steps=2, bc_left=2, bc_right=1, ts_method=3, dx=0.01, seed=20261104,
                         use_correction=True, use_mg_equilib=True, efirst=0.01, elast=30.0):
    """The synthetic slab of SURVEY 8(d): uniform slab dx=0.01, rho=T=1, log-spaced groups efirst=0.01..elast=30,
    kappa_g = 10^U(-1,2) from PCG64(seed), correction live (V != 0), equilibrium boundary sources."""
    rng = np.random.Generator(np.random.PCG64(seed))
    kap = 10.0 ** rng.uniform(-1.0, 2.0, size=G)
    p = rt.Parameters(M=M, G=G, N=N, X=N * dx, efirst=efirst, elast=elast, bc_left_indicator=bc_left,
                      bc_right_indicator=bc_right, use_mg_equilib=use_mg_equilib, V=V, use_correction=use_correction,
                      ts_method=ts_method, dt=dt, max_timesteps=steps, include_validation=False, rho=1.0, T=1.0)
    p.have_group_absorption_opacities = True
    p.group_kappa = kap
    return p
