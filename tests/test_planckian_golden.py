"""The reference's own stand-alone test program tests/test_planckian.cpp (compiled unmodified, driven through stdin by
tests/golden/make_golden_planckian.py) as a second set of golden vectors for the G-length pieces of the path: Planck
group integrals, multigroup opacities (correction.cpp:66-97), group-edge opacities, dEB, dkapEdE, dkapEB, cor1..3.
The program prints 6 significant digits, so the comparison is at that precision.

CPU part: the oracle (oracle-2) against those tables.  GPU part: the device kernels through the C ABI."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN

CASES = json.load(open(os.path.join(GOLDEN, "planckian.json")))


def six_digits(a, b):
    """b was printed with 6 significant digits: |a-b| <= 1 unit in the 6th digit of b"""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return bool(np.all(np.abs(a - b) <= 1.0e-5 * np.abs(b) + 1e-300))


def _derived(B, kappa, gt):
    """cor1..3 as test_planckian.cpp:533-538 / correction.cpp:328-340 define them (rho = 1)"""
    cor1 = gt["dsigEdE"]
    cor2 = 3.0 * kappa * B - gt["dkapEB"]
    cor3 = cor1 * (4.0 * B - gt["dEB"])
    return cor1, cor2, cor3


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_against_test_planckian(rt, oracle, name):
    c, t = CASES[name]["inputs"], CASES[name]["tables"]
    G = c["G"]
    e = rt.grid.group_edges(G, c["efirst"], c["elast"])
    assert six_digits(0.5 * (e[:-1] + e[1:]), t["e_ave"])
    B, dB = oracle.planck(c["T"], e)
    assert six_digits(B, t["B"]) and six_digits(dB, t["dBdT"])
    k6 = oracle.multigroup_opacities(c["T"], c["kappa_grey"], e, B)
    for k in ("ukappa", "ckappa", "kappa", "emis_spec"):
        assert six_digits(k6[k], t[k]), k
    ac = rt.grid.RADIATION_CONSTANT_A * rt.grid.SPEED_OF_LIGHT
    assert abs(np.sum(k6["kappa"] * B) - c["kappa_grey"] * ac * c["T"] ** 4) < 1e-9  # test_planckian.cpp:338-349
    mu, wt = rt.grid.gauss_legendre(c["M"])
    o = oracle.Oracle(M=c["M"], N=c["N"], G=G, ts_method=3, max_timesteps=1, bc_left=2, bc_right=1, use_correction=True,
                      use_mg_equilib=True, dx=c["X"] / c["N"], dt=1e-5, rho=1.0, T=c["T"], V=0.0, kappa_grey=c["kappa_grey"],
                      mu=mu, wt=wt, e_edge=e, kappa=k6["kappa"])
    o.run_stages(1)  # like the reference, the edge stencils are first evaluated by compute_correction (solver.cpp:608)
    gt = o.group_terms()
    ok, sums = o.validate()
    assert ok  # with these opacities all three identities of correction.cpp:39-63,100-122 hold
    o.close()
    assert six_digits(gt["dEB"], t["dEB"]) and six_digits(gt["kappa_edge"], t["kappa_edge"])
    assert six_digits(gt["dsigEdE"], t["dkapEdE"]) and six_digits(gt["dkapEB"], t["dkapEB"])
    cor1, cor2, cor3 = _derived(B, k6["kappa"], gt)
    assert six_digits(cor1, t["cor1"]) and six_digits(cor2, t["cor2"]) and six_digits(cor3, t["cor3"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_device_against_test_planckian(rt, oracle, name):
    c, t = CASES[name]["inputs"], CASES[name]["tables"]
    G = c["G"]
    e = rt.grid.group_edges(G, c["efirst"], c["elast"])
    k6 = rt.multigroup_opacities(e, c["T"], c["kappa_grey"])
    for k in ("ukappa", "ckappa", "kappa", "emis_spec"):
        assert six_digits(k6[k], t[k]), k
    B, _ = oracle.planck(c["T"], e)
    ref = oracle.multigroup_opacities(c["T"], c["kappa_grey"], e, B)
    for k in ("ukappa", "ckappa", "kappa", "emis_spec"):
        # the last group's B is a cancellation remainder (Planck.cpp:58-76): everything divided by it inherits that
        rel = np.abs(k6[k] - ref[k]) / np.abs(ref[k])
        assert np.all(rel[:-1] <= 1e-10) and rel[-1] <= 1e-6, (k, rel.max())
    p = rt.Parameters(M=c["M"], G=G, N=c["N"], X=c["X"], efirst=c["efirst"], elast=c["elast"], use_mg_equilib=True, V=0.0,
                      use_correction=True, ts_method=3, dt=1e-5, max_timesteps=1, include_validation=True, rho=1.0,
                      T=c["T"], kappa_grey=c["kappa_grey"])
    p.have_group_absorption_opacities = True
    p.group_kappa = k6["kappa"]
    with rt.Solver(p) as s:
        ok, _ = s.validate()
        assert ok
        gt = s.group_terms()
        s.solve()  # include_validation is on: rt_solve would fail if the identities did not hold
    assert six_digits(gt["B"], t["B"]) and six_digits(gt["dBdT"], t["dBdT"])
    assert six_digits(gt["dEB"], t["dEB"]) and six_digits(gt["kappa_edge"], t["kappa_edge"])
    assert six_digits(gt["dsigEdE"], t["dkapEdE"]) and six_digits(gt["dkapEB"], t["dkapEB"])
    cor1, cor2, cor3 = _derived(gt["B"], k6["kappa"], gt)
    assert six_digits(cor1, t["cor1"]) and six_digits(cor2, t["cor2"]) and six_digits(cor3, t["cor3"])
