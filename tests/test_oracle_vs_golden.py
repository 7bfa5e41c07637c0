"""CPU tests: the oracle (TEST INFRASTRUCTURE) against the golden vectors dumped by the UNMODIFIED reference.

tests/golden/*.npz were produced by tests/golden/make_golden.py from oracle/_ref/ref_dump (the reference's own sources
compiled in place).  oracle-2 (oracle/rt_oracle.cpp, the CPU restatement the GPU parity tests and the bench's
cpu_baseline use) must reproduce them BIT FOR BIT -- that is what pins it.  Also checked: the reference's own
known-answer tests for this path (tests/test_gray.cpp:89, correction.cpp:53-54,113) and the 6-digit anchors of
BASELINE.md.
"""
import os

import numpy as np
import pytest

from conftest import GOLDEN_CASES, PRM_DIR
from parity_util import load_case, load_golden, make_oracle

STATE_KEYS = ("psi", "ends", "prev_ends", "half_ends")


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle2_reproduces_reference_bitwise(rt, oracle, name):
    p = load_case(rt, name)
    g = load_golden(name)
    o = make_oracle(oracle, rt, p)
    assert o.total_stages == p.max_timesteps * (4 if p.ts_method == 3 else 1)
    assert np.array_equal(o.psi(), np.broadcast_to(g["B_init"][None, :, None], g["psi"].shape))
    o.solve()
    assert o.stages_done == o.total_stages
    assert np.array_equal(o.psi(), g["psi"])
    for which, key in enumerate(("ends", "prev_ends", "half_ends")):
        assert np.array_equal(o.ends(which), g[key]), key
    phi, pp, F = o.moments()
    assert np.array_equal(phi, g["phi"]) and np.array_equal(pp, g["phi_plus"]) and np.array_equal(F, g["F"])
    assert np.array_equal(o.balance(), g["balance"])
    l, r = o.group_ends()
    assert np.array_equal(l, g["left_ends"]) and np.array_equal(r, g["right_ends"])
    gt = o.group_terms()
    for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB", "kappa_edge"):
        assert np.array_equal(gt[k], g[k]), k
    assert np.array_equal(o.psi_source(), g["psi_source"])
    o.close()


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_oracle2_step_snapshots_bitwise(rt, oracle, name):
    """per-step parity: reference runs with max_timesteps = k against oracle-2 stopped after k steps"""
    p = load_case(rt, name)
    g = load_golden(name)
    per_step = 4 if p.ts_method == 3 else 1
    ks = sorted(int(k[len("psi_steps"):]) for k in g if k.startswith("psi_steps"))
    assert ks
    o = make_oracle(oracle, rt, p)
    done = 0
    for k in ks:
        o.run_stages((k - done) * per_step)
        done = k
        assert np.array_equal(o.psi(), g[f"psi_steps{k}"]), f"{name}: psi after {k} steps"
        assert np.array_equal(o.ends(0), g[f"ends_steps{k}"]), f"{name}: ends after {k} steps"
    o.close()


def test_gray_test_invariant_on_golden():
    """tests/test_gray.cpp:89 -- abs(F.maxCoeff()) < 1e-6 on single_group.prm (max F = -2.84307e-08 in the survey run)"""
    g = load_golden("single_group")
    assert abs(float(np.max(g["F"]))) < 1e-6
    assert f"{float(np.max(g['F'])):.6g}" == "-2.84307e-08"


def test_baseline_anchors():
    """6-significant-digit anchors of BASELINE.md (from the reference's own CSV writers)"""
    s = lambda v: f"{float(v):.6g}"
    c1 = load_golden("single_group")
    assert all(s(v) == "51.4354" for v in c1["phi"].ravel())
    assert all(s(v) == "4.0931" for v in c1["psi"].ravel())
    assert s(c1["left_ends"][0]) == "0.136531" and s(c1["right_ends"][0]) == "0.136531"
    c2 = load_golden("multi_group_equilibrium")
    assert s(c2["phi"][0, 0]) == "0.00254249" and s(c2["phi"][3, 0]) == "40.6125" and s(c2["phi"][3, 99]) == "40.4892"
    assert s(c2["F"].min()) == "1.09661e-05" and s(c2["F"].max()) == "0.14639"
    c3 = load_golden("llnl_slab_test")
    assert s(c3["phi"][0, 0]) == "2.2596e-09" and s(c3["phi"][123, 0]) == "8.12797e-06"
    assert s(c3["F"].min()) == "-4.67685" and s(c3["F"].max()) == "6.63603"
    assert s(c3["e_ave"][0]) == "0.0011145"


@pytest.mark.parametrize("name", ["single_group", "multi_group_equilibrium"])
def test_validation_identities(rt, oracle, name):
    """correction.cpp:39-63,100-122: sum B = a c T^4, sum dB/dT = 4 a c T^3, sum kappa_g B_g = kappa_grey a c T^4 (1e-6)"""
    p = load_case(rt, name)
    o = make_oracle(oracle, rt, p)
    ok, sums = o.validate()
    ac = rt.grid.RADIATION_CONSTANT_A * rt.grid.SPEED_OF_LIGHT
    assert ok
    assert abs(sums[0] - ac * p.T ** 4) < 1e-6 and abs(sums[1] - 4 * ac * p.T ** 3) < 1e-6
    assert abs(sums[2] - p.kappa_grey * ac * p.T ** 4) < 1e-6
    o.close()


def test_single_group_B_is_acT4(rt, oracle):
    """prm/single_group.prm:26 pins B_grey = a c T^4 = 4.0931"""
    p = load_case(rt, "single_group")
    o = make_oracle(oracle, rt, p)
    assert f"{o.group_terms()['B'][0]:.5g}" == "4.0931"
    o.close()


def test_oracle2_threads_do_not_change_results(rt, oracle):
    """the OpenMP-over-groups mode used for the all-host-cores CPU baseline is bit-identical to 1 thread"""
    p = load_case(rt, "bdf2_tables_ragged")
    a = make_oracle(oracle, rt, p, num_threads=1)
    b = make_oracle(oracle, rt, p, num_threads=4)
    a.solve()
    b.solve()
    assert np.array_equal(a.psi(), b.psi()) and np.array_equal(a.ends(0), b.ends(0))
    a.close()
    b.close()


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference checkout (build container only)")
@pytest.mark.parametrize("name", ["llnl_slab_test", "be_reflect_source", "bdf2_one_cell"])
def test_live_reference_matches_golden(oracle, name):
    """regenerate a golden case from oracle/_ref/ref_dump right now: the committed fixtures are what the reference gives"""
    if not oracle.reference_available():
        pytest.skip("oracle/_ref not built")
    d = oracle.run_reference(os.path.join(PRM_DIR, name + ".prm"), timeout=300)
    g = load_golden(name)
    for k in ("psi", "ends", "phi", "F"):
        assert np.array_equal(d[k], g[k]), k
