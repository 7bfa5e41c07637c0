"""CPU tests of the host-side logic: .prm parsing, angular quadrature, energy grid, group sharding / batching."""
import os

import numpy as np
import pytest

from conftest import GOLDEN_CASES, PRM_DIR, host_setup
from parity_util import load_case, load_golden


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_host_inputs_bit_identical_to_reference(rt, name):
    """mu, wt (GLQuad.cpp:4-44), e_edge / e_ave (solver.cpp:6-32) and kappa as the reference's Solver ctor derives them"""
    p = load_case(rt, name)
    g = load_golden(name)
    mu, wt, e, kap = host_setup(rt, p)
    assert np.array_equal(mu, g["mu"]) and np.array_equal(wt, g["wt"])
    assert np.array_equal(e, g["e_edge"]) and np.array_equal(kap, g["kappa"])
    e_ave, de = rt.grid.group_averages(e)
    assert np.array_equal(e_ave, g["e_ave"])
    assert np.all(mu[: p.M // 2] < 0) and np.all(mu[p.M // 2:] > 0) and np.all(np.diff(mu) > 0)
    assert abs(wt.sum() - rt.grid.FOUR_PI) < 1e-9  # Newton tolerance is 1e-12 on the nodes (GLQuad.h:13)


def test_grid_matches_oracle_restatement(rt, oracle):
    for M in (2, 4, 8, 16, 32, 64):
        mu, wt = rt.grid.gauss_legendre(M)
        omu, owt = oracle.glquad(M, rt.grid.FOUR_PI)
        assert np.array_equal(mu, omu) and np.array_equal(wt, owt)
    for G in (1, 2, 5, 124, 256, 4096):
        assert np.array_equal(rt.grid.group_edges(G, 0.01, 30.0), oracle.group_edges(G, 0.01, 30.0))


def test_single_group_edges(rt):
    """G = 1: the reference's `assert(logfac = 1.)` is an assignment; edges are [0, efirst]"""
    assert np.array_equal(rt.grid.group_edges(1, 0.1, 10.0), np.array([0.0, 0.1]))


def test_prm_reference_configs(rt):
    p = load_case(rt, "single_group")
    assert (p.M, p.G, p.N, p.ts_method, p.max_timesteps) == (2, 1, 100, 3, 1000)
    assert p.bc_left_indicator == 2 and p.bc_right_indicator == 1
    p = load_case(rt, "multi_group_equilibrium")
    assert (p.M, p.G, p.N, p.max_timesteps) == (2, 4, 100, 500)
    p = load_case(rt, "llnl_slab_test")
    assert (p.M, p.G, p.N, p.max_timesteps) == (2, 124, 50, 2)
    assert p.have_group_bounds and p.have_group_absorption_opacities and not p.include_validation
    assert p.group_bounds.size == 125 and p.group_kappa.size == 124
    assert p.bc_left_indicator == 0 and p.bc_right_indicator == 0


def test_prm_parsing_rules(rt, tmp_path):
    """param.h:61-75 / param.cpp:5-66: '#' lines skipped, first key wins, numeric prefix conversion, exact-string bools"""
    f = tmp_path / "t.prm"
    f.write_text("# comment\nM=4   # trailing text\nM=8\nG=2\nX=2.5e0 cm\nN=10\nuse_correction=true \nuse_mg_equilib=True\n"
                 "psi_source=1 2 3 4 5 6 7 8\nV=.5\nnot a key value line\n")
    p = rt.prm.load(str(f))
    assert p.M == 4 and p.G == 2 and p.X == 2.5 and p.N == 10 and p.dx == 0.25
    assert p.use_correction is False  # "true " (trailing blank) is not one of the accepted strings
    assert p.use_mg_equilib is True
    assert np.all(p.psi_source == 0)  # sources are not read when use_mg_equilib (ParameterHandler.cpp:120)
    assert p.V == 0.5 and p.ts_method == 3 and p.dt == 1e-5 and p.max_timesteps == 1000 and p.include_validation


def test_prm_psi_source_row_major_fill(rt, tmp_path):
    f = tmp_path / "s.prm"
    f.write_text("M=2\nG=3\npsi_source=1 2 3 4 5 6\n")
    p = rt.prm.load(str(f))
    assert np.array_equal(p.psi_source, np.array([[1.0, 2.0, 3.0], [4.0, 5.0, 6.0]]))


def test_prm_missing_file_gives_defaults(rt, tmp_path, capsys):
    """ParameterHandler.cpp:13-15 / param.h:56-59: a missing file only prints; every parameter takes its default"""
    p = rt.prm.load(str(tmp_path / "nope.prm"))
    assert (p.M, p.G, p.N, p.ts_method) == (2, 1, 100, 3)


def test_prm_dump_round_trip(rt, tmp_path):
    p = load_case(rt, "bdf2_reflect_equilib_s8")
    out = tmp_path / "rt.prm"
    rt.prm.dump(p, str(out))
    q = rt.prm.load(str(out))
    for k in ("M", "G", "N", "X", "efirst", "elast", "bc_left_indicator", "bc_right_indicator", "use_mg_equilib", "rho",
              "kappa_grey", "T", "V", "use_correction", "ts_method", "dt", "max_timesteps", "include_validation"):
        assert getattr(p, k) == getattr(q, k), k


def test_missing_table_is_loud(rt, tmp_path):
    f = tmp_path / "t.prm"
    f.write_text("G=3\nhave_group_bounds=true\nfilename_group_bounds=does_not_exist.txt\n")
    with pytest.raises(FileNotFoundError):
        rt.prm.load(str(f))


def test_group_ranges_cover_and_balance(rt):
    import importlib
    sh = importlib.import_module("radiative-transfer_b200.sharding")
    for G, W in ((4096, 8), (4096, 1), (256, 4), (7, 3), (9, 2), (8, 8)):
        r = sh.all_ranges(G, W)
        assert r[0][0] == 0 and sum(n for _, n in r) == G
        assert all(r[k][0] + r[k][1] == r[k + 1][0] for k in range(W - 1))
        assert max(n for _, n in r) - min(n for _, n in r) <= 1
    with pytest.raises(ValueError):
        sh.group_range(3, 4, 0)
    assert sh.batches(7, 3) == [(0, 3), (3, 3), (6, 1)]
    assert sh.batches(4, 8) == [(0, 4)]
