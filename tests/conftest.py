import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
ORACLE_DIR = os.path.join(ROOT, "oracle")
if ORACLE_DIR not in sys.path:
    sys.path.insert(0, ORACLE_DIR)

GOLDEN = os.path.join(ROOT, "tests", "golden")
PRM_DIR = os.path.join(GOLDEN, "prm")

GOLDEN_CASES = [
    "single_group", "multi_group_equilibrium", "llnl_slab_test", "be_reflect_source", "cn_vacuum_vacuum",
    "bdf2_reflect_equilib_s8", "bdf2_tables_ragged", "bdf2_nocorr_vacuum_left", "bdf2_one_cell",
    "bdf2_right_reflective_s4",
]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def rt():
    """the product package (directory name has a hyphen)"""
    return importlib.import_module("radiative-transfer_b200")


@pytest.fixture(scope="session")
def oracle():
    """TEST INFRASTRUCTURE: the CPU oracle binding (oracle/oracle_py.py)"""
    import oracle_py
    oracle_py.build()
    return oracle_py


def host_setup(rt, p):
    """mu, wt, e_edge, kappa the way the Solver constructor derives them (solver.cpp:76-78,96-100,145-152)"""
    import numpy as np
    mu, wt = rt.grid.gauss_legendre(p.M)
    e = p.group_bounds if p.have_group_bounds else rt.grid.group_edges(p.G, p.efirst, p.elast)
    kap = p.group_kappa if p.have_group_absorption_opacities else np.full(p.G, p.kappa_grey)
    return mu, wt, np.asarray(e, dtype=np.float64), np.asarray(kap, dtype=np.float64)
