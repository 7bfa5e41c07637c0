"""GPU tests of the C++ drop-in: the reference's UNMODIFIED drivers (src/main.cc, tests/test_gray.cpp), its own
ParameterHandler / param / GLQuad, linked against radiative-transfer_b200/host/solver.cpp + librt_cuda.so
(binaries prebuilt by __graft_entry__.build() in the container that has the reference checkout).

`transfer_b200 <file.prm>` writes the same 8 CSV files as the reference's `transfer` (main.cc:117-133), at the default
ostream precision (6 significant digits); they are compared with the full-precision golden dumps of the reference at
that precision.  `gray_test_b200` is the reference's only CTest test (tests/CMakeLists.txt:7-9) and must pass."""
import os
import shutil
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, PRM_DIR, ROOT
from parity_util import load_golden

pytestmark = pytest.mark.gpu

BUILD = os.path.join(ROOT, "radiative-transfer_b200", "host", "build")
TRANSFER = os.path.join(BUILD, "transfer_b200")
GRAY = os.path.join(BUILD, "gray_test_b200")


def _need(path):
    if not os.path.exists(path):
        pytest.fail(f"{path} missing: __graft_entry__.build() makes it where the reference checkout is present")


def _csv(path):
    rows = [ln.split() for ln in open(path).read().strip().splitlines()]
    return np.array(rows, dtype=np.float64)


def _six_digits_equal(a, b, scale):
    """values printed with 6 significant digits: agree to half a unit in the 6th digit (plus 1e-9 of the array's scale)"""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return bool(np.all(np.abs(a - b) <= 1e-5 * np.abs(b) + 1e-9 * scale))


@pytest.mark.parametrize("name", ["single_group", "multi_group_equilibrium", "llnl_slab_test"])
def test_transfer_driver_writes_reference_csvs(name, tmp_path):
    _need(TRANSFER)
    # the reference looks its table files up as ../prm/<name> relative to the working directory
    run = tmp_path / "run"
    run.mkdir()
    shutil.copytree(PRM_DIR, tmp_path / "prm")
    out = subprocess.run([TRANSFER, str(tmp_path / "prm" / (name + ".prm"))], cwd=run, capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    g = load_golden(name)
    M, G, N = g["psi"].shape
    scale = float(np.max(np.abs(g["phi"])))
    phi = _csv(run / "phi.csv")
    assert phi.shape == (G, N) and _six_digits_equal(phi, g["phi"], scale)
    assert _six_digits_equal(_csv(run / "phi_plus.csv"), g["phi_plus"], scale)
    assert np.all(np.abs(_csv(run / "F.csv") - g["F"]) <= 1e-5 * np.abs(g["F"]) + 2e-9 * scale)
    psi = _csv(run / "psi.csv")  # tensor printed as M rows x (G*N) columns
    assert psi.shape == (M, G * N)
    assert _six_digits_equal(psi, g["psi"].reshape(M, G * N, order="F"), float(np.max(np.abs(g["psi"]))))
    assert _six_digits_equal(_csv(run / "e_ave.csv").ravel(), g["e_ave"], 1.0)
    assert _six_digits_equal(_csv(run / "left_ends.csv").ravel(), g["left_ends"], float(np.max(np.abs(g["left_ends"]))))
    assert _six_digits_equal(_csv(run / "right_ends.csv").ravel(), g["right_ends"], float(np.max(np.abs(g["right_ends"]))))
    x = _csv(run / "x.csv").ravel()
    assert x.size == N


def test_gray_test_passes(tmp_path):
    """tests/test_gray.cpp:89 -- the reference's own pass criterion, through the C++ drop-in"""
    _need(GRAY)
    out = subprocess.run([GRAY], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "Gray test passed." in out.stdout
    F = _csv(tmp_path / "gray-test-F.csv")
    assert abs(F.max()) < 1e-6


@pytest.mark.parametrize("name", ["multi_group_equilibrium", "llnl_slab_test"])
def test_python_mirror_writes_the_same_files_as_the_cpp_driver(rt, name, tmp_path):
    """Solver.write_outputs (Python mirror of main.cc:117-133) against the files transfer_b200 writes: same 8 files, same
    numbers as printed (6 significant digits)"""
    _need(TRANSFER)
    from parity_util import load_case
    run = tmp_path / "run"
    run.mkdir()
    shutil.copytree(PRM_DIR, tmp_path / "prm")
    out = subprocess.run([TRANSFER, str(tmp_path / "prm" / (name + ".prm"))], cwd=run, capture_output=True, text=True,
                         timeout=600)
    assert out.returncode == 0
    py = tmp_path / "py"
    py.mkdir()
    p = load_case(rt, name)
    with rt.Solver(p) as s:
        s.solve()
        s.write_outputs(str(py))
    names = ["phi", "phi_plus", "psi", "x", "F", "e_ave", "left_ends", "right_ends"]
    for n in names:
        a, b = _csv(run / (n + ".csv")), _csv(py / (n + ".csv"))
        assert a.shape == b.shape, n
        scale = float(np.max(np.abs(a))) or 1.0
        # both print the SAME device results with 6 digits; F is a difference of O(phi) terms whose last bits may differ
        tol = 2e-6 * scale if n != "F" else 1e-9 * float(np.max(np.abs(_csv(run / "phi.csv"))))
        assert np.all(np.abs(a - b) <= tol + 1e-5 * np.abs(a)), n
