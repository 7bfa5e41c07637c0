"""Correction::Print() (reference src/correction.cpp:404-603): the report formatted by the library from the
DEVICE's per-group tables against the text the UNMODIFIED reference prints (tests/golden/correction_print_*.txt, made by
tests/golden/make_golden_print.py from oracle/_ref/ref_dump --print).

The reference prints 6 significant digits.  Layout (every title, rule, column start, blank line) must be identical;
numbers are compared as numbers: equal to the printed precision (half a unit of the 6th digit, plus one unit for the
rounding of a value that sits on a boundary), except the one line that prints a cancellation -- Sum dkapEB / Sum |dkapEB|,
O(1e-17) in the reference itself -- which must be < 1e-12 in both."""
import os
import re
import shutil
import subprocess

import pytest

from conftest import GOLDEN, PRM_DIR, ROOT
from parity_util import load_case

CASES = ("single_group", "multi_group_equilibrium", "llnl_slab_test")
NUMBER = re.compile(r"^[-+]?(\d+\.?\d*|\.\d+)([eE][-+]?\d+)?$")


def _golden(case):
    with open(os.path.join(GOLDEN, f"correction_print_{case}.txt")) as f:
        return f.read()


def _tokens(line):
    return [(m.start(), m.group()) for m in re.finditer(r"\S+", line)]


def assert_same_report(got: str, want: str):
    gl, wl = got.split("\n"), want.split("\n")
    assert len(gl) == len(wl), (len(gl), len(wl))
    n_numbers = 0
    for k, (g, w) in enumerate(zip(gl, wl)):
        if w.startswith("Sum dkapEB/Sum |dkapEB| = "):
            assert g.startswith("Sum dkapEB/Sum |dkapEB| = ")
            x, y = float(g.split("=")[1]), float(w.split("=")[1])
            assert x == y or (abs(x) < 1e-12 and abs(y) < 1e-12), (x, y)  # G = 1: one term, the ratio is -1 or 1 in both
            continue
        tg, tw = _tokens(g), _tokens(w)
        assert [p for p, _ in tg] == [p for p, _ in tw], f"line {k + 1}: columns differ\n got: {g!r}\nwant: {w!r}"
        for (_, a), (_, b) in zip(tg, tw):
            if a == b:
                continue
            assert NUMBER.match(a) and NUMBER.match(b), f"line {k + 1}: {a!r} != {b!r}"
            x, y = float(a), float(b)
            assert abs(x - y) <= 1.5e-6 * abs(y) + 1e-300, f"line {k + 1}: {a} vs {b}"
            n_numbers += 1
        assert len(g) == len(w) or any(a != b for (_, a), (_, b) in zip(tg, tw)), f"line {k + 1}: padding differs"
    return n_numbers


def test_the_comparison_itself_catches_differences():
    want = _golden("multi_group_equilibrium")
    assert assert_same_report(want, want) == 0
    with pytest.raises(AssertionError):
        assert_same_report(want.replace("Group Width", "Group  Width", 1), want)  # a title
    with pytest.raises(AssertionError):
        assert_same_report(want.replace("\n\n\n", "\n\n", 1), want)  # a blank line
    with pytest.raises(AssertionError):
        assert_same_report(want.replace("kappa_grey: 1", "kappa_grey: 2", 1), want)  # a number
    row = [l for l in want.split("\n") if l.startswith("1 ")][0]
    num = row.split()[1]
    close = f"{float(num) * (1 + 5e-7):.7g}"  # within the printed precision: accepted, and counted
    assert assert_same_report(want.replace(row, row.replace(num, close[:len(num)].ljust(len(num)), 1), 1), want) <= 1


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_report_matches_the_reference_print(rt, case):
    p = load_case(rt, case)
    with rt.Solver(p) as s:
        got = s.correction_report()
        flips = assert_same_report(got, _golden(case))
        print(f"{case}: {len(got.splitlines())} lines, {flips} numbers differ in the last printed digit")
    if p.G >= 2:
        with rt.Solver(p, devices=[0, 0]) as s:  # the multi-device handle answers with the same text
            assert s.correction_report() == got


@pytest.mark.gpu
def test_buffer_contract(rt):
    import ctypes as C
    p = load_case(rt, "multi_group_equilibrium")
    with rt.Solver(p) as s:
        text = s.correction_report()
        lib, n = rt.capi.lib(), C.c_size_t()
        small = C.create_string_buffer(100)
        assert s._f("correction_print")(s._h, small, 100, C.byref(n)) != 0  # too small: an error, the length still reported
        assert n.value == len(text) and b"too small" in lib.rt_last_error()
        assert small.value.decode() == text[:99]  # truncated and NUL-terminated, never overrun


@pytest.mark.gpu
def test_reference_driver_on_the_dropin_prints_it(tmp_path):
    """RT_CUDA_PRINT_CORRECTION=1 + the reference's unmodified main.cc on the drop-in: the report appears in its stdout"""
    exe = os.path.join(ROOT, "radiative-transfer_b200", "host", "build", "transfer_b200")
    if not os.path.exists(exe):
        pytest.fail(f"{exe} missing: __graft_entry__.build() makes it where the reference checkout is present")
    shutil.copytree(PRM_DIR, tmp_path / "prm")
    run = tmp_path / "run"
    run.mkdir()
    env = dict(os.environ, RT_CUDA_PRINT_CORRECTION="1")
    out = subprocess.run([exe, str(tmp_path / "prm" / "llnl_slab_test.prm")], cwd=run, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    want = _golden("llnl_slab_test")
    a = out.stdout.index("========== Correction::Print() ==========")
    got = out.stdout[a:a + len(want) + 4096]
    got = "\n".join(got.split("\n")[: len(want.split("\n")) - 1]) + "\n"
    assert_same_report(got, want)
