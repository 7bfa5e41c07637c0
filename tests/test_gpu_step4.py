"""GPU parity of the one-scan fused BDF2-step kernel (csrc/step4_kernel.cu, default) through the C ABI.

One 4-vector scan per sub-tile instead of four scalar scans (tools/step4_form_check.py is the numpy statement of the
formulation).  Covered here: both fused kernels == oracle, chained-warp shapes, rows that end anywhere in the CTA tile,
CTAs that walk hundreds of tiles with idle warps (the back-pressure bug of round 1, ADVICE), the per-step scalar intensity
the kernel accumulates itself (PHI) against the separate moments pass for every run length, full-length rows in C5's
shape / the local regime dt = 1e-5 / V = 0 (reference src/solver.cpp:706-809, 191-204)."""
import ctypes as C

import numpy as np
import pytest

from parity_util import make_oracle, state_err, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _one_scan_kernel(monkeypatch):
    """the library picks the kernel by job (four-scan for plain steps, one-scan when the step sums phi): these tests are
    about the one-scan kernel, so it is forced; the tests that compare both set the variable themselves"""
    monkeypatch.setenv("RT_CUDA_STEP_KERNEL", "4")


def _oracle_state(oracle, rt, p, stages, threads=2):
    o = make_oracle(oracle, rt, p, num_threads=threads)
    o.run_stages(stages)
    out = dict(B=o.group_terms()["B"], psi=o.psi(), ends=o.ends(0), phi=o.moments()[0])
    o.close()
    return out


def _check(s, ref):
    ok1, w1, _ = state_ok(s.psi_mat_ref, ref["psi"], ref["B"])
    ok2, w2, _ = state_ok(s.ends(0), ref["ends"], ref["B"])
    return ok1 and ok2, max(w1, w2)


@pytest.mark.parametrize("kernel", [4, 1])
@pytest.mark.parametrize("bcl,bcr,V", [(2, 1, 1.0), (1, 1, 1.0), (0, 0, 1.0), (2, 1, 0.0), (1, 2, 0.5)])
def test_both_fused_kernels_match_the_oracle(rt, oracle, monkeypatch, kernel, bcl, bcr, V):
    monkeypatch.setenv("RT_CUDA_STEP_KERNEL", str(kernel))
    p = synthetic_parameters(rt, M=8, G=5, N=5000, dt=1e-3, V=V, steps=3, bc_left=bcl, bc_right=bcr)
    ref = _oracle_state(oracle, rt, p, 12)
    with rt.Solver(p) as s:
        assert s.plan().step_kernel == kernel
        s.solve()
        assert s.timing().step_launches == 3 and s.plan().time_levels == 2
        ok, worst = _check(s, ref)
    assert ok, f"kernel {kernel} bc {bcl}{bcr} V {V}: err/tol {worst:.3e}"


@pytest.mark.parametrize("dt", [1e-5, 1e-2])
def test_local_and_long_range_regimes(rt, oracle, dt):
    """alpha ~ 0.01 (dt = 1e-5: the transfer matrices' powers underflow towards zero) and alpha -> 1 (dt = 1e-2)"""
    p = synthetic_parameters(rt, M=4, G=6, N=4099, dt=dt, V=1.0, steps=2)
    ref = _oracle_state(oracle, rt, p, 8)
    with rt.Solver(p) as s:
        s.solve()
        ok, worst = _check(s, ref)
    assert ok, f"dt {dt}: err/tol {worst:.3e}"


@pytest.mark.parametrize("chain", [2, 4, 8])
@pytest.mark.parametrize("bcl", [2, 1])
def test_chained_warp_shapes(rt, oracle, monkeypatch, chain, bcl):
    monkeypatch.setenv("RT_CUDA_STEP_CHAIN", str(chain))
    tile = 256 * chain
    p = synthetic_parameters(rt, M=8, G=3, N=2 * tile + 37, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=1)
    ref = _oracle_state(oracle, rt, p, 8)
    with rt.Solver(p) as s:
        assert s.plan().step_kernel == 4 and s.plan().step_chain == chain
        s.solve()
        ok, worst = _check(s, ref)
    assert ok, f"chain {chain}: err/tol {worst:.3e}"


@pytest.mark.parametrize("N", [1, 7, 8, 9, 255, 256, 257, 511, 513, 767, 768, 769, 1023, 1024, 1025, 1279, 2047, 2048, 2049, 3333])
@pytest.mark.parametrize("bcl", [2, 1])
def test_rows_ending_anywhere_in_the_tile(rt, oracle, N, bcl):
    """the row ends in warp 0/1/2/3's sub-tile, at a sub-tile or tile boundary, one cell past it, inside the first lane"""
    p = synthetic_parameters(rt, M=4, G=3, N=N, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=1)
    ref = _oracle_state(oracle, rt, p, 8)
    with rt.Solver(p) as s:
        s.solve()
        ok, worst = _check(s, ref)
    assert ok, f"N={N}: err/tol {worst:.3e}"


@pytest.mark.parametrize("kernel", [4, 1])
@pytest.mark.parametrize("N", [300, 700])
def test_many_tiles_per_cta_with_idle_warps(rt, oracle, monkeypatch, kernel, N):
    """32768 short rows: every CTA walks dozens of tiles in which its last warp(s) have no cells.  With one `done` counter
    written by the last warp, warp 0 ran ahead of the working warps and overwrote their carry slots (round-1 ADVICE)."""
    monkeypatch.setenv("RT_CUDA_STEP_KERNEL", str(kernel))
    p = synthetic_parameters(rt, M=16, G=2048, N=N, dt=1e-3, V=1.0, steps=2, bc_left=1, bc_right=1)
    o = make_oracle(oracle, rt, p, num_threads=8)
    o.run_stages(8)
    B, ophi = o.group_terms()["B"], o.moments()[0]
    o.close()
    with rt.Solver(p, fetch_psi=False) as s:
        s.solve()
        ok, worst, _ = state_ok(s.moments()[0], ophi, B, group_axis=0)
    assert ok, f"kernel {kernel} N={N}: phi err/tol {worst:.3e}"


def _solve_with_step_output(rt, p, fused_phi, monkeypatch, **kw):
    monkeypatch.setenv("RT_CUDA_STEP_PHI", "1" if fused_phi else "0")
    pin = C.c_void_p()
    n = p.G * p.N
    rt.capi.check(rt.capi.lib().rt_host_alloc(C.byref(pin), n * 8))
    try:
        with rt.Solver(p, fetch_psi=False, **kw) as s:
            s.set_step_output(pin, True)
            s.solve()
            out = np.ctypeslib.as_array(C.cast(pin, C.POINTER(C.c_double)), shape=(n,)).reshape(p.G, p.N).copy()
            phi = s.moments_group_major()[0]
            tm = s.timing()
    finally:
        rt.capi.lib().rt_host_free(pin)
    return out, phi, tm


@pytest.mark.parametrize("M", [2, 4, 6, 8, 16, 32])
@pytest.mark.parametrize("bcl,N", [(2, 2100), (1, 1024), (1, 700)])
def test_step_phi_from_the_step_kernel(rt, monkeypatch, M, bcl, N):
    """per-step scalar intensity accumulated by the step kernel (runs of 1/2/3/4/8 rows, 1 or 2 slots per sign) == the
    separate moments pass over the final state, up to the summation order (reference: ascending m, solver.cpp:191-204)"""
    p = synthetic_parameters(rt, M=M, G=5, N=N, dt=1e-3, V=1.0, steps=3, bc_left=bcl, bc_right=1)
    fused, phi, tm = _solve_with_step_output(rt, p, True, monkeypatch)
    staged, phi2, _ = _solve_with_step_output(rt, p, False, monkeypatch)
    # the state itself does not depend on who sums phi (the PHI / non-PHI instantiations are separate compilations of the
    # kernel and may contract multiply-adds differently: agreement to rounding, not bitwise)
    assert state_err(phi, phi2, group_axis=0) <= 1e-13, state_err(phi, phi2, group_axis=0)
    assert np.array_equal(staged, phi2)       # the moments pass of the last step is the final phi
    assert state_err(fused, phi, group_axis=0) <= 1e-13, state_err(fused, phi, group_axis=0)
    assert np.isfinite(fused).all()


def test_step_phi_batched(rt, monkeypatch):
    """group batches: every batch leaves its rows of the per-step output"""
    p = synthetic_parameters(rt, M=8, G=6, N=3000, dt=1e-3, V=1.0, steps=2)
    fused, phi, _ = _solve_with_step_output(rt, p, True, monkeypatch, groups_per_batch=4)
    assert state_err(fused, phi, group_axis=0) <= 1e-13


def test_full_size_rows_c5_shape(rt, oracle):
    """N = 2^20 cells, S32, both boundaries sources (C5's shape), two BDF2 steps: 1024 chained tiles per row"""
    p = synthetic_parameters(rt, M=32, G=2, N=1 << 20, dt=1e-3, V=1.0, steps=2, bc_left=1, bc_right=1)
    o = make_oracle(oracle, rt, p, num_threads=2)
    o.solve()
    B, ophi, oends = o.group_terms()["B"], o.moments()[0], o.ends(0)
    o.close()
    with rt.Solver(p, fetch_psi=False) as s:
        s.solve()
        assert s.stages_done == 8
        ok, worst, _ = state_ok(s.moments()[0], ophi, B, group_axis=0)
        assert ok, f"phi err/tol {worst:.3e}"
        ok, worst, _ = state_ok(s.ends(0), oends, B)
        assert ok, f"ends err/tol {worst:.3e}"


@pytest.mark.parametrize("dt,V", [(1e-5, 1.0), (1e-3, 0.0)])
def test_full_size_rows_local_regime_and_no_correction(rt, oracle, dt, V):
    """N = 2^20, S16 x 2 groups: dt = 1e-5 (alpha ~ 0.01) and V = 0 (the CORR = false instantiation, stage 2 skipped)"""
    p = synthetic_parameters(rt, M=16, G=2, N=1 << 20, dt=dt, V=V, steps=1)
    o = make_oracle(oracle, rt, p, num_threads=2)
    o.solve()
    B, ophi, oends = o.group_terms()["B"], o.moments()[0], o.ends(0)
    o.close()
    with rt.Solver(p, fetch_psi=False) as s:
        s.solve()
        ok, worst, _ = state_ok(s.moments()[0], ophi, B, group_axis=0)
        assert ok, f"phi err/tol {worst:.3e}"
        ok, worst, _ = state_ok(s.ends(0), oends, B)
        assert ok, f"ends err/tol {worst:.3e}"


def test_kernel_is_chosen_by_the_job(rt, monkeypatch):
    """default (no RT_CUDA_STEP_KERNEL): plain steps run the four-scan kernel; once the per-step scalar intensity is asked
    for, whole steps run the one-scan kernel, which sums it in the same pass"""
    monkeypatch.delenv("RT_CUDA_STEP_KERNEL")
    p = synthetic_parameters(rt, M=8, G=4, N=2100, dt=1e-3, V=1.0, steps=2)
    pin = C.c_void_p()
    rt.capi.check(rt.capi.lib().rt_host_alloc(C.byref(pin), p.G * p.N * 8))
    try:
        with rt.Solver(p, fetch_psi=False) as s:
            assert s.plan().step_kernel == 1
            s.set_step_output(pin, True)
            assert s.plan().step_kernel == 4
            s.solve()
            out = np.ctypeslib.as_array(C.cast(pin, C.POINTER(C.c_double)), shape=(p.G * p.N,)).reshape(p.G, p.N).copy()
            assert state_err(out, s.moments_group_major()[0], group_axis=0) <= 1e-13
            s.set_step_output(None, True)
            assert s.plan().step_kernel == 1
    finally:
        rt.capi.lib().rt_host_free(pin)
