"""GPU edge cases and full-size properties of the sweep kernel (through the C ABI).

Edge cases the reference's own cases do not reach: ragged rows (N not a multiple of 8, shorter than a tile, one cell
past a tile boundary), every supported tile shape, the aggregate walk-back look-back mode, run-to-run bitwise
reproducibility.  Full size (N = 2^20, BASELINE.json's cell count): parity against the oracle on a 2-group slice,
equilibrium preservation, batched == resident."""
import numpy as np
import pytest

from parity_util import RTOL, make_oracle, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu


def _run_both(rt, oracle, p, stages, threads=2):
    o = make_oracle(oracle, rt, p, num_threads=threads)
    o.run_stages(stages)
    B = o.group_terms()["B"]
    with rt.Solver(p) as s:
        s.run_stages(stages)
        ok_p, w_p, _ = state_ok(s.psi_mat_ref, o.psi(), B)
        ok_e, w_e, _ = state_ok(s.ends(0), o.ends(0), B)
    o.close()
    return ok_p and ok_e, max(w_p, w_e)


@pytest.mark.parametrize("N", [1, 2, 7, 8, 9, 511, 512, 513, 1025, 4097])
@pytest.mark.parametrize("bcl", [2, 1])
def test_ragged_rows(rt, oracle, N, bcl):
    """rows that end inside a 128-byte chunk / inside a tile / one cell into the next tile (TMA clips and zero-fills)"""
    p = synthetic_parameters(rt, M=4, G=3, N=N, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=1)
    ok, worst = _run_both(rt, oracle, p, 8)
    assert ok, f"N={N}: err/tol {worst:.3e}"


@pytest.mark.parametrize("threads,cpt", [(64, 8), (64, 4), (128, 8), (128, 4), (256, 8), (256, 4)])
def test_tile_shapes(rt, oracle, monkeypatch, threads, cpt):
    monkeypatch.setenv("RT_CUDA_SWEEP_THREADS", str(threads))
    monkeypatch.setenv("RT_CUDA_SWEEP_CPT", str(cpt))
    p = synthetic_parameters(rt, M=8, G=3, N=2 * threads * cpt + 37, dt=1e-3, V=1.0, steps=2)
    ok, worst = _run_both(rt, oracle, p, 8)
    assert ok, f"{threads}x{cpt}: err/tol {worst:.3e}"


@pytest.mark.parametrize("ring", [2, 3])
def test_ring_depths(rt, oracle, monkeypatch, ring):
    monkeypatch.setenv("RT_CUDA_SWEEP_RING", str(ring))
    p = synthetic_parameters(rt, M=4, G=2, N=6000, dt=1e-3, V=1.0, steps=2)
    ok, worst = _run_both(rt, oracle, p, 8)
    assert ok, f"ring {ring}: err/tol {worst:.3e}"


def test_walk_back_mode(rt, oracle, monkeypatch):
    """few chains, many tiles: predecessors are still running -> aggregate walk-back over several tiles"""
    monkeypatch.setenv("RT_CUDA_LOOKBACK", "walk")
    p = synthetic_parameters(rt, M=2, G=1, N=60000, dt=5e-3, V=0.5, steps=2, bc_left=1, bc_right=1)
    ok, worst = _run_both(rt, oracle, p, 8, threads=1)
    assert ok, f"walk-back: err/tol {worst:.3e}"


def test_bitwise_reproducible(rt):
    """default look-back mode: the same launch sequence gives bit-identical results run to run"""
    p = synthetic_parameters(rt, M=8, G=4, N=30000, dt=1e-3, V=1.0, steps=2)
    outs = []
    for _ in range(3):
        with rt.Solver(p) as s:
            s.solve()
            outs.append((s.psi_mat_ref.copy(order="F"), s.ends(0)))
    for a, b in outs[1:]:
        assert np.array_equal(a, outs[0][0]) and np.array_equal(b, outs[0][1])


def test_full_size_row_vs_oracle(rt, oracle):
    """N = 2^20 cells (BASELINE.json's slab): 2048 tiles per row chained by the look-back, S16, 2 groups, one BDF2 step"""
    p = synthetic_parameters(rt, M=16, G=2, N=1 << 20, dt=1e-3, V=1.0, steps=1)
    o = make_oracle(oracle, rt, p, num_threads=2)
    o.solve()
    B = o.group_terms()["B"]
    ophi = o.moments()[0]
    with rt.Solver(p, fetch_psi=False) as s:
        s.solve()
        assert s.stages_done == 4
        phi = s.moments()[0]
        ok, worst, _ = state_ok(phi, ophi, B, group_axis=0)
        assert ok, f"phi err/tol {worst:.3e}"
        ok, worst, _ = state_ok(s.ends(0), o.ends(0), B)
        assert ok, f"ends err/tol {worst:.3e}"
    o.close()


def test_full_size_equilibrium_is_preserved(rt):
    """size-independent property: psi = B_g with Planckian boundary sources and V = 0 is a fixed point of every stepper
    (GrayTest's physics, tests/test_gray.cpp:30-45); checked at N = 2^20, S16, 8 groups over 2 BDF2 steps"""
    p = synthetic_parameters(rt, M=16, G=8, N=1 << 20, dt=1e-3, V=0.0, steps=2, bc_left=1, bc_right=1)
    with rt.Solver(p, fetch_psi=False) as s:
        B = s.group_terms()["B"]
        s.solve()
        phi, _, F = s.moments_group_major()
        want = float(np.sum(s.m_wt)) * B
        rel = np.max(np.abs(phi / want[:, None] - 1.0))
        assert rel < 1e-12, rel
        assert np.max(np.abs(F)) < 1e-10 * float(np.max(want))


def test_full_size_batched_equals_resident(rt):
    p = synthetic_parameters(rt, M=16, G=6, N=1 << 19, dt=1e-3, V=1.0, steps=1)
    with rt.Solver(p, fetch_psi=False) as a:
        a.solve()
        pa = a.moments_group_major()[0]
    with rt.Solver(p, groups_per_batch=4, fetch_psi=False) as b:
        assert b.plan().n_batches == 2
        b.solve()
        pb = b.moments_group_major()[0]
    assert np.array_equal(pa, pb)
