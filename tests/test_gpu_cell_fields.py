"""GPU parity of the cell-dependent material extension (csrc/var_kernel.cu) through the C ABI.

The reference's slab is uniform (solver.cpp:150-157 TODO), so: (1) with UNIFORM per-cell arrays the cell-dependent
kernels must agree with the uniform kernels and with the reference-pinned oracle; (2) with varying fields they are
checked against the oracle's cell-dependent restatement (tests/test_oracle_cell_fields.py pins that one to the uniform
oracle bit for bit for uniform arrays) at the same 1e-9 tolerance, every time-stepping method and boundary condition."""
import os

import numpy as np
import pytest

from parity_util import CELL_FLOOR, make_oracle, state_err, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu


def _fields(p, seed=11, T_amp=0.3, rho_amp=0.5, kap_dec=1.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    x = (np.arange(p.N) + 0.5) / p.N
    p.T_cell = 1.0 + T_amp * np.sin(2 * np.pi * x) ** 2 + (0.2 if T_amp else 0.0) * (x > 0.6)
    p.rho_cell = 1.0 + rho_amp * np.cos(3 * np.pi * x)
    p.kappa_cell = np.asarray(p.group_kappa)[:, None] * 10.0 ** rng.uniform(-kap_dec, kap_dec, size=(p.G, p.N))
    return p


def _compare(rt, oracle, p, stages, threads=4):
    o = make_oracle(oracle, rt, p, num_threads=threads)
    o.run_stages(stages)
    B = np.maximum(o.group_terms()["B"], 1e-300)
    ref = dict(psi=o.psi(), ends=o.ends(0), prev=o.ends(1), half=o.ends(2), phi=o.moments()[0], bal=o.balance(),
               ge=o.group_ends(), src=o.psi_source())
    o.close()
    with rt.Solver(p) as s:
        s.run_stages(stages)
        # whole BDF2 steps run as ONE fused launch each (var_step_kernel.cu) unless RT_CUDA_VAR_FUSED=0 / the chain-walk kernel
        fused = p.ts_method == 3 and os.environ.get("RT_CUDA_VAR_FUSED", "1") != "0" and os.environ.get("RT_CUDA_VAR_THREADS", "64") != "0" \
            and os.environ.get("RT_CUDA_VAR_CPT", "4") == "4" and os.environ.get("RT_CUDA_FUSED", "1") != "0"
        assert s.timing().step_launches == (stages // 4 if fused else 0)
        worst = 0.0
        for got, want in ((s.psi_mat_ref, ref["psi"]), (s.ends(0), ref["ends"]), (s.ends(1), ref["prev"]), (s.ends(2), ref["half"])):
            ok, w, _ = state_ok(got, want, B, floor=CELL_FLOOR)
            assert ok, f"err/tol {w:.3e}"
            worst = max(worst, w)
        ok, w, _ = state_ok(s.moments()[0], ref["phi"], B, group_axis=0, floor=CELL_FLOOR)
        assert ok, f"phi err/tol {w:.3e}"
        s.compute_balance()
        s.compute_group_ends()
        assert np.allclose(s.psi_source(), ref["src"], rtol=1e-12, atol=0)
        scale = np.abs(ref["phi"]).max()
        assert np.all(np.abs(s.get_ends("left") - ref["ge"][0]) <= 1e-9 * np.abs(ref["ge"][0]).max())
        assert np.all(np.abs(s.get_ends("right") - ref["ge"][1]) <= 1e-9 * np.abs(ref["ge"][1]).max())
        assert np.all(np.abs(s.get_balance() - ref["bal"]) <= 1e-9 * np.maximum(1.0, np.abs(ref["bal"])))
        assert np.isfinite(scale)
    return worst


@pytest.mark.parametrize("bcl,bcr", [(2, 1), (1, 1), (0, 0)])
def test_uniform_arrays_match_the_uniform_kernels_and_the_oracle(rt, oracle, bcl, bcr):
    p = synthetic_parameters(rt, M=8, G=5, N=3000, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=bcr)
    with rt.Solver(p) as s:
        s.solve()
        uni = (s.psi_mat_ref.copy(order="F"), s.ends(0))
    p.rho_cell = np.full(p.N, p.rho)
    p.T_cell = np.full(p.N, p.T)
    p.kappa_cell = np.repeat(np.asarray(p.group_kappa)[:, None], p.N, axis=1)
    _compare(rt, oracle, p, 8)
    with rt.Solver(p) as s:
        s.solve()
        assert state_err(s.psi_mat_ref, uni[0]) <= 1e-11 and state_err(s.ends(0), uni[1]) <= 1e-11


@pytest.mark.parametrize("ts,bcl,bcr", [(3, 2, 1), (3, 1, 1), (3, 0, 0), (1, 1, 1), (2, 2, 1)])
def test_varying_fields_vs_oracle(rt, oracle, ts, bcl, bcr):
    p = synthetic_parameters(rt, M=8, G=6, N=2500, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=bcr, ts_method=ts)
    _fields(p)
    _compare(rt, oracle, p, 2 * (4 if ts == 3 else 1))


@pytest.mark.parametrize("N", [1, 3, 4, 5, 511, 512, 513, 1025])
def test_varying_fields_ragged_rows(rt, oracle, N):
    p = synthetic_parameters(rt, M=4, G=3, N=N, dt=1e-3, V=1.0, steps=2, bc_left=2, bc_right=1)
    _fields(p)
    _compare(rt, oracle, p, 8, threads=2)


def test_no_correction_and_static_medium(rt, oracle):
    for V, corr in ((0.0, True), (1.0, False)):
        p = synthetic_parameters(rt, M=4, G=4, N=1500, dt=1e-4, V=V, steps=2, bc_left=1, bc_right=1, use_correction=corr)
        _fields(p)
        _compare(rt, oracle, p, 8, threads=2)


def test_only_temperature_given_and_group_shards(rt, oracle):
    """missing fields are filled with the scalars; a shard of the groups equals the same rows of the whole solve"""
    p = synthetic_parameters(rt, M=4, G=6, N=900, dt=1e-3, V=1.0, steps=1, bc_left=2, bc_right=1)
    p.T_cell = 1.0 + 0.4 * (np.arange(p.N) > 300)
    _compare(rt, oracle, p, 4, threads=2)
    with rt.Solver(p) as s:
        s.solve()
        whole = s.ends(0)
    with rt.Solver(p, group_begin=2, group_count=3) as s:
        s.solve()
        assert np.array_equal(s.ends(0), whole[:, 2:5])


def test_per_cell_tables_are_sharded_with_the_groups(rt):
    """a handle that owns groups [g0, g0+n) keeps kappa_g(c) of those groups and ONE neighbour on either side (edge opacities
    interpolate between group centres, correction.cpp:125-139) plus all groups at the two boundary cells (equilibrium
    boundary sources) -- nothing else of the G x N input is read: poison it and the shard's result does not change;
    the owner of the last group still gets the grey remainder over ALL groups (integrated on the fly, not tabulated)"""
    p = synthetic_parameters(rt, M=4, G=12, N=700, dt=1e-3, V=1.0, steps=2, bc_left=2, bc_right=1)
    _fields(p)
    with rt.Solver(p) as s:
        s.solve()
        whole = s.ends(0)
    clean = p.kappa_cell.copy()
    for g0, n in ((4, 3), (0, 2), (9, 3)):
        p.kappa_cell = clean.copy()
        outside = [g for g in range(p.G) if g < g0 - 1 or g > g0 + n]
        p.kappa_cell[outside, 1:-1] = np.nan
        with rt.Solver(p, group_begin=g0, group_count=n) as s:
            s.solve()
            got = s.ends(0)
            assert np.isfinite(got).all()
            assert np.array_equal(got, whole[:, g0:g0 + n]), (g0, n)
    p.kappa_cell = clean


def test_short_scan_kernel_for_small_time_steps(rt, oracle, monkeypatch):
    """at small time steps no row can feel more than the first two steps of the warp scan (alpha(sigma_min)^16 < 2^-70 for
    every group and direction): the solver picks the instantiation that runs two (rt_plan::var_scan_steps) -- checked
    against the oracle like every other case and against the full scan; at dt = 1e-3 it must NOT be picked"""
    # N = 3000: the rows end inside a tile, whose identity padding must still carry the aggregate to the reflective junction
    p = synthetic_parameters(rt, M=8, G=6, N=3000, dt=1e-5, V=1.0, steps=2, bc_left=2, bc_right=1)
    _fields(p)
    with rt.Solver(p) as s:
        assert s.plan().var_scan_steps == 2
        s.solve()
        short = s.ends(0)
    _compare(rt, oracle, p, 8)
    for N, bcl in ((2048, 2), (4097, 1), (300, 2)):
        q = synthetic_parameters(rt, M=4, G=3, N=N, dt=1e-5, V=1.0, steps=2, bc_left=bcl, bc_right=1)
        _fields(q)
        _compare(rt, oracle, q, 8, threads=2)
    monkeypatch.setenv("RT_CUDA_VAR_SCAN_STEPS", "5")
    with rt.Solver(p) as s:
        assert s.plan().var_scan_steps == 5
        s.solve()
        assert state_err(short, s.ends(0)) <= 1e-13
    monkeypatch.delenv("RT_CUDA_VAR_SCAN_STEPS")
    p = synthetic_parameters(rt, M=8, G=6, N=3000, dt=1e-3, V=1.0, steps=1, bc_left=2, bc_right=1)
    _fields(p)
    with rt.Solver(p) as s:
        assert s.plan().var_scan_steps == 5


def test_validation_is_refused_with_cell_fields(rt):
    p = synthetic_parameters(rt, M=4, G=3, N=100, dt=1e-3, V=1.0, steps=1)
    p.include_validation = True
    p.T_cell = np.full(p.N, 1.0)
    with pytest.raises(rt.capi.RtError):
        rt.Solver(p)


@pytest.mark.parametrize("var_threads", ["64", "128", "0"])
@pytest.mark.parametrize("ts,bcl,bcr", [(3, 2, 1), (2, 1, 1)])
def test_every_cell_dependent_kernel_variant(rt, oracle, monkeypatch, var_threads, ts, bcl, bcr):
    """the tile kernel in both shapes (64 / 128 threads x 4 cells: var_tile_kernel.cu) and the chain-walk cross-check
    kernel (RT_CUDA_VAR_THREADS=0: var_kernel.cu) against the oracle, rows longer than several tiles and ending inside one"""
    monkeypatch.setenv("RT_CUDA_VAR_THREADS", var_threads)
    p = synthetic_parameters(rt, M=8, G=5, N=2 * 512 + 77, dt=1e-3, V=1.0, steps=2, bc_left=bcl, bc_right=bcr, ts_method=ts)
    _fields(p)
    _compare(rt, oracle, p, 2 * (4 if ts == 3 else 1))


def test_tile_kernel_is_bitwise_reproducible(rt):
    """the look-back always takes the predecessor's inclusive value, so two runs agree bit for bit"""
    p = synthetic_parameters(rt, M=8, G=6, N=5000, dt=1e-3, V=1.0, steps=2, bc_left=2, bc_right=1)
    _fields(p)
    runs = []
    for _ in range(2):
        with rt.Solver(p) as s:
            s.solve()
            runs.append(s.ends(0).copy(order="F"))
    assert np.array_equal(runs[0], runs[1])


@pytest.mark.parametrize("bcl,bcr,V", [(2, 1, 1.0), (1, 1, 1.0), (0, 0, 0.0)])
def test_fused_step_equals_the_staged_step(rt, oracle, monkeypatch, bcl, bcr, V):
    """cell-dependent fields: the fused BDF2 step (var_step_kernel.cu) against the same steps run stage by stage
    (var_tile_kernel.cu) -- same stage algebra, same per-cell constants, so they agree to rounding -- after whole steps,
    after a partial step on top (stage-by-stage continuation from a fused state), and prev_ends / half_ends after a fused step"""
    p = synthetic_parameters(rt, M=8, G=4, N=3 * 256 + 100, dt=1e-3, V=V, steps=3, bc_left=bcl, bc_right=bcr)
    _fields(p)
    out = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("RT_CUDA_VAR_FUSED", mode)
        with rt.Solver(p) as s:
            s.run_stages(8)
            assert s.timing().step_launches == (2 if mode == "1" else 0)
            a = dict(ends=s.ends(0), prev=s.ends(1), half=s.ends(2))
            s.run_stages(2)  # a partial step continues stage by stage
            a["partial"] = s.ends(0)
            s.run_stages(2)
            a["final"] = s.ends(0)
            out[mode] = a
    for k in out["1"]:
        ref = out["0"][k]
        scale = np.abs(ref).max()
        assert np.max(np.abs(out["1"][k] - ref)) <= 1e-12 * scale, k
