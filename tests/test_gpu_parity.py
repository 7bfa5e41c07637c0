"""GPU parity tests proper: the CUDA path (through the C ABI, via the Python mirror of rt::Solver) against
 (a) the golden vectors dumped by the UNMODIFIED reference (tests/golden/*.npz), and
 (b) the CPU oracle (oracle-2) on seeded synthetic slabs at sizes the oracle finishes in seconds.

Tolerance (north_star): 1e-9 relative on psi / ends / scalar intensity at every BDF2 step, with the SAME number of
stage sweeps as the reference.  The radiative flux F and the balance are differences of O(phi) terms; they are held to
1e-9 of that scale.
"""
import numpy as np
import pytest

from conftest import GOLDEN_CASES
from parity_util import RTOL, flux_err, load_case, load_golden, make_oracle, rel_err, scaled_err, state_err, state_ok, synthetic_parameters

pytestmark = pytest.mark.gpu


def _compare_state(tag, s, ref_psi, ref_ends, B):
    ok_p, w_p, pg = state_ok(s.psi_mat_ref, ref_psi, B)
    ok_e, w_e, _ = state_ok(s.ends(0), ref_ends, B)
    print(f"{tag}: psi err/tol {w_p:.2e} ends err/tol {w_e:.2e}; worst group err {pg.max():.2e} "
          f"(pure elementwise rel: {rel_err(s.psi_mat_ref, ref_psi):.2e})")
    assert ok_p, f"{tag}: psi off, err/tol = {w_p:.3e}"
    assert ok_e, f"{tag}: ends off, err/tol = {w_e:.3e}"


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden_final_state(rt, name):
    """every buffer the reference exposes or owns, after the reference's full stage count"""
    p = load_case(rt, name)
    g = load_golden(name)
    with rt.Solver(p) as s:
        assert np.array_equal(s.m_mu, g["mu"]) and np.array_equal(s.m_wt, g["wt"])
        assert np.array_equal(s.e_edge, g["e_edge"]) and np.array_equal(s.kappa_vec, g["kappa"])
        assert state_ok(s.psi_mat_ref, np.broadcast_to(g["B_init"][None, :, None], s.psi_mat_ref.shape), g["B"])[0]
        assert s.total_stages == p.max_timesteps * (4 if p.ts_method == 3 else 1)
        s.solve()
        assert s.stages_done == s.total_stages  # identical iteration count
        _compare_state(name, s, g["psi"], g["ends"], g["B"])
        for which, key in ((1, "prev_ends"), (2, "half_ends")):
            assert state_ok(s.ends(which), g[key], g["B"])[0], f"{name}: {key}"
        s.compute_angle_integrated_intensity()
        s.compute_positive_angle_integrated_intensity()
        s.compute_radiative_flux()
        s.compute_balance()
        s.compute_group_ends()
        scale = float(np.max(np.abs(g["phi"])))
        assert state_ok(s.phi_ref, g["phi"], g["B"], group_axis=0)[0]
        assert state_ok(s.get_phi_plus(), g["phi_plus"], g["B"], group_axis=0)[0]
        assert flux_err(s.F_ref, g["F"], g["phi"]) <= RTOL
        assert scaled_err(s.get_balance(), g["balance"], 1.0) <= 1e-10  # observed <= 1e-12 (tools/gpu_tolerance_audit.py)
        assert state_ok(s.get_ends("left"), g["left_ends"], g["B"], group_axis=0, floor=0.0)[0]
        assert state_ok(s.get_ends("right"), g["right_ends"], g["B"], group_axis=0, floor=0.0)[0]
        assert np.array_equal(s.get_e_ave(), g["e_ave"])
        gt = s.group_terms()
        for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB", "kappa_edge"):
            bscale = float(np.max(np.abs(g[k]))) or 1.0
            assert scaled_err(gt[k], g[k], bscale) <= 1e-12, f"{name}: {k}"
        # B and dB/dT group by group (the small groups too); the last group is a remainder of the grey total
        assert rel_err(gt["B"][:-1], g["B"][:-1]) <= RTOL and rel_err(gt["dBdT"][:-1], g["dBdT"][:-1]) <= RTOL
        assert scaled_err(s.psi_source(), g["psi_source"], float(np.max(np.abs(g["psi_source"]))) or 1.0) <= 1e-12


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_golden_step_snapshots(rt, name):
    """parity at intermediate time steps against reference runs with max_timesteps = k"""
    p = load_case(rt, name)
    g = load_golden(name)
    per_step = 4 if p.ts_method == 3 else 1
    ks = sorted(int(k[len("psi_steps"):]) for k in g if k.startswith("psi_steps"))
    assert ks
    with rt.Solver(p) as s:
        done = 0
        for k in ks:
            s.run_stages((k - done) * per_step)
            done = k
            _compare_state(f"{name}@step{k}", s, g[f"psi_steps{k}"], g[f"ends_steps{k}"], g["B"])


SYNTH = [
    # name,                         M,  G,  N,     dt,   V,  steps, bcl, bcr
    ("c4like_local_reflect",        16, 6,  4096,  1e-5, 1.0, 2,    2,   1),
    ("c4like_longrange_reflect",    16, 6,  5000,  1e-3, 1.0, 2,    2,   1),
    ("c5like_longrange_source",     32, 4,  6151,  1e-3, 1.0, 2,    1,   1),
    ("very_long_range_vacuum",      4,  3,  20000, 5e-2, 0.5, 1,    0,   0),
    ("v0_reflect_multitile",        8,  5,  2049,  1e-4, 0.0, 2,    2,   1),
    ("mixed_local_and_long_range",  16, 24, 5000,  1e-4, 1.0, 2,    2,   1),
]


def _local_rows(rt, p):
    """rows the fused step kernel treats as LOCAL (step_warp_kernel.cu SC_LOCAL: alpha^16 < 2^-70 for all three stage kinds,
    so its warp scan runs one step instead of five) -- the stepper constants of SURVEY 8(a) S1-S3 / S*"""
    from parity_util import host_setup
    c = 299.792458
    mu, _, _, kap = host_setup(rt, p)
    sig, amu, tau, dx = p.rho * np.asarray(kap)[None, :], np.abs(np.asarray(mu))[:, None], p.dt / 2, p.dx
    worst = 0.0
    for k, d in ((c * tau * amu, 0.5 * ((1 + c * tau * sig) * dx + c * tau * amu)),
                 (0.5 * c * amu * tau, 0.5 * ((1 + 0.5 * c * tau * sig) * dx + 0.5 * c * amu * tau)),
                 (c * amu * p.dt / 6, 0.5 * ((1 + c * sig * tau / 6) * dx + c * amu * p.dt / 6))):
        worst = np.maximum(worst, 0.5 * k * k / (d * d + 0.25 * k * k))
    return worst ** 16 < 2.0 ** -70


def test_the_mixed_case_has_rows_of_both_kinds(rt):
    """mixed_local_and_long_range above puts local and long-range rows into the SAME launch (and the same CTA: tickets run
    over the rows of a tile position), c4like_local_reflect is all local, very_long_range_vacuum has none"""
    frac = {}
    for name, M, G, N, dt, V, steps, bcl, bcr in SYNTH:
        frac[name] = float(np.mean(_local_rows(rt, synthetic_parameters(rt, M=M, G=G, N=N, dt=dt, V=V, steps=steps, bc_left=bcl, bc_right=bcr))))
    assert 0.05 < frac["mixed_local_and_long_range"] < 0.95, frac
    assert frac["c4like_local_reflect"] == 1.0 and frac["very_long_range_vacuum"] == 0.0, frac


@pytest.mark.parametrize("name,M,G,N,dt,V,steps,bcl,bcr", SYNTH)
def test_synthetic_vs_oracle_each_step(rt, oracle, name, M, G, N, dt, V, steps, bcl, bcr):
    """multi-tile rows (N > 2048), ragged N, alpha close to 1 (long-range coupling through the look-back), every BC"""
    p = synthetic_parameters(rt, M=M, G=G, N=N, dt=dt, V=V, steps=steps, bc_left=bcl, bc_right=bcr)
    o = make_oracle(oracle, rt, p, num_threads=8)
    B = o.group_terms()["B"]
    with rt.Solver(p) as s:
        for step in range(steps):
            for stage in range(4):
                o.run_stages(1)
                s.run_stages(1)
                _compare_state(f"{name}@step{step}.{stage}", s, o.psi(), o.ends(0), B)
        phi, pp, F = s.moments()
        ophi, opp, oF = o.moments()
        assert state_ok(phi, ophi, B, group_axis=0)[0] and state_ok(pp, opp, B, group_axis=0)[0]
        assert flux_err(F, oF, ophi) <= RTOL
        gm = s.moments_group_major()
        assert np.array_equal(np.asfortranarray(gm[0]), phi)
    o.close()


@pytest.mark.parametrize("ts_method", [1, 2])
def test_synthetic_be_cn_only(rt, oracle, ts_method):
    p = synthetic_parameters(rt, M=8, G=3, N=4500, dt=2e-4, V=2.0, steps=5, ts_method=ts_method)
    o = make_oracle(oracle, rt, p, num_threads=8)
    o.solve()
    B = o.group_terms()["B"]
    with rt.Solver(p) as s:
        s.solve()
        _compare_state(f"ts{ts_method}", s, o.psi(), o.ends(0), B)
        assert state_ok(s.ends(1), o.ends(1), B)[0]
    o.close()


def test_batched_equals_resident(rt):
    """group batching inside one GPU is exact (groups never couple): bit-identical moments"""
    p = synthetic_parameters(rt, M=8, G=7, N=3000, dt=1e-3, steps=2)
    with rt.Solver(p) as a:
        a.solve()
        a.compute_balance()
        ra = a.moments() + (a.get_balance(),)
        a.compute_group_ends()
        la, rra = a.get_ends("left"), a.get_ends("right")
    with rt.Solver(p, groups_per_batch=3, fetch_psi=False) as b:
        assert b.plan().n_batches == 3 and not b.resident
        b.solve()
        rb = b.moments()
        b.compute_balance()
        b.compute_group_ends()
        for x, y in zip(ra[:3], rb):
            assert np.array_equal(x, y)
        assert np.array_equal(ra[3], b.get_balance())
        assert np.array_equal(la, b.get_ends("left")) and np.array_equal(rra, b.get_ends("right"))


def test_group_shards_equal_whole(rt):
    """a handle that owns groups [g0, g0+n) reproduces exactly those rows of the whole-problem solve"""
    p = synthetic_parameters(rt, M=4, G=9, N=2500, dt=1e-3, steps=1)
    with rt.Solver(p) as whole:
        whole.solve()
        phi = whole.moments()[0]
        psi = whole.psi_mat_ref.copy(order="F")
    for g0, n in ((0, 4), (4, 5)):
        with rt.Solver(p, group_begin=g0, group_count=n) as part:
            part.solve()
            assert np.array_equal(part.moments()[0], phi[g0:g0 + n])
            assert np.array_equal(part.psi_mat_ref, psi[:, g0:g0 + n])


def test_step_output_matches_moments(rt):
    import ctypes as C
    p = synthetic_parameters(rt, M=4, G=5, N=2100, dt=1e-3, steps=3)
    with rt.Solver(p, groups_per_batch=2, fetch_psi=False) as s:
        buf = np.zeros((p.G, p.N))
        s.set_step_output(buf.ctypes.data_as(C.c_void_p), True)
        s.solve()
        # the step kernel adds up w_m psi itself (per run of rows and per sign, then a combine): the same M terms as the
        # moments pass (ascending m, solver.cpp:191-204) in another association -> equal to rounding, not bit for bit
        from parity_util import state_err
        assert state_err(buf, s.moments_group_major()[0], group_axis=0) <= 1e-13


def test_validation_failure_is_loud(rt):
    """include_validation with opacities that violate the emission identity -> rt_solve reports it (solver.cpp:609-612)"""
    p = synthetic_parameters(rt, M=2, G=4, N=64, dt=1e-4, steps=1)
    p.include_validation = True
    with rt.Solver(p) as s:
        ok, sums = s.validate()
        assert not ok
        with pytest.raises(rt.capi.RtError, match="Invalid Correction Terms"):
            s.solve()


def test_planck_kernel_vs_oracle(rt, oracle):
    """device Planck integrals vs the CPU restatement over a temperature sweep (all three branches of Planck.cpp:131-150)"""
    e = rt.grid.group_edges(64, 0.001, 60.0)
    T = np.array([0.05, 0.3, 1.0, 2.5, 10.0])
    B, dB, _ = rt.planck_groups(e, T)
    for k, t in enumerate(T):
        oB, odB = oracle.planck(float(t), e)
        tot, dtot = float(np.sum(oB)), float(np.sum(odB))
        assert scaled_err(B[k], oB, tot) <= 1e-12 and scaled_err(dB[k], odB, dtot) <= 1e-12
        big = oB[:-1] > 1e-250
        assert rel_err(B[k][:-1][big], oB[:-1][big]) <= RTOL
        bigd = odB[:-1] > 1e-250
        assert rel_err(dB[k][:-1][bigd], odB[:-1][bigd]) <= RTOL
