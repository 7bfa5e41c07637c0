"""CPU tests of the two scan FORMULATIONS the kernels implement, stated in plain numpy and run serially:

 * scan_form.py  -- every stepper (solver.cpp:319-587) as the affine recurrence v_j = alpha v_(j-1) + beta_j with the
   sweep-ordered (u, v) layout, the boundary multipliers and the buffer aliasing of prev / half (csrc/sweep_kernel.cu,
   csrc/rt_capi.cu run_one_stage), against the oracle on every golden case;
 * step4_form.py -- a whole BDF2 step as ONE recurrence in the 4-vector of the stages' outflow nodes with a chunk transfer
   matrix that is a row constant (csrc/step4_kernel.cu), in the kernel's own structure (lanes, sub-tiles, carries,
   junction, ragged rows), against the serial stage sweeps.

No GPU: these pin the algebra the device code transliterates, so that a kernel bug and a formulation bug can be told apart."""
import importlib

import numpy as np

from conftest import GOLDEN_CASES, host_setup
from parity_util import load_case, rel_err


def test_stage_sweeps_as_affine_recurrences_match_the_oracle(oracle):
    import scan_form
    rt = importlib.import_module("radiative-transfer_b200")
    for name in GOLDEN_CASES:
        p = load_case(rt, name)
        if p.max_timesteps > 20:
            p.max_timesteps = 3
        mu, wt, e, kap = host_setup(rt, p)
        n = p.max_timesteps * (4 if p.ts_method == 3 else 1)
        o = oracle.oracle_from_parameters(p, mu, wt, e, kap)
        o.run_stages(n)
        ref = o.ends(0)
        o.close()
        got = scan_form.run(rt, oracle, p, mu, wt, e, kap, n)
        assert rel_err(got, ref) <= 1e-9, (name, rel_err(got, ref))


def test_one_scan_formulation_of_a_whole_step():
    import step4_form
    assert step4_form.main() < 1e-10


# ---------------------------------------------------------------------------------------------------------------------
# the two scan shortcuts of round 2, stated on the CPU (numpy models of the warp scans; the kernels are checked on the GPU)
def _seeded_scan(alpha_chunk, incl, carry, steps):
    """step_warp_kernel.cu `scan`: 32 lanes, lane-chunk aggregates `incl` with zero inflow, the chunk factor alpha^CPT uniform;
    the carry is folded into lane 0, then `steps` Kogge-Stone steps with the factors alpha^(CPT 2^s)"""
    import numpy as np
    v = incl.copy()
    v[0] = alpha_chunk * carry + v[0]
    f = alpha_chunk
    for s in range(steps):
        d = 1 << s
        up = np.concatenate([np.zeros(d), v[:-d]])
        v = np.where(np.arange(32) >= d, f * up + v, v)
        f = f * f
    return v


def test_local_rows_need_one_scan_step_only():
    """SC_LOCAL (pack_step_coef_kernel): with alpha^16 < 2^-70 the steps s >= 1 of the seeded scan cannot change any lane's
    value by more than 2^-70 of a neighbouring value -- one step reproduces the serial recurrence to the last bits"""
    import numpy as np
    rng = np.random.Generator(np.random.PCG64(5))
    for alpha in (0.0478, 0.03, 0.0126):  # 0.0478^16 = 2^-70.2: the largest alpha that still counts as local
        assert alpha ** 16 < 2.0 ** -70
        a8 = alpha ** 8
        incl = rng.uniform(0.5, 2.0, 32) * 10.0 ** rng.integers(-3, 4, 32)  # neighbouring chunks of very different size
        carry = 7.0
        serial = np.zeros(32)
        c = carry
        for i in range(32):
            c = a8 * c + incl[i]
            serial[i] = c
        one, five = _seeded_scan(a8, incl, carry, 1), _seeded_scan(a8, incl, carry, 5)
        assert np.max(np.abs(one - serial) / np.abs(serial)) < 1e-15
        assert np.max(np.abs(one - five) / np.abs(five)) < 1e-15
    # and a row that is NOT local does need them
    a8 = 0.9 ** 8
    incl = rng.uniform(0.5, 2.0, 32)
    assert np.max(np.abs(_seeded_scan(a8, incl, 7.0, 1) - _seeded_scan(a8, incl, 7.0, 5))) > 0.1


def _pair_scan(A, W, steps):
    """var_step_kernel.cu `scan`: inclusive Kogge-Stone scan of the pairs (A, W) = (product of a lane's alphas, its outflow
    with zero inflow) under (A2, W2) o (A1, W1) = (A2 A1, A2 W1 + W2)"""
    import numpy as np
    A, W = A.copy(), W.copy()
    for s in range(steps):
        d = 1 << s
        Au, Wu = np.concatenate([np.ones(d), A[:-d]]), np.concatenate([np.zeros(d), W[:-d]])
        take = np.arange(32) >= d
        W = np.where(take, A * Wu + W, W)
        A = np.where(take, A * Au, A)
    return A, W


def test_short_pair_scan_and_the_ragged_tile():
    """var_step_kernel<..., NSTEPS = 2>: with alpha_max^16 < 2^-70 two steps are enough for every REAL lane -- but a row's
    last tile is padded with identity elements (alpha = 1, beta = 0) whose products do not die out: after two steps lane 31
    does not hold the warp's aggregate, the warp's last real lane does (what the kernel publishes since the fix)"""
    import numpy as np
    rng = np.random.Generator(np.random.PCG64(6))
    cpt, real = 4, 15  # 15 lanes of real cells, 17 lanes of padding
    alpha = rng.uniform(0.01, 0.045, (32, cpt))
    beta = rng.uniform(0.5, 2.0, (32, cpt))
    alpha[real:], beta[real:] = 1.0, 0.0
    assert 0.045 ** 16 < 2.0 ** -70
    A = np.prod(alpha, axis=1)
    W = np.zeros(32)
    for j in range(cpt):
        W = alpha[:, j] * W + beta[:, j]
    v, serial = 0.0, np.zeros(32)  # the serial recurrence over all cells, zero inflow
    for i in range(32):
        for j in range(cpt):
            v = alpha[i, j] * v + beta[i, j]
        serial[i] = v
    _, W5 = _pair_scan(A, W, 5)
    _, W2 = _pair_scan(A, W, 2)
    assert np.max(np.abs(W5 - serial) / np.abs(serial)) < 1e-14
    assert np.max(np.abs(W2[:real] - serial[:real]) / np.abs(serial[:real])) < 1e-14  # every real lane is right ...
    assert abs(W2[real - 1] - serial[31]) / abs(serial[31]) < 1e-14                   # ... the last real lane holds the aggregate ...
    assert abs(W2[31] - serial[31]) / abs(serial[31]) > 0.5                           # ... and lane 31 does not (the first version's bug)
