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
