#!/usr/bin/env python
"""GPU-side audit behind the tolerances written in tests/: the errors actually observed, under stricter measures than
the tests used in round 1 (floor 1e-6 of the group's scale instead of 1e-3; F against max|F|; balance absolute)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
from conftest import GOLDEN_CASES
from parity_util import load_case, load_golden, make_oracle, state_ok, synthetic_parameters, group_rtol
import oracle_py
oracle_py.build()
rt = importlib.import_module("radiative-transfer_b200")

def ratio(a, b, B, axis, floor):
    return state_ok(a, b, B, group_axis=axis, floor=floor)[1]

print(f"{'case':32s} {'psi@1e-3':>9s} {'psi@1e-6':>9s} {'ends@1e-6':>9s} {'phi@1e-6':>9s} {'F/maxF':>9s} {'maxF/maxphi':>11s} {'bal abs':>9s}   (state columns: err/tol, tol = 1e-9)")
for name in GOLDEN_CASES:
    p = load_case(rt, name); g = load_golden(name)
    with rt.Solver(p) as s:
        s.solve()
        s.compute_angle_integrated_intensity(); s.compute_radiative_flux(); s.compute_balance()
        mF = float(np.max(np.abs(g["F"]))); mphi = float(np.max(np.abs(g["phi"])))
        fe = float(np.max(np.abs(s.F_ref - g["F"]))) / mF if mF > 0 else 0.0
        print(f"{name:32s} {ratio(s.psi_mat_ref, g['psi'], g['B'], 1, 1e-3):9.2e} {ratio(s.psi_mat_ref, g['psi'], g['B'], 1, 1e-6):9.2e} "
              f"{ratio(s.ends(0), g['ends'], g['B'], 1, 1e-6):9.2e} {ratio(s.phi_ref, g['phi'], g['B'], 0, 1e-6):9.2e} {fe:9.2e} {mF / mphi:11.2e} "
              f"{float(np.max(np.abs(s.get_balance() - g['balance']))):9.2e}", flush=True)
for (nm, M, G, N, dt, V, steps, bl, br) in [("c4like_local", 16, 6, 4096, 1e-5, 1.0, 2, 2, 1), ("c4like_long", 16, 6, 5000, 1e-3, 1.0, 2, 2, 1),
                                             ("c5like_long", 32, 4, 6151, 1e-3, 1.0, 2, 1, 1), ("vacuum_very_long", 4, 3, 20000, 0.05, 0.5, 1, 0, 0),
                                             ("V0", 8, 5, 3000, 1e-3, 0.0, 2, 2, 1)]:
    p = synthetic_parameters(rt, M=M, G=G, N=N, dt=dt, V=V, steps=steps, bc_left=bl, bc_right=br)
    o = make_oracle(oracle_py, rt, p, num_threads=4); o.solve()
    B = o.group_terms()["B"]; ophi, _, oF = o.moments()
    with rt.Solver(p) as s:
        s.solve()
        phi, _, F = s.moments(); s.compute_balance()
        mF = float(np.max(np.abs(oF))); mphi = float(np.max(np.abs(ophi)))
        print(f"{nm:32s} {ratio(s.psi_mat_ref, o.psi(), B, 1, 1e-3):9.2e} {ratio(s.psi_mat_ref, o.psi(), B, 1, 1e-6):9.2e} "
              f"{ratio(s.ends(0), o.ends(0), B, 1, 1e-6):9.2e} {ratio(phi, ophi, B, 0, 1e-6):9.2e} {float(np.max(np.abs(F - oF))) / mF:9.2e} {mF / mphi:11.2e} "
              f"{float(np.max(np.abs(s.get_balance() - o.balance()))) if hasattr(o, 'balance') else float('nan'):9.2e}", flush=True)
    o.close()
