#!/usr/bin/env python
"""Quick parity table on the GPU box (diagnostic, no asserts): every golden case, final state vs the reference dump."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
from conftest import GOLDEN_CASES  # noqa: E402
from parity_util import load_case, load_golden, rel_err, scaled_err  # noqa: E402
rt = importlib.import_module("radiative-transfer_b200")
for name in GOLDEN_CASES:
    p = load_case(rt, name)
    g = load_golden(name)
    t0 = time.time()
    try:
        with rt.Solver(p) as s:
            s.solve()
            dt = time.time() - t0
            gt = s.group_terms()
            phi, pp, F = s.moments()
            print(f"{name:28s} M={p.M} G={p.G} N={p.N} stages={s.stages_done:5d} {dt*1e3:8.1f} ms  "
                  f"psi {rel_err(s.psi_mat_ref, g['psi']):.1e} ends {rel_err(s.ends(0), g['ends']):.1e} "
                  f"prev {rel_err(s.ends(1), g['prev_ends']):.1e} half {rel_err(s.ends(2), g['half_ends']):.1e} "
                  f"phi {rel_err(phi, g['phi']):.1e} F/phi {scaled_err(F, g['F'], np.max(np.abs(g['phi']))):.1e} "
                  f"B {rel_err(gt['B'], g['B']):.1e} dBdT {rel_err(gt['dBdT'], g['dBdT']):.1e} dEB {rel_err(gt['dEB'], g['dEB']):.1e} "
                  f"dsig {rel_err(gt['dsigEdE'], g['dsigEdE']):.1e} dkap {rel_err(gt['dkapEB'], g['dkapEB']):.1e}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"{name}: FAILED {type(e).__name__}: {e}", flush=True)
