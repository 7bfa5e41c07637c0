#!/usr/bin/env python
"""Per-cell, per-group Planck integrals on the device (rt_planck_groups; reference Planck.cpp:50-229 evaluated for
`ncell` temperatures): kernel time, group integrals per second, and a parity spot check against the CPU oracle.
Usage: python tools/planck_bench.py [ncell] [G]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
rt = importlib.import_module("radiative-transfer_b200")
ncell = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 16
G = int(sys.argv[2]) if len(sys.argv) > 2 else 256
e = rt.grid.group_edges(G, 0.01, 30.0)
rng = np.random.Generator(np.random.PCG64(20261104))
mode = sys.argv[3] if len(sys.argv) > 3 else "random"
if mode == "smooth":  # a temperature profile as a slab has it: neighbouring cells (= neighbouring lanes) are close in T
    x = (np.arange(ncell) + 0.5) / ncell
    T = 10.0 ** (-1.0 + 2.0 * np.sin(np.pi * x) ** 2)
else:
    T = 10.0 ** rng.uniform(-1.0, 1.0, size=ncell)  # 0.1 .. 10 keV: all three branches of Planck.cpp:131-150
rt.planck_groups(e, T[:64])  # warm-up (context, constant upload)
best = None
for _ in range(3):
    B, dB, ms = rt.planck_groups(e, T)
    best = ms if best is None else min(best, ms)
import oracle_py  # TEST INFRASTRUCTURE: spot check only
worst = 0.0
for k in rng.integers(0, ncell, size=8):
    oB, odB = oracle_py.planck(float(T[k]), e)
    big = oB[:-1] > 1e-200
    worst = max(worst, float(np.max(np.abs(B[k][:-1][big] - oB[:-1][big]) / oB[:-1][big])))
out = {"ncell": ncell, "G": G, "T_profile": mode, "kernel_ms": best, "group_integrals_per_s": 2.0 * ncell * (G - 1) / (best * 1e-3),
       "max_rel_err_vs_oracle_8_cells": worst}
print(json.dumps(out))
