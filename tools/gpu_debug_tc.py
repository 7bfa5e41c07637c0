#!/usr/bin/env python
"""GPU-side diagnosis aid for the temperature-coupled sub-run of bench.py: how T(c) and the table-rebuild time evolve
over the steps of c5r, and which group carries the phi error of the small parity problem."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import bench
rt = importlib.import_module("radiative-transfer_b200")
sharding = importlib.import_module("radiative-transfer_b200.sharding")

M, G, N, l, r = bench.WORKLOADS["c5r"]
if len(sys.argv) > 1: G = int(sys.argv[1])
p = bench.synthetic_parameters(rt, M, G, N, l, r, 1e-3, 24)
p.cv = 1.0
bench.cell_fields(p, N)
with rt.Solver(p, fetch_psi=False) as s:
    prev = 0.0
    for k in range(24):
        s.run_stages(4)
        tm = s.timing()
        T = s.temperature()
        print(f"step {k+1:2d} tables_ms {tm.tables_ms - prev:7.2f}  T min {T.min():.4f} max {T.max():.4f} nan {int(np.isnan(T).sum())} "
              f"max|dT/dc| {np.max(np.abs(np.diff(T))):.3e}", flush=True)
        prev = tm.tables_ms

# the small parity problem, per-group phi error
import oracle_py
oracle_py.build()
M, G, N, steps = 8, 24, 1500, 3
p = bench.synthetic_parameters(rt, M, G, N, 2, 1, 1e-3, steps)
x = (np.arange(N) + 0.5) / N
p.T_cell = 0.8 + 0.4 * np.sin(2 * np.pi * x) ** 2
p.rho_cell = 1.0 + 0.5 * np.cos(3 * np.pi * x)
p.cv = 0.7
mu, wt = rt.grid.gauss_legendre(M)
o = oracle_py.oracle_from_parameters(p, mu, wt, rt.grid.group_edges(G, p.efirst, p.elast), p.group_kappa, num_threads=8)
with rt.Solver(p, fetch_psi=False) as s:
    for _ in range(steps):
        s.run_stages(4); o.run_stages(4)
    phi = s.moments_group_major()[0]
    phi_ref_order = s.moments()[0]
ref = o.moments()[0]
print("oracle moments shape/flags", ref.shape, ref.flags.f_contiguous, "device ref-order", phi_ref_order.shape)
refc = np.ascontiguousarray(ref)
gmax = np.max(np.abs(refc), axis=1, keepdims=True)
pg = np.max(np.abs(phi - refc) / (np.abs(refc) + 1e-3 * gmax), axis=1)
pg2 = np.max(np.abs(phi_ref_order - ref) / (np.abs(ref) + 1e-3 * gmax), axis=1)
B = np.asarray(o.group_terms()["B"])
print("per-group err (group-major):", " ".join(f"{v:.1e}" for v in pg))
print("per-group err (ref order)  :", " ".join(f"{v:.1e}" for v in pg2))
print("B:", " ".join(f"{v:.1e}" for v in B))
o.close()
