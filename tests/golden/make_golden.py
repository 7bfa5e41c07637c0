#!/usr/bin/env python
"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle-1, oracle/_ref/ref_dump).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py

What it does
  1. copies the reference's INPUT files for the three BASELINE configs (prm/*.prm and the llnl
     table files -- data, not source) into tests/golden/prm/ ;
  2. writes extra small `.prm` cases into tests/golden/prm/ that exercise what the three configs do
     not: ts_method 1 and 2, every BC combination incl. the vacuum fall-through,
     V != 0 with the correction live, S4/S8 quadratures, ragged N, N=1, tabulated bounds/kappa;
  3. runs ref_dump on every case (and on `_steps<k>` variants with max_timesteps=k for per-step
     parity) and stores the dumped buffers as compressed .npz ;
  4. cross-checks oracle-2 (oracle/librt_oracle.so) against each dump and refuses to write a golden
     file it cannot reproduce bit-for-bit.
"""
import importlib
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402

prm = importlib.import_module("radiative-transfer_b200.prm")
grid = importlib.import_module("radiative-transfer_b200.grid")

REF_PRM = "/root/reference/prm"
PRM_DIR = os.path.join(HERE, "prm")
KEEP = ["mu", "wt", "e_edge", "kappa", "B_init", "psi", "ends", "prev_ends", "half_ends", "phi", "F", "phi_plus",
        "balance", "e_ave", "left_ends", "right_ends", "psi_source", "B", "dBdT", "dEB", "dsigEdE", "dkapEB",
        "kappa_edge"]


def write_table(name, values):
    with open(os.path.join(PRM_DIR, name), "w") as fh:
        fh.write("\n".join(repr(float(v)) for v in values) + "\n")


def extra_cases():
    P = prm.Parameters
    rng = np.random.Generator(np.random.PCG64(20261104))
    cases = {}
    cases["be_reflect_source"] = P(M=4, G=3, N=37, X=0.37, bc_left_indicator=2, bc_right_indicator=1, V=0.7,
                                   use_correction=True, ts_method=1, dt=1e-3, max_timesteps=6, include_validation=False,
                                   psi_source=np.array([[1.0, 2.0, 3.0], [0.5, 0.25, 0.125], [9.0, 9.0, 9.0], [9.0, 9.0, 9.0]]))
    cases["cn_vacuum_vacuum"] = P(M=4, G=2, N=40, X=0.8, bc_left_indicator=0, bc_right_indicator=0, V=0.5,
                                  use_correction=True, ts_method=2, dt=2e-4, max_timesteps=5, include_validation=False)
    cases["bdf2_reflect_equilib_s8"] = P(M=8, G=5, N=64, X=0.64, efirst=0.05, elast=20.0, bc_left_indicator=2,
                                         bc_right_indicator=1, use_mg_equilib=True, V=1.0, use_correction=True,
                                         ts_method=3, dt=1e-3, max_timesteps=3, include_validation=False)
    kap = 10.0 ** rng.uniform(-1.0, 2.0, size=6)
    bounds = np.concatenate([[0.0], np.geomspace(0.02, 25.0, 6)])
    write_table("syn6_group_bounds.txt", bounds)
    write_table("syn6_group_kappa.txt", kap)
    cases["bdf2_tables_ragged"] = P(M=4, G=6, N=33, X=0.5, bc_left_indicator=1, bc_right_indicator=1, V=2.0,
                                    use_correction=True, ts_method=3, dt=5e-4, max_timesteps=3,
                                    include_validation=False, have_group_bounds=True,
                                    filename_group_bounds="syn6_group_bounds.txt",
                                    have_group_absorption_opacities=True, filename_group_kappa="syn6_group_kappa.txt",
                                    psi_source=rng.uniform(0.0, 3.0, size=(4, 6)))
    cases["bdf2_nocorr_vacuum_left"] = P(M=2, G=2, N=16, X=1.0, bc_left_indicator=0, bc_right_indicator=1,
                                         use_correction=False, V=3.0, ts_method=3, dt=2e-3, max_timesteps=2,
                                         include_validation=False, psi_source=np.array([[0.3, 0.0], [0.0, 0.7]]))
    cases["bdf2_one_cell"] = P(M=2, G=1, N=1, X=0.01, bc_left_indicator=2, bc_right_indicator=1, V=1.0,
                               use_correction=True, ts_method=3, dt=1e-4, max_timesteps=2, include_validation=False,
                               psi_source=np.array([[2.5], [2.5]]))
    cases["bdf2_right_reflective_s4"] = P(M=4, G=2, N=9, X=0.3, bc_left_indicator=2, bc_right_indicator=2, V=0.25,
                                          use_correction=True, ts_method=3, dt=1e-3, max_timesteps=2,
                                          include_validation=False)
    return cases


def host_setup(p):
    mu, wt = grid.gauss_legendre(p.M)
    e = p.group_bounds if p.have_group_bounds else grid.group_edges(p.G, p.efirst, p.elast)
    kap = p.group_kappa if p.have_group_absorption_opacities else np.full(p.G, p.kappa_grey)
    return mu, wt, np.asarray(e, dtype=np.float64), np.asarray(kap, dtype=np.float64)


def crosscheck(name, p, ref):
    mu, wt, e, kap = host_setup(p)
    for k, v in (("mu", mu), ("wt", wt), ("e_edge", e), ("kappa", kap)):
        assert np.array_equal(v, ref[k]), f"{name}: host {k} differs from the reference"
    o = O.oracle_from_parameters(p, mu, wt, e, kap)
    o.solve()
    got = {"psi": o.psi(), "ends": o.ends(0), "prev_ends": o.ends(1), "half_ends": o.ends(2)}
    got["phi"], got["phi_plus"], got["F"] = o.moments()
    got["balance"] = o.balance()
    got["left_ends"], got["right_ends"] = o.group_ends()
    got.update(o.group_terms())
    got["psi_source"] = o.psi_source()
    for k, v in got.items():
        assert np.array_equal(v, ref[k]), f"{name}: oracle-2 {k} is not bit-identical to the reference"
    o.close()


def run_case(name, prm_path, steps_variants=()):
    p = prm.load(prm_path)
    ref = O.run_reference(prm_path, data_dir=PRM_DIR)
    crosscheck(name, p, ref)
    out = {k: ref[k] for k in KEEP}
    for k in steps_variants:
        pk = prm.load(prm_path)
        pk.max_timesteps = k
        tmp_prm = os.path.join(PRM_DIR, f"_tmp_{name}_steps{k}.prm")
        prm.dump(pk, tmp_prm)
        try:
            refk = O.run_reference(tmp_prm, data_dir=PRM_DIR)
        finally:
            os.remove(tmp_prm)
        crosscheck(f"{name}@{k}", pk, refk)
        out[f"psi_steps{k}"] = refk["psi"]
        out[f"ends_steps{k}"] = refk["ends"]
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(f"golden {name}: M={p.M} G={p.G} N={p.N} ts={p.ts_method} steps={p.max_timesteps} "
          f"ref solve {float(ref['solve_seconds'][0]):.3f}s  (+{len(steps_variants)} step snapshots)")


def main():
    os.chdir(HERE)  # prm.load's ../prm lookup falls back to the .prm's own directory
    os.makedirs(PRM_DIR, exist_ok=True)
    for f in ("single_group.prm", "multi_group_equilibrium.prm", "llnl_slab_test.prm",
              "llnl_slab_test_group_bounds.txt", "llnl_slab_test_group_kappa_a.txt"):
        shutil.copy(os.path.join(REF_PRM, f), os.path.join(PRM_DIR, f))
    O.build(force=True)
    run_case("single_group", os.path.join(PRM_DIR, "single_group.prm"), steps_variants=(1, 2, 10))
    run_case("multi_group_equilibrium", os.path.join(PRM_DIR, "multi_group_equilibrium.prm"), steps_variants=(1, 3))
    run_case("llnl_slab_test", os.path.join(PRM_DIR, "llnl_slab_test.prm"), steps_variants=(1,))
    for name, p in extra_cases().items():
        path = os.path.join(PRM_DIR, name + ".prm")
        prm.dump(p, path, extra_comment=f"({name}; written by tests/golden/make_golden.py)")
        steps = tuple(range(1, p.max_timesteps)) if p.ts_method == 3 else (1,)
        run_case(name, path, steps_variants=steps)


if __name__ == "__main__":
    main()
