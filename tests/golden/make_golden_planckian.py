#!/usr/bin/env python
"""Generate tests/golden/planckian.json from the reference's own stand-alone test program tests/test_planckian.cpp
(compiled UNMODIFIED by `make -C oracle ref` -> oracle/_ref/test_planckian, driven through stdin).

That program is the reference author's second statement of the whole Correction pipeline -- Planck group integrals,
multigroup opacities (correction.cpp:66-97, never called by the solver itself), group-edge opacities, dEB, dkapEdE,
dkapEB, cor1..3 -- and prints every table with 6 significant digits.  Run in the build container only."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O  # noqa: E402

CASES = {
    # name: (M, X, N, G, efirst, elast, T, kappa_grey)
    "g4_T1": (2, 1.0, 10, 4, 0.5, 10.0, 1.0, 1.0),
    "g12_T0p3": (4, 2.0, 8, 12, 0.05, 20.0, 0.3, 2.5),
    "g32_T2": (2, 1.0, 4, 32, 0.01, 60.0, 2.0, 0.7),
    "g64_T1": (8, 1.0, 4, 64, 0.01, 30.0, 1.0, 1.0),
}

out = {}
for name, args in CASES.items():
    r = O.run_test_planckian(*args)
    out[name] = {"inputs": dict(zip(("M", "X", "N", "G", "efirst", "elast", "T", "kappa_grey"), args)),
                 "tables": {k: [float(x) for x in v] for k, v in r.items()}}
    print(name, "ok", len(r["B"]), "groups")
with open(os.path.join(HERE, "planckian.json"), "w") as fh:
    json.dump(out, fh, indent=1)
