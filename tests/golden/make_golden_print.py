#!/usr/bin/env python
"""Generate tests/golden/correction_print_<case>.txt: the text the UNMODIFIED reference's Correction::Print()
(/root/reference/src/correction.cpp:404-603) writes for the reference's own .prm files.

No driver of the reference calls Print(), so oracle/ref_dump.cpp (the oracle-1 harness, which already lifts access
control for its dumps) has a `--print` mode: construct rt::Solver as main.cc does, one compute_correction, Print().
Run in the build container (needs /root/reference and `make -C oracle ref`); the texts are committed because
/root/reference does not exist on the GPU box.

    python tests/golden/make_golden_print.py
"""
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_DUMP = os.path.join(ROOT, "oracle", "_ref", "ref_dump")
PRM = os.path.join(HERE, "prm")  # the committed copies of the reference's inputs
CASES = ("single_group", "multi_group_equilibrium", "llnl_slab_test")
BEGIN, END = "@@@ Correction::Print begin", "@@@ Correction::Print end"


def reference_print(prm_file: str) -> str:
    with tempfile.TemporaryDirectory() as tmp:
        os.symlink(PRM, os.path.join(tmp, "prm"))  # the reference reads ../prm/<data file> relative to its cwd
        run = os.path.join(tmp, "run")
        os.mkdir(run)
        out = subprocess.run([REF_DUMP, prm_file, "--print"], cwd=run, capture_output=True, text=True, check=True).stdout
    a, b = out.index(BEGIN) + len(BEGIN) + 1, out.index(END)
    return out[a:b]


if __name__ == "__main__":
    for case in CASES:
        text = reference_print(os.path.join(PRM, case + ".prm"))
        with open(os.path.join(HERE, f"correction_print_{case}.txt"), "w") as f:
            f.write(text)
        print(case, len(text.splitlines()), "lines")
