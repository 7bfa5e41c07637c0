"""CPU tests of the drop-in boundary: librt_cuda.so loads, exports every symbol include/rt_cuda.h declares, its
structs have the layout the bindings assume, and -- with no GPU -- it fails loudly instead of falling back."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rt_cuda.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", src)))


def test_header_and_binding_agree(rt):
    assert declared_functions() == sorted(rt.capi.EXPORTS)


def test_library_exports_every_declared_symbol(rt):
    if not os.path.exists(rt.capi.LIB_PATH):
        rt.capi.build()
    lib = rt.capi.lib()
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/rt_cuda.h but not exported"
    assert lib.rt_abi_version() == 5
    out = subprocess.check_output(["nm", "-D", "--defined-only", rt.capi.LIB_PATH], text=True)
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    assert set(declared_functions()) <= exported


def test_library_is_sm100a_only(rt):
    out = subprocess.run(["cuobjdump", "-lelf", rt.capi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    archs = set(re.findall(r"sm_(\d+a?)", out.stdout))
    assert archs == {"100a"}, archs


def test_sweep_kernels_have_no_divergence_fallbacks(rt):
    """The scans of the sweep kernels are warp shuffles in loops without a CTA barrier.  When ptxas cannot prove the
    warp converged it wraps every shuffle in a BRA.DIV check with an out-of-line WARPSYNC.COLLECTIVE path (and pays two
    register moves per 64-bit shuffle in the main path).  None of the library's kernels may contain one."""
    out = subprocess.run(["cuobjdump", "-sass", rt.capi.LIB_PATH], capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump not available")
    per_kernel, name = {}, None
    for ln in out.stdout.splitlines():
        m = re.search(r"Function : (\S+)", ln)
        if m:
            name = m.group(1)
        elif "BRA.DIV" in ln and name:
            per_kernel[name] = per_kernel.get(name, 0) + 1
    hot = {k: v for k, v in per_kernel.items() if re.search(r"(sweep|step|var_tile)_?\w*kernel", k)}
    assert not hot, hot


def test_struct_layout_matches_header(rt, tmp_path):
    """compile a C program against include/rt_cuda.h that prints sizeof/offsetof; compare with the ctypes mirrors"""
    prog = tmp_path / "layout.c"
    prog.write_text(
        '#include <stdio.h>\n#include <stddef.h>\n#include "rt_cuda.h"\n'
        "int main(void){printf(\"%zu %zu %zu %zu %zu %zu %zu %zu\\n\", sizeof(rt_params), offsetof(rt_params, dx),"
        " offsetof(rt_params, mu), offsetof(rt_params, psi_source), sizeof(rt_timing), offsetof(rt_timing, stages),"
        " sizeof(rt_plan), offsetof(rt_plan, bytes_per_group)); return 0;}\n")
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    P, T, L = rt.capi.RtParams, rt.capi.RtTiming, rt.capi.RtPlan
    want = [C.sizeof(P), P.dx.offset, P.mu.offset, P.psi_source.offset, C.sizeof(T), T.stages.offset, C.sizeof(L),
            L.bytes_per_group.offset]
    assert got == want


def _no_gpu():
    try:
        return subprocess.run(["nvidia-smi", "-L"], capture_output=True).returncode != 0
    except OSError:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful on a box without a GPU")
def test_no_cpu_fallback(rt):
    """the product path must fail loudly when CUDA is unusable"""
    p = rt.Parameters(M=2, G=1, N=10, max_timesteps=1)
    with pytest.raises(rt.capi.RtError, match="no CUDA device"):
        rt.Solver(p)
    with pytest.raises(rt.capi.RtError, match="no CUDA device"):
        rt.planck_groups(np.array([0.0, 1.0, 2.0]), 1.0)


def test_bad_parameters_are_rejected_before_touching_the_device(rt):
    lib = rt.capi.lib()
    h = C.c_void_p()
    p = rt.capi.RtParams()
    p.M, p.N, p.G, p.group_count = 3, 10, 1, 1  # odd M
    assert lib.rt_create(C.byref(p), C.byref(h)) != 0
    assert b"M must be even" in lib.rt_last_error()
    assert lib.rt_create(None, C.byref(h)) != 0
    assert lib.rt_solve(None) != 0 and b"null handle" in lib.rt_last_error()


def test_product_never_touches_the_oracle():
    """nothing under radiative-transfer_b200/ or include/rt_cuda.h may import, link or execute oracle/"""
    pkg = os.path.join(ROOT, "radiative-transfer_b200")
    bad = []
    for base, _, files in os.walk(pkg):
        if os.sep + "build" in base or "__pycache__" in base:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(base, f), errors="replace").read()
                if re.search(r"oracle_py|librt_oracle|rt_oracle\.h|rto_[a-z]+\(", txt):
                    bad.append(os.path.join(base, f))
    assert not bad, bad
    out = subprocess.run(["ldd", os.path.join(pkg, "librt_cuda.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out
