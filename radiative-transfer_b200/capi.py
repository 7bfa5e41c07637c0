"""ctypes binding of librt_cuda.so (the C ABI of include/rt_cuda.h).

This is the ONLY compute path of the package: if the shared library is missing or no CUDA device is
usable, calls raise -- there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_CUDA_LIB") or os.path.join(HERE, "librt_cuda.so")  # RT_CUDA_LIB: A/B builds of the same ABI
CSRC = os.path.join(HERE, "csrc")

_dp = C.POINTER(C.c_double)

#: every symbol include/rt_cuda.h declares
EXPORTS = (
    "rt_last_error", "rt_abi_version", "rt_create", "rt_destroy", "rt_get_plan", "rt_solve", "rt_run_stages",
    "rt_stages_done", "rt_total_stages", "rt_reset", "rt_get_psi", "rt_get_ends", "rt_moments",
    "rt_moments_group_major", "rt_balance", "rt_group_ends", "rt_get_group_terms", "rt_get_psi_source",
    "rt_validate", "rt_set_step_output", "rt_get_timing", "rt_reset_timing", "rt_planck_groups", "rt_bench_steps",
    "rt_host_alloc", "rt_host_free", "rt_multigroup_opacities", "rt_set_allreduce", "rt_get_temperature",
    "rt_planck_groups_noalloc", "rt_nccl_unique_id", "rt_set_nccl",
    "rt_multi_create", "rt_multi_destroy", "rt_multi_shards", "rt_multi_solve", "rt_multi_run_stages", "rt_multi_reset",
    "rt_multi_stages_done", "rt_multi_get_psi", "rt_multi_get_ends", "rt_multi_moments", "rt_multi_balance", "rt_multi_group_ends",
    "rt_multi_get_group_terms", "rt_multi_validate", "rt_multi_get_temperature", "rt_multi_bench_steps", "rt_multi_get_timing",
    "rt_multi_get_plan", "rt_correction_print", "rt_multi_correction_print", "rt_set_step_history", "rt_multi_set_step_history",
)

#: int fn(void *device_buf, size_t count, void *cuda_stream, void *user)  (include/rt_cuda.h: rt_allreduce_fn)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p)


class RtParams(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("G", C.c_int), ("group_begin", C.c_int), ("group_count", C.c_int),
        ("ts_method", C.c_int), ("max_timesteps", C.c_int), ("bc_left", C.c_int), ("bc_right", C.c_int),
        ("use_correction", C.c_int), ("use_mg_equilib", C.c_int), ("validate", C.c_int), ("device", C.c_int),
        ("groups_per_batch", C.c_int),
        ("dx", C.c_double), ("dt", C.c_double), ("rho", C.c_double), ("T", C.c_double), ("V", C.c_double),
        ("kappa_grey", C.c_double),
        ("mu", _dp), ("wt", _dp), ("e_edge", _dp), ("kappa", _dp), ("psi_source", _dp),
        ("rho_cell", _dp), ("T_cell", _dp), ("kappa_cell", _dp),  # extension (ABI 2): cell-dependent material fields
        ("cv", C.c_double),  # extension (ABI 3): material-temperature update when > 0
    ]


class RtTiming(C.Structure):
    _fields_ = [
        ("sweep_ms", C.c_double * 3), ("sweep_launches", C.c_longlong * 3), ("moments_ms", C.c_double),
        ("stages", C.c_longlong), ("unknown_stages", C.c_longlong), ("kernel_launches", C.c_longlong),
        ("h2d_bytes", C.c_longlong), ("d2h_bytes", C.c_longlong),
        ("step_ms", C.c_double), ("step_launches", C.c_longlong),
        ("tupdate_ms", C.c_double), ("tables_ms", C.c_double), ("tupdate_count", C.c_longlong),
    ]


class RtPlan(C.Structure):
    _fields_ = [
        ("groups_per_batch", C.c_int), ("n_batches", C.c_int), ("resident", C.c_int), ("tile_cells", C.c_int),
        ("tiles_per_row", C.c_int), ("sweep_grid", C.c_int), ("sweep_block", C.c_int),
        ("bytes_per_group", C.c_size_t), ("state_bytes", C.c_size_t), ("free_bytes", C.c_size_t),
        ("total_bytes", C.c_size_t),
        ("fused", C.c_int), ("step_tile_cells", C.c_int), ("step_block", C.c_int), ("step_chain", C.c_int),
        ("time_levels", C.c_int), ("step_kernel", C.c_int), ("var_scan_steps", C.c_int),
    ]


class RtError(RuntimeError):
    pass


_lib = None


def build(verbose: bool = False) -> str:
    """Compile librt_cuda.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", CSRC], stdout=out)
    return LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(f"{LIB_PATH} is missing: build it with `make -C {CSRC}` (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.rt_last_error.restype = C.c_char_p
        vp = C.c_void_p
        L.rt_create.argtypes = [C.POINTER(RtParams), C.POINTER(vp)]
        L.rt_destroy.argtypes = [vp]
        L.rt_get_plan.argtypes = [vp, C.POINTER(RtPlan)]
        L.rt_solve.argtypes = [vp]
        L.rt_run_stages.argtypes = [vp, C.c_int]
        L.rt_stages_done.argtypes = [vp, C.POINTER(C.c_int)]
        L.rt_total_stages.argtypes = [vp, C.POINTER(C.c_int)]
        L.rt_reset.argtypes = [vp]
        L.rt_get_psi.argtypes = [vp, _dp]
        L.rt_get_ends.argtypes = [vp, C.c_int, _dp]
        L.rt_moments.argtypes = [vp, _dp, _dp, _dp]
        L.rt_moments_group_major.argtypes = [vp, _dp, _dp, _dp]
        L.rt_balance.argtypes = [vp, _dp]
        L.rt_group_ends.argtypes = [vp, _dp, _dp]
        L.rt_get_group_terms.argtypes = [vp] + [_dp] * 6
        L.rt_get_psi_source.argtypes = [vp, _dp]
        L.rt_validate.argtypes = [vp, _dp, C.POINTER(C.c_int)]
        L.rt_set_step_output.argtypes = [vp, C.c_void_p, C.c_int]
        L.rt_get_timing.argtypes = [vp, C.POINTER(RtTiming)]
        L.rt_reset_timing.argtypes = [vp]
        L.rt_bench_steps.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.rt_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
        L.rt_host_free.argtypes = [C.c_void_p]
        L.rt_planck_groups.argtypes = [C.c_int, C.c_int, _dp, C.c_int, _dp, _dp, _dp, _dp]
        L.rt_planck_groups_noalloc.argtypes = [C.c_int, C.c_int, _dp, C.c_int, _dp, _dp, _dp]
        L.rt_multigroup_opacities.argtypes = [C.c_int, C.c_int, _dp, C.c_double, C.c_double, _dp, _dp, _dp, _dp]
        L.rt_set_allreduce.argtypes = [vp, ALLREDUCE_FN, C.c_void_p]
        L.rt_get_temperature.argtypes = [vp, _dp, _dp]
        ip = C.POINTER(C.c_int)
        L.rt_multi_create.argtypes = [C.POINTER(RtParams), C.c_int, ip, C.POINTER(vp)]
        L.rt_multi_destroy.argtypes = [vp]
        L.rt_multi_shards.argtypes = [vp, ip, ip, ip, ip]
        L.rt_multi_solve.argtypes = [vp]
        L.rt_multi_run_stages.argtypes = [vp, C.c_int]
        L.rt_multi_reset.argtypes = [vp]
        L.rt_multi_stages_done.argtypes = [vp, ip]
        L.rt_multi_get_psi.argtypes = [vp, _dp]
        L.rt_multi_get_ends.argtypes = [vp, C.c_int, _dp]
        L.rt_multi_moments.argtypes = [vp, _dp, _dp, _dp]
        L.rt_multi_balance.argtypes = [vp, _dp]
        L.rt_multi_group_ends.argtypes = [vp, _dp, _dp]
        L.rt_multi_get_group_terms.argtypes = [vp] + [_dp] * 6
        L.rt_multi_validate.argtypes = [vp, _dp, ip]
        L.rt_multi_get_temperature.argtypes = [vp, _dp, _dp]
        for f in (L.rt_set_step_history, L.rt_multi_set_step_history):
            f.argtypes = [vp, _dp, _dp, C.c_int]
        for f in (L.rt_correction_print, L.rt_multi_correction_print):
            f.argtypes = [vp, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]
        L.rt_multi_bench_steps.argtypes = [vp, C.c_int, C.c_int, _dp]
        L.rt_multi_get_timing.argtypes = [vp, C.c_int, C.POINTER(RtTiming)]
        L.rt_multi_get_plan.argtypes = [vp, C.c_int, C.POINTER(RtPlan)]
        L.rt_nccl_unique_id.argtypes = [C.c_void_p]
        L.rt_set_nccl.argtypes = [vp, C.c_void_p, C.c_int, C.c_int]
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise RtError(lib().rt_last_error().decode(errors="replace"))


def ptr(a: np.ndarray):
    return a.ctypes.data_as(_dp)
