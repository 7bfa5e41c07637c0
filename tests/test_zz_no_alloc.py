"""GPU test of rt_planck_groups_noalloc (include/rt_cuda.h): the Planck group integrals without any allocation.

The entry point has no counterpart in the reference; it was added for callers that
cannot call cudaMalloc.  Its scratch is one static device array that every call shares and that keeps the last call's
contents, so the check is made on TWO calls in a row -- the second one smaller and at other temperatures, where stale
contents of the first would show -- against the existing entry point rt_planck_groups (same kernels, allocating), at the
tolerance the Planck tests use (1e-9 relative, tests/parity_util.py RTOL; in practice the results are bit-identical).
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from parity_util import RTOL, rel_err

pytestmark = pytest.mark.gpu


def _noalloc(rt, e_edge, T, device=0):
    e_edge = np.ascontiguousarray(e_edge, dtype=np.float64)
    T = np.ascontiguousarray(np.atleast_1d(T), dtype=np.float64)
    G = e_edge.size - 1
    B, dB = np.zeros((T.size, G)), np.zeros((T.size, G))
    rt.capi.check(rt.capi.lib().rt_planck_groups_noalloc(device, G, rt.capi.ptr(e_edge), T.size, rt.capi.ptr(T), rt.capi.ptr(B),
                                                         rt.capi.ptr(dB)))
    return B, dB


def test_two_calls_in_a_row_match_the_allocating_entry_point(rt):
    from importlib import import_module
    planck_groups = import_module("radiative-transfer_b200.solver").planck_groups
    cases = [
        # (groups, temperatures): the first call fills a large part of the array, the second is smaller and hotter/colder,
        # the third is the reference's own shape (one temperature) and takes the untrimmed series
        (rt.grid.group_edges(64, 0.01, 30.0), 0.2 + 1.8 * np.linspace(0.0, 1.0, 700) ** 2),
        (rt.grid.group_edges(12, 0.05, 20.0), np.array([3.0, 0.05, 0.7, 1.0, 2.2])),
        (rt.grid.group_edges(32, 0.01, 30.0), np.array([1.0])),
    ]
    for e_edge, T in cases:
        B, dB = _noalloc(rt, e_edge, T)
        B_ref, dB_ref, _ = planck_groups(e_edge, T)
        assert np.isfinite(B).all() and np.isfinite(dB).all()
        assert rel_err(B, B_ref) <= RTOL, rel_err(B, B_ref)
        assert rel_err(dB, dB_ref) <= RTOL, rel_err(dB, dB_ref)
    # and the very same request twice gives the very same answer (the tables are cleared at the start of every call)
    e_edge, T = cases[1]
    a = _noalloc(rt, e_edge, T)
    b = _noalloc(rt, e_edge, T)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_sizes_beyond_the_static_array_are_rejected(rt):
    G, ncell = 256, 2048  # 3 * G * ncell doubles > RT_NOALLOC_MAX_DOUBLES = 2^20
    e_edge = rt.grid.group_edges(G, 0.01, 30.0)
    T = np.ones(ncell)
    B, dB = np.full((ncell, G), -1.0), np.full((ncell, G), -1.0)
    rc = rt.capi.lib().rt_planck_groups_noalloc(0, G, rt.capi.ptr(e_edge), ncell, rt.capi.ptr(T), rt.capi.ptr(B), rt.capi.ptr(dB))
    assert rc != 0
    msg = rt.capi.lib().rt_last_error().decode()
    assert "RT_NOALLOC_MAX_DOUBLES" in msg and "scratch" in msg, msg
    assert (B == -1.0).all() and (dB == -1.0).all()  # nothing was computed
    # the largest documented shape still fits
    Bk, _ = _noalloc(rt, e_edge, np.linspace(0.5, 1.5, 1024))
    assert np.isfinite(Bk).all() and (Bk[:, :-1] > 0).all()
