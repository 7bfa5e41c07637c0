"""Host-side setup that stays on the CPU: angular quadrature and the energy-group grid.

These are O(M^2) / O(G) one-time computations whose results (mu, wt, e_edge) are INPUTS of
the CUDA path (rt_params in include/rt_cuda.h).  Every floating-point operation is done in the
reference's order with the same libm calls, so the arrays are bit-identical to the reference's
(checked against the golden dumps in tests/test_host_setup.py).
"""
from __future__ import annotations

import math

import numpy as np

# reference include/Constants.h:15-16 -- PI is truncated on purpose
PI = 3.1415926546
FOUR_PI = 4.0 * PI
SPEED_OF_LIGHT = 299.79245800
RADIATION_CONSTANT_A = 1.3653104e-2


def gauss_legendre(num_points: int, norm: float = FOUR_PI, tolerance: float = 1.0e-12):
    """Gauss-Legendre nodes (ascending mu) and weights normalised to `norm`.

    Reference: GLQuad::build, src/GLQuad.cpp:4-44 (Newton iteration on the Legendre three-term
    recurrence, started from cos(PI*(i-1/4)/(n+1/2)), stopped when |dz| <= 1e-12).
    """
    if num_points <= 0 or norm <= 0.0:
        raise AssertionError("num_points > 0 and norm > 0.0 required")
    n = num_points
    mu = np.zeros(n)
    wt = np.zeros(n)
    xm, xl = 0.0, 1.0  # mid-point and half-length of [-1, 1]
    dnp = float(n)
    for i in range(1, (n + 1) // 2 + 1):
        z = math.cos(PI * (float(i) - 0.25) / (dnp + 0.5))
        while True:
            p1, p2 = 1.0, 0.0
            for j in range(1, n + 1):
                dj = float(j)
                p3 = p2
                p2 = p1
                p1 = ((2.0 * dj - 1.0) * z * p2 - (dj - 1.0) * p3) / dj
            pp = dnp * (z * p1 - p2) / (z * z - 1.0)
            z1 = z
            z = z1 - p1 / pp
            if not (abs(z - z1) > tolerance):
                break
        mu[i - 1] = xm - xl * z
        mu[n - i] = xm + xl * z
        wt[i - 1] = norm * xl / ((1.0 - z * z) * pp * pp)
        wt[n - i] = wt[i - 1]
    return mu, wt


def group_edges(G: int, efirst: float, elast: float) -> np.ndarray:
    """Log-spaced group edges e_0=0, e_1=efirst, e_{g+1}=e_g*logfac (src/solver.cpp:6-19)."""
    e = np.zeros(G + 1)
    e[1] = efirst
    if G > 1:
        logfac = math.exp((math.log(elast) - math.log(efirst)) / (G - 1.0))
        for g in range(1, G):
            e[g + 1] = e[g] * logfac
    return e


def group_averages(e_edge: np.ndarray):
    """e_ave, de_ave (src/solver.cpp:22-32)."""
    e_edge = np.asarray(e_edge, dtype=np.float64)
    return 0.5 * (e_edge[:-1] + e_edge[1:]), e_edge[1:] - e_edge[:-1]
