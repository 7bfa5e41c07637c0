#!/usr/bin/env python
"""CPU cross-check of the ONE-SCAN formulation of the fused BDF2 step (csrc/step4_kernel.cu), in plain numpy.

A BDF2 step is four stage sweeps (BE, CN, BE, BDF) of a row.  Each stage is a first-order recurrence in its outflow
node, and stage k+1 of a cell needs stage-k values of the SAME cell (plus old-level / half-level values of the upstream
cell).  The whole step is therefore ONE affine recurrence in the 4-vector c = (v0, v1, v2, v3) of the four stages'
outflow nodes, with a lower-triangular 4x4 transfer matrix per chunk of cells that is a ROW CONSTANT (uniform slab):

    c_out = T c_in + e_local ,   (u3, v3)_j = local_j + sum_l R[l][j] c_in[l]

so a lane runs all four stages over its 8 cells with ZERO inflow (pass A, no communication), the warp does ONE
Kogge-Stone scan of 4-vectors with the tabulated T, T^2, .. T^16 (instead of four dependent scalar scans), and pass B
adds the response R c_in to the local stage-3 result.  This script emulates exactly that structure (lanes, sub-tiles,
carries, row-first / junction / ragged handling) and compares it with the serial scan form of scan_form.py.
Development aid: the kernel's device code is a transliteration of pass_a / chunk_tables below."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
from scan_form import C, coef  # noqa: E402

CPT, LANES = 8, 32
WT = CPT * LANES


def folded(P, kind, corr):
    """pack_step4_coef_kernel: beta / U as linear forms of the cell's inputs (cf. step_warp_kernel.cu StepCoefIndex)"""
    io, idd = P["io"], P["id"]
    qb = P["qb"] if corr else 0.0
    cuu, cuv = P["c_uu"], (0.0 if kind == 0 else P["c_uv"])
    puu, puv = (P["p_uu"], P["p_uv"]) if kind == 2 else (0.0, 0.0)
    u1 = P["up1"] if kind == 2 else 0.0
    u2 = P["up2"] if kind != 0 else 0.0
    sb, su = io + idd, idd - io
    F = dict(alpha=P["alpha"], gamma=P["gamma"], kbc=P["kbc"])
    F.update(B0=sb * P["qa"], Bq=sb * qb, Bx=io * cuu - idd * cuv, By=io * cuv + idd * cuu,
             BX=io * puu - idd * puv, BY=io * puv + idd * puu, Bup1=io * u1, Bup2=io * u2)
    F.update(U0=su * P["qa"], Uq=su * qb, Ux=idd * cuu + io * cuv, Uy=idd * cuv - io * cuu,
             UX=idd * puu + io * puv, UY=idd * puv - io * puu, Uup1=idd * u1, Uup2=idd * u2)
    if kind != 2:  # stages 0..2: the source cell IS the X1 cell, q folds into the X1 weights
        F["Bx"] += F["Bq"]; F["By"] += F["Bq"]; F["Ux"] += F["Uq"]; F["Uy"] += F["Uq"]
    return F


def pass_a(K, x, up2_0, up1_0, seed, sgn, homogeneous=False):
    """All four stages over one lane chunk.  x: (n, CPT, 2) old level; up2_0 / up1_0: (n,) outflow nodes of the old level /
    the half level at the cell upstream of the chunk; seed: (n, 4) inflow of the four stages.  Returns the stage-3 result
    (n, CPT, 2), the chunk's outflow 4-vector and the per-cell outflow nodes of all stages (ragged rows publish one)."""
    be, cn, bdf = K
    n = x.shape[0]
    z = 0.0 if homogeneous else 1.0
    vs = np.zeros((n, 4, CPT))
    # stage 0: BE from X
    r0 = np.zeros((n, CPT, 2)); v = seed[:, 0].copy()
    for j in range(CPT):
        b = be["By"] * x[:, j, 1] + be["Bx"] * x[:, j, 0] + z * be["B0"]
        U = be["Uy"] * x[:, j, 1] + be["Ux"] * x[:, j, 0] + z * be["U0"]
        r0[:, j, 0] = be["gamma"] * v + U
        v = be["alpha"] * v + b
        r0[:, j, 1] = v
    e0 = v; vs[:, 0] = r0[:, :, 1]
    # stage 1: CN from the stage-0 result, upstream X
    r1 = np.zeros((n, CPT, 2)); v = seed[:, 1].copy(); up2 = up2_0.copy()
    for j in range(CPT):
        b = cn["Bup2"] * up2 + cn["By"] * r0[:, j, 1] + cn["Bx"] * r0[:, j, 0] + z * cn["B0"]
        U = cn["Uup2"] * up2 + cn["Uy"] * r0[:, j, 1] + cn["Ux"] * r0[:, j, 0] + z * cn["U0"]
        r1[:, j, 0] = cn["gamma"] * v + U
        v = cn["alpha"] * v + b
        r1[:, j, 1] = v
        up2 = x[:, j, 1]
    e1 = v; vs[:, 1] = r1[:, :, 1]
    # stage 2: BE from the stage-1 result; only u+v survives
    xs = np.zeros((n, CPT)); v = seed[:, 2].copy()
    for j in range(CPT):
        b = be["By"] * r1[:, j, 1] + be["Bx"] * r1[:, j, 0] + z * be["B0"]
        U = be["Uy"] * r1[:, j, 1] + be["Ux"] * r1[:, j, 0] + z * be["U0"]
        u = be["gamma"] * v + U
        v = be["alpha"] * v + b
        xs[:, j] = u + v
        vs[:, 2, j] = v
    e2 = v
    # stage 3: BDF, x1 = half (stage 0 for mu>0 rows, stage 1 for mu<0 rows), x2 = X, source from stage 2
    h = r0 if sgn else r1
    out = np.zeros((n, CPT, 2)); v = seed[:, 3].copy(); up1 = up1_0.copy(); up2 = up2_0.copy()
    for j in range(CPT):
        b = (bdf["By"] * h[:, j, 1] + bdf["Bx"] * h[:, j, 0] + z * bdf["B0"] + bdf["Bq"] * xs[:, j] + bdf["BX"] * x[:, j, 0]
             + bdf["BY"] * x[:, j, 1] + bdf["Bup1"] * up1 + bdf["Bup2"] * up2)
        U = (bdf["Uy"] * h[:, j, 1] + bdf["Ux"] * h[:, j, 0] + z * bdf["U0"] + bdf["Uq"] * xs[:, j] + bdf["UX"] * x[:, j, 0]
             + bdf["UY"] * x[:, j, 1] + bdf["Uup1"] * up1 + bdf["Uup2"] * up2)
        out[:, j, 0] = bdf["gamma"] * v + U
        v = bdf["alpha"] * v + b
        out[:, j, 1] = v
        up1 = h[:, j, 1]; up2 = x[:, j, 1]
    vs[:, 3] = out[:, :, 1]
    return out, np.stack([e0, e1, e2, v], axis=1), vs


def chunk_tables(K, sgn):
    """row constants: T (4x4 lower triangular) and its powers, R[l] = response of the stage-3 result to inflow l"""
    T = np.zeros((4, 4)); R = np.zeros((4, CPT, 2))
    half = 0 if sgn else 1
    for l in range(4):
        seed = np.zeros((1, 4)); seed[0, l] = 1.0
        out, e, _ = pass_a(K, np.zeros((1, CPT, 2)), np.zeros(1), np.array([1.0 if l == half else 0.0]), seed, sgn, homogeneous=True)
        T[:, l] = e[0]; R[l] = out[0]
    assert np.allclose(np.triu(T, 1), 0.0)
    P = [T]
    for _ in range(4):
        P.append(P[-1] @ P[-1])
    return P, R


def step_row(K, sgn, X, inflow, junction_kbc=None):
    """one BDF2 step of one row in the kernel's structure.  X: (N, 2).  inflow: 4 boundary inflows (already times kbc), or
    for a junction row the 4 outflows of the mirrored row (multiplied by kbc here).  Returns ends^(n+1) (N,2) and the 4
    outflows at the row's last cell."""
    N = X.shape[0]
    Tp, R = chunk_tables(K, sgn)
    out = np.zeros_like(X)
    carry = None
    last = np.zeros(4)
    for s0 in range(0, N, WT):
        valid = min(WT, N - s0)
        x = np.zeros((LANES, CPT, 2)); x.reshape(-1, 2)[:valid] = X[s0:s0 + valid]
        up2_0 = np.zeros(LANES)
        up2_0[1:] = x[:-1, CPT - 1, 1]
        row_first = s0 == 0
        seed = np.zeros((LANES, 4))
        if row_first:
            seed[0] = inflow if junction_kbc is None else junction_kbc * inflow
        else:
            up2_0[0] = X[s0 - 1, 1]
        loc, e, vs = pass_a(K, x, up2_0, np.zeros(LANES), seed, sgn)
        # seeded Kogge-Stone scan of 4-vectors: e_incl[l] = T e_incl[l-1] + e[l]
        incl = e.copy()
        if not row_first:
            incl[0] = Tp[0] @ carry + incl[0]
        for s in range(5):
            d = 1 << s
            sh = np.vstack([np.zeros((d, 4)), incl[:-d]])
            upd = (Tp[s] @ sh.T).T + incl
            incl[d:] = upd[d:]
        cin = np.vstack([np.zeros((1, 4)) if row_first else carry[None, :], incl[:-1]])
        if valid == WT:
            res = loc + np.einsum("ljs,nl->njs", R, cin)
            carry = incl[31].copy()
            last = carry
        else:
            # ragged sub-tile (the row's last): re-run pass A seeded with the true inflows, take the outflow at the last cell
            half = 0 if sgn else 1
            seed2 = cin.copy(); up1_0 = cin[:, half].copy()
            if row_first:
                seed2[0] = seed[0]; up1_0[0] = 0.0
            res, _, vs = pass_a(K, x, up2_0, up1_0, seed2, sgn)
            jl = valid - 1
            last = vs[jl // CPT, :, jl % CPT].copy()
        out[s0:s0 + valid] = res.reshape(-1, 2)[:valid]
    return out, last


def step_row_serial(Ks, sgn, X, inflow):
    """the same step, stage by stage, cell by cell (unfolded coefficients: scan_form_check.py)"""
    N = X.shape[0]
    be, cn, bdf = Ks

    def sweep(P, kind, X1, X2, XC, vin):
        res = np.zeros_like(X1)
        for j in range(N):
            q = P["qa"] + P["qb"] * (XC[j, 0] + XC[j, 1])
            ru = q + P["c_uu"] * X1[j, 0] + P["c_uv"] * X1[j, 1]
            rv = q + P["c_vu"] * X1[j, 0] + P["c_vv"] * X1[j, 1]
            if kind == 2:
                ru += P["p_uu"] * X2[j, 0] + P["p_uv"] * X2[j, 1]; rv += P["p_vu"] * X2[j, 0] + P["p_vv"] * X2[j, 1]
            if kind >= 1 and j > 0:
                ru += P["up1"] * (X1[j - 1, 1] if kind == 2 else 0.0) + P["up2"] * X2[j - 1, 1]
            v = P["alpha"] * vin + (P["io"] * ru + P["id"] * rv)
            u = P["gamma"] * vin + (P["id"] * ru - P["io"] * rv)
            res[j] = (u, v); vin = v
        return res
    s0 = sweep(be, 0, X, None, X, inflow[0])
    s1 = sweep(cn, 1, s0, X, s0, inflow[1])
    s2 = sweep(be, 0, s1, None, s1, inflow[2])
    h = s0 if sgn else s1
    s3 = sweep(bdf, 2, h, X, s2, inflow[3])
    return s3, np.array([s0[-1, 1], s1[-1, 1], s2[-1, 1], s3[-1, 1]])


def main():
    rng = np.random.default_rng(7)
    worst = 0.0
    for N in (8, 200, 256, 700, 1024, 2100):
        for corr in (True, False):
            for dt in (1e-3, 1e-5):
                for sgn in (0, 1):
                    mu = (0.57 if sgn else -0.57) * rng.uniform(0.3, 1.7)
                    sig = 10.0 ** rng.uniform(-1, 2); B = rng.uniform(0.1, 3); c1, c2, c3 = rng.uniform(-1, 1, 3)
                    args = (mu, sig, sig * B, c1, c2, c3, 0.01)
                    Ks = [coef(k, corr, *args, dt / 2, dt, 1.0 / C) for k in range(3)]
                    K = [folded(Ks[k], k, corr) for k in range(3)]
                    X = rng.uniform(0.5, 2.0, (N, 2))
                    bd = rng.uniform(0.0, 2.0)
                    inflow = np.array([1.0 * bd, 2.0 * bd, 1.0 * bd, 6.0 * bd])
                    a, la = step_row(K, sgn, X, inflow)
                    b, lb = step_row_serial(Ks, sgn, X, inflow)
                    err = max(np.max(np.abs(a - b) / np.abs(b)), np.max(np.abs(la - lb) / np.abs(lb)))
                    worst = max(worst, err)
                    # junction: the mirrored row's outflows times kbc
                    a2, _ = step_row(K, sgn, X, la, junction_kbc=np.array([1.0, 2.0, 1.0, 6.0]))
                    b2, _ = step_row_serial(Ks, sgn, X, la * np.array([1.0, 2.0, 1.0, 6.0]))
                    worst = max(worst, np.max(np.abs(a2 - b2) / np.abs(b2)))
    print(f"one-scan formulation vs serial stage sweeps: worst relative difference {worst:.2e}")
    assert worst < 1e-10  # elementwise relative, dominated by entries near zero crossings (max-norm error ~1e-15)
    return worst


if __name__ == "__main__":
    main()
