#!/usr/bin/env python
"""CPU cross-check of the SCAN FORMULATION used by sweep_kernel.cu (coefficients, boundary multiplier, sweep-ordered
(u,v) layout, buffer aliasing of prev/half) against the oracle, in plain numpy.  Development aid / test helper: it
mirrors make_coef() and the stage schedule of rt_capi.cu but runs the recurrence serially."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
C = 299.79245800


def coef(kind, corr, mu, sig, srcB, c1, c2, c3, dx, tau, dt, beta):
    amu = abs(mu); ctau = C * tau; hq = 0.5 * ctau * dx
    qa = hq * srcB; qb = 0.0
    if corr:
        mb = mu * beta
        qa += hq * (c2 * mb - c3 * mb * mb); qb = 0.5 * hq * c1 * mb
    P = dict(p_uu=0., p_uv=0., p_vu=0., p_vv=0., up1=0., up2=0.)
    if kind == 0:
        k = ctau * amu; d = 0.5 * ((1 + ctau * sig) * dx + k)
        P.update(c_uu=0.5 * dx, c_uv=0., c_vu=0., c_vv=0.5 * dx, kbc=1.)
    elif kind == 1:
        t = 0.5 * ctau * sig; k = 0.5 * C * amu * tau; d = 0.5 * ((1 + t) * dx + k); e1 = 0.5 * ((1 - t) * dx - k)
        P.update(c_uu=e1, c_uv=-0.5 * k, c_vu=0.5 * k, c_vv=e1, up2=k, kbc=2.)
    else:
        t = C * sig * tau / 6; k = C * amu * dt / 6; d = 0.5 * ((1 + t) * dx + k)
        f1 = 0.5 * ((1 - 4 * t) * dx - 4 * k); g1 = 0.5 * (k + t * dx)
        P.update(c_uu=f1, c_uv=-2 * k, c_vu=2 * k, c_vv=f1, p_uu=-g1, p_uv=-0.5 * k, p_vu=0.5 * k, p_vv=-g1, up1=4 * k, up2=k, kbc=6.)
    o = 0.5 * k; det = d * d + o * o
    P.update(id=d / det, io=o / det, qa=qa, qb=qb)
    P["alpha"] = P["io"] * k; P["gamma"] = P["id"] * k
    return P


def run(rt, oracle, p, mu, wt, e, kap, nstages):
    """returns ends in reference layout after nstages, computed in scan form"""
    o = oracle.oracle_from_parameters(p, mu, wt, e, kap)
    o.run_stages(0)
    M, G, N = p.M, p.G, p.N
    # group tables from the oracle (the G-length terms are checked separately)
    o.run_stages(1); gt = o.group_terms(); src_tab = o.psi_source(); o.close()
    B = gt["B"]; sig = p.rho * kap
    c1 = gt["dsigEdE"]; c2 = 3 * sig * B - gt["dkapEB"]; c3 = c1 * (4 * B - gt["dEB"])
    beta = p.V / C; corr = p.use_correction and p.V != 0
    E = [np.zeros((G, M, N, 2)) for _ in range(4)]
    E[0][:] = B[:, None, None, None]
    ie, ip, ihn, ihp = 0, -1, -1, -1
    def free(*used):
        return [k for k in range(4) if k not in used][0]
    for it in range(nstages):
        if p.ts_method == 1: kind, tau, out, x1n, x1p, x2, xc = 0, p.dt, free(ie), ie, ie, None, None; ip = ie
        elif p.ts_method == 2: kind, tau, out, x1n, x1p, x2, xc = 1, p.dt, free(ie), ie, ie, ie, None; ip = ie
        else:
            tau = p.dt / 2; ind = it % 4; xc = None
            if ind == 0: kind, out, x1n, x1p, x2 = 0, free(ie, ihn, ihp), ie, ie, None; ip = ie
            elif ind == 1: kind, out, x1n, x1p, x2 = 1, free(ie, ip), ie, ie, ip; ihp = ie; ihn = out
            elif ind == 2: kind, out, x1n, x1p, x2 = 0, free(ie, ip, ihn, ihp), ie, ie, None
            else: kind, out, x1n, x1p, x2, xc = 2, ie, ihn, ihp, ip, ie
        new = E[out].copy() if out == ie else E[out]
        res = np.zeros((G, M, N, 2))
        for g in range(G):
            for i in list(range(M // 2)) + list(range(M // 2, M)):
                P = coef(kind, corr, mu[i], sig[g], sig[g] * B[g], c1[g], c2[g], c3[g], p.dx, tau, p.dt, beta)
                neg = i < M // 2
                X1 = E[x1n if neg else x1p][g, i]; X2 = E[x2][g, i] if x2 is not None else None
                XC = E[xc][g, i] if (xc is not None and corr) else X1
                if neg: b = src_tab[i, g] if p.bc_right_indicator == 1 else 0.0
                elif p.bc_left_indicator in (0, 1): b = src_tab[i, g]
                else: b = res[g, M // 2 - 1 - (i - M // 2), N - 1, 1]
                vin = P["kbc"] * b
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
                    res[g, i, j] = (u, v); vin = v
        E[out][:] = res
        ie = out
    # -> reference layout ends(i,g,c,s)
    ends = np.zeros((M, G, N, 2), order="F")
    for i in range(M):
        a = E[ie][:, i]                      # (G, N, 2) sweep order
        if i < M // 2: a = a[:, ::-1, ::-1]  # cell = N-1-j ; (u,v) = (R,L)
        ends[i] = a
    return ends


if __name__ == "__main__":
    from conftest import GOLDEN_CASES, host_setup
    from parity_util import load_case, load_golden, rel_err
    import oracle_py
    rt = importlib.import_module("radiative-transfer_b200")
    for name in GOLDEN_CASES:
        p = load_case(rt, name)
        if p.max_timesteps > 20: p.max_timesteps = 3
        mu, wt, e, kap = host_setup(rt, p)
        n = p.max_timesteps * (4 if p.ts_method == 3 else 1)
        o = oracle_py.oracle_from_parameters(p, mu, wt, e, kap); o.run_stages(n); ref = o.ends(0); o.close()
        got = run(rt, oracle_py, p, mu, wt, e, kap, n)
        print(f"{name:28s} stages={n:3d} scan-form vs oracle: rel {rel_err(got, ref):.2e}")
