// var_device.cuh -- device helpers shared by the cell-dependent kernels (var_tile_kernel.cu, var_step_kernel.cu).
#pragma once
#include "rt_common.cuh"
#include "sweep_device.cuh"

namespace rtk
{
// What coef_from_material (sweep_device.cuh) needs of the cell enters linearly except for ONE reciprocal: everything
// that depends on the row only (mu, tau, dt, dx, beta) is formed once per tile, the per-cell part is a handful of FMAs:
//    d = d0 + d1 sigma ;  1/det = 1/(d^2 + o^2) ;  id = d/det, io = o/det, alpha = io k, gamma = id k
//    qa = hq sigma B + qc2 cor2 + qc3 cor3 ;  qb = qb1 cor1 ;  c_uu = cu0 + cu1 sigma ;  p_uu = pu0 + pu1 sigma
// (same formulas, solver.cpp:324-334 BE, 411-425 CN, 498-512 BDF; rounding differs from the division form at 1e-16).
// 1/x for a normal, positive x: the hardware's 2^-23 seed (MUFU.RCP64H) and two Newton steps (2^-46, 2^-92): within an
// ulp of the correctly rounded reciprocal in 5 instructions, without the special-case paths of a full division
// (det = d^2 + o^2 is always a normal positive number).  __drcp_rn costs ~11 instructions: 12 % of the fused step (ncu).
__device__ __forceinline__ double rcp_normal(double x)
{
   double r;
   asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
   double e = fma(-x, r, 1.0);
   r = fma(r, e, r);
   e = fma(-x, r, 1.0);
   return fma(r, e, r);
}

template <int KIND, bool CORR>
struct RowConst
{
   double k, o, o2, d0, d1, hq, qc2, qc3, qb1, cu0, cu1, pu0, pu1, c_uv, p_uv, up1, up2;

   __device__ __forceinline__ RowConst(double mu, double dx, double tau, double dt, double beta)
   {
      const double amu = fabs(mu), ctau = C_LIGHT * tau;
      hq = 0.5 * ctau * dx;
      const double mb = mu * beta;
      qc2 = hq * mb; qc3 = -(hq * mb) * mb; qb1 = 0.5 * hq * mb;
      pu0 = pu1 = p_uv = up1 = up2 = 0.0;
      if (KIND == KIND_BE)
      {
         k = ctau * amu;
         d0 = 0.5 * (dx + k); d1 = 0.5 * ctau * dx;
         cu0 = 0.5 * dx; cu1 = 0.0; c_uv = 0.0;
      }
      else if (KIND == KIND_CN)
      {
         k = 0.5 * C_LIGHT * amu * tau;
         d0 = 0.5 * (dx + k); d1 = 0.25 * ctau * dx;
         cu0 = 0.5 * (dx - k); cu1 = -0.25 * ctau * dx; c_uv = -0.5 * k;
         up2 = k;
      }
      else
      {
         const double t1 = C_LIGHT * tau / 6.0; // t = t1 * sigma
         k = C_LIGHT * amu * dt / 6.0;          // the FULL dt
         d0 = 0.5 * (dx + k); d1 = 0.5 * dx * t1;
         cu0 = 0.5 * (dx - 4.0 * k); cu1 = -2.0 * dx * t1; c_uv = -2.0 * k;
         pu0 = -0.5 * k; pu1 = -0.5 * dx * t1; p_uv = -0.5 * k;
         up1 = 4.0 * k; up2 = k;
      }
      o = 0.5 * k;
      o2 = o * o;
   }

   // the source terms q = qa + qb (u+v) of a cell: they depend on tau, mu, beta and the material only, NOT on the stage
   // kind -- all four stages of a BDF2 step (tau = dt/2) share them
   __device__ __forceinline__ void source_terms(double srcB, double cor1, double cor2, double cor3, double &qa, double &qb) const
   {
      qa = hq * srcB;
      qb = 0.0;
      if (CORR)
      {
         qa = fma(qc2, cor2, fma(qc3, cor3, qa));
         qb = qb1 * cor1;
      }
   }

   // 1/det of a cell: the one expensive part of its constants (shared by the two BE stages of a BDF2 step)
   __device__ __forceinline__ double inv_det(double sig) const
   {
      const double d = fma(d1, sig, d0);
      return rcp_normal(fma(d, d, o2));
   }
   __device__ __forceinline__ RowCoef cell_q(double sig, double qa, double qb) const { return cell_from_inv(sig, inv_det(sig), qa, qb); }
   // stepper constants of a cell from sigma, 1/det and its (precomputed) source terms
   __device__ __forceinline__ RowCoef cell_from_inv(double sig, double inv, double qa, double qb) const
   {
      RowCoef c;
      const double d = fma(d1, sig, d0);
      c.id = d * inv;
      c.io = o * inv;
      c.alpha = c.io * k;
      c.gamma = c.id * k;
      c.qa = qa;
      c.qb = CORR ? qb : 0.0;
      c.c_uu = (KIND == KIND_BE) ? cu0 : fma(cu1, sig, cu0);
      c.c_vv = c.c_uu;
      c.c_uv = c_uv; c.c_vu = -c_uv;
      c.p_uu = (KIND == KIND_BDF) ? fma(pu1, sig, pu0) : 0.0;
      c.p_vv = c.p_uu;
      c.p_uv = p_uv; c.p_vu = -p_uv;
      c.up1 = up1; c.up2 = up2;
      c.kbc = (KIND == KIND_BE) ? 1.0 : (KIND == KIND_CN ? 2.0 : 6.0);
      c.inflow0 = 0.0;
      return c;
   }

   __device__ __forceinline__ RowCoef cell(double sig, double srcB, double cor1, double cor2, double cor3) const
   {
      RowCoef c;
      const double d = fma(d1, sig, d0);
      const double inv = rcp_normal(fma(d, d, o2));
      c.id = d * inv;
      c.io = o * inv;
      c.alpha = c.io * k;
      c.gamma = c.id * k;
      c.qa = hq * srcB;
      c.qb = 0.0;
      if (CORR)
      {
         c.qa = fma(qc2, cor2, fma(qc3, cor3, c.qa));
         c.qb = qb1 * cor1;
      }
      c.c_uu = (KIND == KIND_BE) ? cu0 : fma(cu1, sig, cu0);
      c.c_vv = c.c_uu;
      c.c_uv = c_uv; c.c_vu = -c_uv;
      c.p_uu = (KIND == KIND_BDF) ? fma(pu1, sig, pu0) : 0.0;
      c.p_vv = c.p_uu;
      c.p_uv = p_uv; c.p_vu = -p_uv;
      c.up1 = up1; c.up2 = up2;
      c.kbc = (KIND == KIND_BE) ? 1.0 : (KIND == KIND_CN ? 2.0 : 6.0);
      c.inflow0 = 0.0;
      return c;
   }
};

} // namespace rtk
