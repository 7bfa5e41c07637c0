// group_terms.cu -- G-length device kernels: Planck group integrals and the per-group pieces of the
// reference's Correction class.  fp64 throughout; sm_100a.
//
// Reference being reproduced (results, not code):
//   Planck::get_Planck / integrate_B / integrate_dBdT        src/Planck.cpp:50-229, include/Planck.h:86-125
//   Correction::generate_planck_integrals                    src/correction.cpp:25-36
//   Correction::compute_group_edge_opacities                 src/correction.cpp:125-139
//   Correction::compute_components_of_correction_source, pf  src/correction.cpp:11-22,162-249
//   Correction::compute_correction_terms                     src/correction.cpp:328-340
//   Correction::validate_planck_integrals / validate_emission src/correction.cpp:39-63,100-122
//   Solver::computeEquilibriumSources                        src/solver.cpp:287-315
#include "rt_common.cuh"
#include "group_terms.cuh"

namespace rtk
{
__constant__ double c_gauss_pt[12];
__constant__ double c_gauss_w[12];

__device__ __forceinline__ double rad_a_long()
{
   // RADIATION_CONSTANT_A_LONG = 8 pi^5 k^4 / (15 h^3 c^3) with k = 1 (Constants.h:21-22)
   const double pi2 = PI_TRUNC * PI_TRUNC;
   return (8.0 * (pi2 * pi2 * PI_TRUNC)) / (15.0 * (H_KEV_SH * H_KEV_SH * H_KEV_SH) * (C_LIGHT * C_LIGHT * C_LIGHT));
}

__device__ __forceinline__ double inv_h3c2() { return 1.0 / ((H_KEV_SH * H_KEV_SH * H_KEV_SH) * (C_LIGHT * C_LIGHT)); }

// The group integrals follow Planck.cpp:85-229 with the pointwise functions of Planck.h:86-125 (k = 1):
//   z2 <= 0.7: 12-point Gauss-Legendre of  B(E) = 2 E^3 / (h^3 c^2 (e^{E/T} - 1)),  dB/dT(E) = 2 E^4 e^{E/T} / (h^3 c^2 T^2 (e^{E/T} - 1)^2)
//   z1 >= 0.5: 2 T^4 / (h^3 c^2) [S(z1) - S(z2)],  S(z) = sum_{n=N..1} e^{-nz}/n^4 P(nz) summed in DESCENDING n
//              (P cubic for B, quartic and T^3 for dB/dT; Planck.cpp:94-120, 170-194); N starts at 32 and grows while the
//              tail estimate exceeds 2.2e-16 relative to the first term
//   otherwise: split at z = 0.6 (Gauss below, series above); the result times 4*pi (truncated pi).
// ---- B and dB/dT of one group together.  Same operations, in the same order, as integrate_B and integrate_dBdT
// evaluated one after the other; what they have in common is evaluated once: exp(-n z) of the series terms (each
// sum keeps its own adaptive term count), the term weights 1/n^4, E/T and expm1(E/T) at the Gauss points.
// TRIM (per-cell tables only; the reference's own G-length evaluation, ncell = 1, keeps every term): the series terms are
// summed in descending n from N >= 32 whatever z is, but term_n / term_1 <= e^{-(n-1) z} (the polynomial ratio is at most
// n^4, cancelled by the 1/n^4 weight), so every term with (n-1) z > 46 is below e^-46 = 2^-66 of the leading term and all of
// them together below 2^-65 of the sum: they cannot move the fp64 result beyond a last-bit rounding flip.  Those terms are
// not evaluated: for z = 5 that leaves 11 of 32 terms, for z = 30 three.
template <bool TRIM>
__device__ void planck_series_pair(double z1, double z2, double &outB, double &outD)
{
   auto polyB = [](double x) { return (((x + 3.0) * x + 6.0) * x + 6.0); };
   auto polyD = [](double x) { return ((((x + 4.0) * x + 12.0) * x + 24.0) * x + 24.0); };
   const double eps = 2.220446049250313e-16;
   const double ez1 = exp(-z1);
   const double firstB = fmax(ez1 * polyB(z1), eps), firstD = fmax(ez1 * polyD(z1), eps);
   const double geo = 1.0 / (1.0 - ez1);
   int NB = 32, ND = 32;
   while (true)
   {
      const double n1 = NB + 1.0, n2 = n1 * n1;
      const double val = exp(-n1 * z1) * geo / (n2 * n2) * polyB(n1 * z1) / firstB;
      if (val > eps && NB < (1 << 20)) ++NB; else break;
   }
   while (true)
   {
      const double n1 = ND + 1.0, n2 = n1 * n1;
      const double val = exp(-n1 * z1) * geo / (n2 * n2) * polyD(n1 * z1) / firstD;
      if (val > eps && ND < (1 << 20)) ++ND; else break;
   }
   double b1 = 0.0, b2 = 0.0, d1 = 0.0, d2 = 0.0;
   int n_top = max(NB, ND), n_top2 = n_top;
   if (TRIM)
   {
      n_top = min(n_top, 2 + (int)(46.0 / z1));  // (n-1) z1 <= 46 (+1 for the truncation)
      n_top2 = min(n_top, 2 + (int)(46.0 / z2)); // z2 > z1: the upper edge's terms die out sooner
   }
   for (int n = n_top; n > 0; --n)
   {
      const double dn = (double)n, n2 = dn * dn, in4 = 1.0 / (n2 * n2);
      const double x1 = dn * z1, x2 = dn * z2;
      const double t1 = exp(-x1) * in4;
      if (n <= NB) b1 += t1 * polyB(x1);
      if (n <= ND) d1 += t1 * polyD(x1);
      if (!TRIM || n <= n_top2)
      {
         const double t2 = exp(-x2) * in4;
         if (n <= NB) b2 += t2 * polyB(x2);
         if (n <= ND) d2 += t2 * polyD(x2);
      }
   }
   outB = b1 - b2;
   outD = d1 - d2;
}

__device__ void gauss12_pair(double T, double lo, double hi, double &outB, double &outD)
{
   const double mid = 0.5 * (hi + lo), map = 0.5 * (hi - lo);
   double accB = 0.0, accD = 0.0;
#pragma unroll
   for (int i = 0; i < 12; ++i)
   {
      const double E = mid + map * c_gauss_pt[i];
      const double x = E / T, em = expm1(x), ex = exp(x), E2 = E * E;
      const double mw = map * c_gauss_w[i];
      accB += mw * (2.0 * (E * E * E) * inv_h3c2() / em);
      accD += mw * (2.0 * inv_h3c2() * (E2 * E2) / (T * T) * ex / (em * em));
   }
   outB = accB;
   outD = accD;
}

template <bool TRIM>
__device__ void planck_group_pair(double T, double E_min, double E_max, double &B, double &D)
{
   B = 0.0; D = 0.0;
   if (T == 0.0 || E_min == E_max) return;
   const double z1 = E_min / T, z2 = E_max / T;
   const double T2 = T * T;
   const double prefB = 2.0 * (T2 * T2) * inv_h3c2(), prefD = 2.0 * (T2 * T) * inv_h3c2();
   double vB, vD;
   if (z2 <= 0.7) gauss12_pair(T, E_min, E_max, vB, vD);
   else if (z1 >= 0.5)
   {
      planck_series_pair<TRIM>(z1, z2, vB, vD);
      vB *= prefB; vD *= prefD;
   }
   else
   {
      double sB, sD;
      gauss12_pair(T, E_min, 0.6 * T, vB, vD);
      planck_series_pair<TRIM>(0.6, z2, sB, sD);
      vB = vB + prefB * sB; vD = vD + prefD * sD;
   }
   B = vB * 4.0 * PI_TRUNC;
   D = vD * 4.0 * PI_TRUNC;
}

// Tables are GROUP-MAJOR: B[g * ncell + c].  One thread per (group, cell), CELL fastest: the lanes of a warp integrate
// the same group at neighbouring temperatures, so they take the same branch (Gauss / series / split) and run the same
// number of series terms -- with the group fastest a warp executed ~10 of 32 lanes per instruction (ncu) -- and the
// stores are coalesced.  g < G-1 only; the last group is the remainder (planck_finalize_kernel).
template <bool TRIM>
__global__ void planck_integrals_kernel(int G, int ncell, int g0, int ng, const double *__restrict__ e_edge,
                                        const double *__restrict__ T, double *__restrict__ B, double *__restrict__ dBdT)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= (long long)ncell * ng) return;
   const int g = g0 + (int)(idx / ncell), c = (int)(idx % ncell);
   double b, d;
   planck_group_pair<TRIM>(T[c], e_edge[g], e_edge[g + 1], b, d);
   B[(size_t)g * ncell + c] = b;
   dBdT[(size_t)g * ncell + c] = d;
}

// ---- group-sharded evaluation (material-temperature update on a group shard, rt_capi.cu): a rank integrates only the
// groups it owns; the last group's remainder needs the sum over ALL other groups, so every rank forms its partial sums
// per cell (ascending over its groups), the caller all-reduces them, and the owner of the last group takes the remainder.
// The sum is then (total - sum_g B_g) instead of the reference's sequential subtraction: same value to rounding.
__global__ void planck_partial_kernel(int ncell, int g0, int ng, double *__restrict__ B, double *__restrict__ dBdT,
                                      double *__restrict__ part)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= ncell) return;
   double bs = 0.0, ds = 0.0;
   for (int g = g0; g < g0 + ng; ++g)
   {
      const size_t o = (size_t)g * ncell + c;
      bs += B[o];
      ds += dBdT[o];
      B[o] *= K_JPK;
      dBdT[o] *= K_JPK;
   }
   part[c] = bs;
   part[ncell + c] = ds;
}

__global__ void planck_last_kernel(int G, int ncell, const double *__restrict__ T, const double *__restrict__ part,
                                   double *__restrict__ B, double *__restrict__ dBdT)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= ncell) return;
   const double t = T[c];
   const double bsum = rad_a_long() * C_LIGHT * (t * t) * (t * t) - part[c];
   const double dsum = 4.0 * rad_a_long() * C_LIGHT * (t * t) * t - part[ncell + c];
   const size_t o = (size_t)(G - 1) * ncell + c;
   B[o] = K_JPK * (bsum > 0.0 ? bsum : 0.0);
   dBdT[o] = K_JPK * (dsum > 0.0 ? dsum : 0.0);
}

// Remainder of the grey totals into the last group (sequential ascending subtraction,
// Planck.cpp:58-76: kept only if positive) and the keV -> jk unit change (correction.cpp:31-35).
// One thread per cell; group-major tables, so every step of the loop is a coalesced access.
__global__ void planck_finalize_kernel(int G, int ncell, const double *__restrict__ T, double *__restrict__ B,
                                       double *__restrict__ dBdT)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= ncell) return;
   const double t = T[c];
   double bsum = rad_a_long() * C_LIGHT * (t * t) * (t * t);
   double dsum = 4.0 * rad_a_long() * C_LIGHT * (t * t) * t;
   double *b = B + c, *d = dBdT + c;
   for (int g = 0; g < G - 1; ++g)
   {
      const size_t o = (size_t)g * ncell;
      bsum -= b[o];
      dsum -= d[o];
      b[o] *= K_JPK;
      d[o] *= K_JPK;
   }
   // last group: the remainder when positive, otherwise the stale previous value re-scaled (reference quirk)
   const size_t o = (size_t)(G - 1) * ncell;
   b[o] = K_JPK * (bsum > 0.0 ? bsum : b[o]);
   d[o] = K_JPK * (dsum > 0.0 ? dsum : d[o]);
}

__device__ __forceinline__ double pf(double E, double T) // correction.cpp:11-22
{
   return K_JPK * (E * E * E) * inv_h3c2() / expm1(E / T);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
   return v;
}

// Everything G-length that follows the Planck integrals.  One block; threads stride over groups.
__global__ void group_terms_kernel(GroupTermsArgs a)
{
   const int G = a.G, tid = threadIdx.x, nt = blockDim.x;
   const double T = a.T;
   // group averages / widths (solver.cpp:22-32) and edge opacities (correction.cpp:125-139)
   for (int g = tid; g <= G; g += nt)
   {
      double ke;
      if (g == 0) ke = a.kappa[0];
      else if (g == G) ke = a.kappa[G - 1];
      else
      {
         const double ea = 0.5 * (a.e_edge[g] + a.e_edge[g + 1]), eam = 0.5 * (a.e_edge[g - 1] + a.e_edge[g]);
         const double wl = (ea - a.e_edge[g]) / (ea - eam), wr = (a.e_edge[g] - eam) / (ea - eam);
         ke = a.kappa[g - 1] * wl + a.kappa[g] * wr;
      }
      a.kappa_edge[g] = ke;
      a.epf[g] = (g == 0) ? 0.0 : a.e_edge[g] * pf(a.e_edge[g], T); // E*pf(E) at each edge (edge 0 is never used)
   }
   __syncthreads();
   for (int g = tid; g < G; g += nt)
   {
      const double de = a.e_edge[g + 1] - a.e_edge[g];
      double dEB, dsig, dkap;
      // 2-point edge stencils (correction.cpp:165-174, 206-211, 241-249); note the last group's
      // dkapEB uses edge G-1 for kappa as well, and for G == 1 the "last" formulas win for dsigEdE only.
      if (g == 0)
      {
         dEB = a.epf[1];
         dsig = a.kappa_edge[1] * a.e_edge[1] / de;
         dkap = a.kappa_edge[1] * a.epf[1];
      }
      else if (g < G - 1)
      {
         dEB = a.epf[g + 1] - a.epf[g];
         dsig = (a.kappa_edge[g + 1] * a.e_edge[g + 1] - a.kappa_edge[g] * a.e_edge[g]) / de;
         dkap = a.kappa_edge[g + 1] * a.epf[g + 1] - a.kappa_edge[g] * a.epf[g];
      }
      else
      {
         dEB = -a.epf[G - 1];
         dsig = 0.0; // set below
         dkap = -a.kappa_edge[G - 1] * a.epf[G - 1];
      }
      if (g == G - 1) dsig = -a.kappa_edge[G] * a.e_edge[G] / de;
      a.dEB[g] = dEB;
      a.dsigEdE[g] = dsig;
      a.dkapEB[g] = dkap;
      const double sig = a.rho * a.kappa[g], Bg = a.B[g];
      a.sigma[g] = sig;
      a.srcB[g] = sig * Bg;
      a.cor1[g] = dsig;
      a.cor2[g] = 3.0 * sig * Bg - dkap;
      a.cor3[g] = dsig * (4.0 * Bg - dEB);
      a.de_ave[g] = de;
   }
   // validation sums (correction.cpp:39-63,100-122) -- warp-shuffle reduction
   double sb = 0.0, sd = 0.0, se = 0.0;
   for (int g = tid; g < G; g += nt)
   {
      sb += a.B[g];
      sd += a.dBdT[g];
      se += a.kappa[g] * a.B[g];
   }
   __shared__ double red[3][32];
   sb = warp_sum(sb); sd = warp_sum(sd); se = warp_sum(se);
   if ((tid & 31) == 0) { red[0][tid >> 5] = sb; red[1][tid >> 5] = sd; red[2][tid >> 5] = se; }
   __syncthreads();
   if (tid < 32)
   {
      const int nw = (nt + 31) >> 5;
      sb = tid < nw ? red[0][tid] : 0.0; sd = tid < nw ? red[1][tid] : 0.0; se = tid < nw ? red[2][tid] : 0.0;
      sb = warp_sum(sb); sd = warp_sum(sd); se = warp_sum(se);
      if (tid == 0) { a.sums[0] = sb; a.sums[1] = sd; a.sums[2] = se; }
   }
}

// Boundary inflow table bdry(i,g) (solver.cpp:635-692) from psi_source, optionally overwritten by the
// equilibrium sources psi_source(i,g) = B_g + mu_i*(V/c)*(4 B_g - dEB_g) (solver.cpp:297-314).
__global__ void boundary_table_kernel(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V,
                                      const double *__restrict__ mu, const double *__restrict__ B,
                                      const double *__restrict__ dEB, double *__restrict__ psi_source,
                                      double *__restrict__ bdry)
{
   const int idx = blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= M * G) return;
   const int i = idx % M, g = idx / M;
   double src = psi_source[idx];
   if (use_mg_equilib)
   {
      double val = 4.0 * B[g] - dEB[g];
      val *= mu[i] * V / C_LIGHT;
      val += B[g];
      src = val;
      psi_source[idx] = val;
   }
   double b;
   if (mu[i] < 0.0) b = (bc_right == 1) ? src : 0.0;          // vacuum and reflective right: no inflow
   else b = (bc_left == 0 || bc_left == 1) ? src : 0.0;        // vacuum falls through to source; reflective is resolved in the sweep
   bdry[idx] = b;
}

// ------------------------------------------------------------------------------------------ host side
void upload_gauss_rule(const double *pt, const double *w)
{
   cudaMemcpyToSymbol(c_gauss_pt, pt, 12 * sizeof(double));
   cudaMemcpyToSymbol(c_gauss_w, w, 12 * sizeof(double));
}

int launch_planck(int G, int ncell, const double *d_e_edge, const double *d_T, double *d_B, double *d_dBdT, cudaStream_t st)
{
   int launches = 0;
   if (G > 1)
   {
      const long long total = (long long)ncell * (G - 1);
      // ncell == 1 is the reference's own evaluation (one T, G groups): every term; per-cell tables: trimmed series
      if (ncell == 1) planck_integrals_kernel<false><<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, ncell, 0, G - 1, d_e_edge, d_T, d_B, d_dBdT);
      else planck_integrals_kernel<true><<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, ncell, 0, G - 1, d_e_edge, d_T, d_B, d_dBdT);
      ++launches;
   }
   planck_finalize_kernel<<<(ncell + 127) / 128, 128, 0, st>>>(G, ncell, d_T, d_B, d_dBdT);
   return launches + 1;
}

int launch_planck_shard_begin(int G, int ncell, int g0, int ng, const double *d_e_edge, const double *d_T, double *d_B,
                              double *d_dBdT, double *d_part, cudaStream_t st)
{
   const int n = min(g0 + ng, G - 1) - g0; // owned groups other than the last one
   if (n > 0)
   {
      const long long total = (long long)ncell * n;
      planck_integrals_kernel<true><<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, ncell, g0, n, d_e_edge, d_T, d_B, d_dBdT);
   }
   planck_partial_kernel<<<(ncell + 127) / 128, 128, 0, st>>>(ncell, g0, n > 0 ? n : 0, d_B, d_dBdT, d_part);
   return n > 0 ? 2 : 1;
}

// ---- a group RANGE without any exchange (a handle that owns part of the groups and stores only those, rt_capi.cu): the
// owner of the last group takes the grey remainder by the reference's sequential subtraction over ALL groups (Planck.cpp:
// 58-76), integrating the groups it does not own on the fly instead of reading them from a table -- same function of
// (T, edges), same order, so the result is bit for bit the one of the whole-table evaluation (planck_finalize_kernel).
// B / dBdT are BASE pointers: entry (g, c) at B[g * ncell + c] with the GLOBAL g, rows outside [g0, g0+ng) do not exist.
__global__ void planck_range_finalize_kernel(int G, int ncell, int g0, int ng, const double *__restrict__ e_edge,
                                             const double *__restrict__ T, double *__restrict__ B, double *__restrict__ dBdT)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= ncell) return;
   const double t = T[c];
   const bool owns_last = g0 + ng == G;
   if (owns_last)
   {
      double bsum = rad_a_long() * C_LIGHT * (t * t) * (t * t);
      double dsum = 4.0 * rad_a_long() * C_LIGHT * (t * t) * t;
      for (int g = 0; g < g0; ++g) // groups this handle does not own: integrated, not stored
      {
         double b, d;
         planck_group_pair<true>(t, e_edge[g], e_edge[g + 1], b, d);
         bsum -= b;
         dsum -= d;
      }
      for (int g = g0; g < G - 1; ++g) // owned groups: one pass, like planck_finalize_kernel
      {
         const size_t o = (size_t)g * ncell + c;
         bsum -= B[o];
         dsum -= dBdT[o];
         B[o] *= K_JPK;
         dBdT[o] *= K_JPK;
      }
      const size_t o = (size_t)(G - 1) * ncell + c;
      B[o] = K_JPK * (bsum > 0.0 ? bsum : B[o]);
      dBdT[o] = K_JPK * (dsum > 0.0 ? dsum : dBdT[o]);
      return;
   }
   for (int g = g0; g < g0 + ng; ++g)
   {
      B[(size_t)g * ncell + c] *= K_JPK;
      dBdT[(size_t)g * ncell + c] *= K_JPK;
   }
}

int launch_planck_range(int G, int ncell, int g0, int ng, const double *d_e_edge, const double *d_T, double *d_B_base, double *d_dBdT_base,
                        cudaStream_t st)
{
   const int n = min(g0 + ng, G - 1) - g0; // owned groups other than the last one
   if (n > 0)
   {
      const long long total = (long long)ncell * n;
      planck_integrals_kernel<true><<<(unsigned)((total + 127) / 128), 128, 0, st>>>(G, ncell, g0, n, d_e_edge, d_T, d_B_base, d_dBdT_base);
   }
   planck_range_finalize_kernel<<<(ncell + 127) / 128, 128, 0, st>>>(G, ncell, g0, ng, d_e_edge, d_T, d_B_base, d_dBdT_base);
   return n > 0 ? 2 : 1;
}

int launch_planck_shard_finish(int G, int ncell, const double *d_T, const double *d_part, double *d_B, double *d_dBdT, cudaStream_t st)
{
   planck_last_kernel<<<(ncell + 127) / 128, 128, 0, st>>>(G, ncell, d_T, d_part, d_B, d_dBdT);
   return 1;
}

int launch_group_terms(const GroupTermsArgs &a, cudaStream_t st)
{
   group_terms_kernel<<<1, 1024, 0, st>>>(a);
   return 1;
}

int launch_boundary_table(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V, const double *mu,
                          const double *B, const double *dEB, double *psi_source, double *bdry, cudaStream_t st)
{
   boundary_table_kernel<<<(M * G + 255) / 256, 256, 0, st>>>(M, G, bc_left, bc_right, use_mg_equilib, V, mu, B, dEB,
                                                             psi_source, bdry);
   return 1;
}
// Planck-weighted multigroup opacities normalised to the grey emission, Correction::generate_multigroup_opacities
// (correction.cpp:66-97; the reference keeps it but never calls it): ukappa_g = 2*4pi*k*T/(h^3 c^2) *
// (e^{-E_g/T} - e^{-E_{g+1}/T}) / B_g, ckappa_g = (1 - e^{-Eave_g/T}) / Eave_g^3, kappa_g = kappa_grey * (a c T^4 /
// sum ukappa*B) * ukappa_g, emis_g = kappa_g * B_g.  One block; the normalisation sum runs in ascending g in one thread
// like the reference.
__global__ void multigroup_opacities_kernel(int G, double T, double kappa_grey, const double *__restrict__ e_edge,
                                            const double *__restrict__ B, double *__restrict__ kappa,
                                            double *__restrict__ emis_spec, double *__restrict__ ukappa,
                                            double *__restrict__ ckappa)
{
   __shared__ double s_nfac;
   const double fourpi = 4.0 * PI_TRUNC;
   const double tmp1 = 2.0 * fourpi * K_JPK * T / ((H_KEV_SH * H_KEV_SH * H_KEV_SH) * (C_LIGHT * C_LIGHT));
   for (int g = threadIdx.x; g < G; g += blockDim.x)
   {
      const double tmp2 = exp(-e_edge[g] / T) - exp(-e_edge[g + 1] / T);
      const double e_ave = 0.5 * (e_edge[g] + e_edge[g + 1]);
      ukappa[g] = tmp1 * tmp2 / B[g];
      ckappa[g] = (1.0 - exp(-e_ave / T)) / (e_ave * e_ave * e_ave);
   }
   __syncthreads();
   if (threadIdx.x == 0)
   {
      double emis_tot = 0.0;
      for (int g = 0; g < G; ++g) emis_tot += ukappa[g] * B[g];
      s_nfac = RAD_A * C_LIGHT * (T * T) * (T * T) / emis_tot;
   }
   __syncthreads();
   for (int g = threadIdx.x; g < G; g += blockDim.x)
   {
      const double k = kappa_grey * s_nfac * ukappa[g];
      kappa[g] = k;
      emis_spec[g] = k * B[g];
   }
}

int launch_multigroup_opacities(int G, double T, double kappa_grey, const double *d_e_edge, const double *d_B, double *d_kappa,
                                double *d_emis, double *d_ukappa, double *d_ckappa, cudaStream_t st)
{
   multigroup_opacities_kernel<<<1, 256, 0, st>>>(G, T, kappa_grey, d_e_edge, d_B, d_kappa, d_emis, d_ukappa, d_ckappa);
   return 1;
}
} // namespace rtk
