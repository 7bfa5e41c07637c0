// group_terms.cuh -- launch interface of the G-length kernels (group_terms.cu).
#pragma once
#include <cuda_runtime.h>

namespace rtk
{
struct GroupTermsArgs
{
   int G;
   double T, rho;
   const double *e_edge; // G+1
   const double *kappa;  // G
   const double *B;      // G   (jk units, after the Planck kernels)
   const double *dBdT;   // G
   double *kappa_edge;   // G+1
   double *epf;          // G+1 scratch: E*pf(E,T) at the group edges
   double *dEB, *dsigEdE, *dkapEB, *cor1, *cor2, *cor3, *sigma, *srcB, *de_ave; // G each
   double *sums;         // 3: sum B, sum dBdT, sum kappa*B
};

// 12-point Gauss-Legendre rule (computed on the host in long double like the reference, rounded to fp64)
void upload_gauss_rule(const double *pt, const double *w);
// B,dBdT: GROUP-MAJOR [G][ncell] (cell fastest; with ncell = 1 simply G long); returns the number of kernels launched
int launch_planck(int G, int ncell, const double *d_e_edge, const double *d_T, double *d_B, double *d_dBdT, cudaStream_t st);
// group-sharded evaluation for the owned groups [g0, g0+ng): integrals + per-cell partial sums of the raw (keV) values into
// d_part[2*ncell] (B sums, then dB/dT sums); after the caller has summed d_part over the ranks, the owner of group G-1
// takes the grey remainder with launch_planck_shard_finish.  Same tables ([G][ncell]), only owned rows are touched.
int launch_planck_shard_begin(int G, int ncell, int g0, int ng, const double *d_e_edge, const double *d_T, double *d_B,
                              double *d_dBdT, double *d_part, cudaStream_t st);
// the owned groups [g0, g0+ng) WITHOUT exchange: integrals of the owned groups and, for the owner of group G-1, the grey
// remainder by sequential subtraction over all groups (the others integrated on the fly) -- bit for bit launch_planck's rows.
// d_B_base / d_dBdT_base: BASE pointers of tables that hold only rows [g0, g0+ng) (entry (g, c) at base[g*ncell + c]).
int launch_planck_range(int G, int ncell, int g0, int ng, const double *d_e_edge, const double *d_T, double *d_B_base, double *d_dBdT_base,
                        cudaStream_t st);
int launch_planck_shard_finish(int G, int ncell, const double *d_T, const double *d_part, double *d_B, double *d_dBdT, cudaStream_t st);
int launch_group_terms(const GroupTermsArgs &a, cudaStream_t st);
// correction.cpp:66-97; B in jk units (after launch_planck); all arrays G long
int launch_multigroup_opacities(int G, double T, double kappa_grey, const double *d_e_edge, const double *d_B, double *d_kappa,
                                double *d_emis, double *d_ukappa, double *d_ckappa, cudaStream_t st);
int launch_boundary_table(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V, const double *mu,
                          const double *B, const double *dEB, double *psi_source, double *bdry, cudaStream_t st);
} // namespace rtk
