// moments.cuh -- launch interface of moments.cu (initial state, moment reductions, diagnostics, exports).
#pragma once
#include <cuda_runtime.h>

namespace rtk
{
void launch_init_state(double2 *buf, const double *B, int M, int N, int Np, int G0, int groups, cudaStream_t st);
// group-major outputs [gl][c]; any of phi/phi_plus/F may be null
void launch_moments(const double2 *ends, const double *mu, const double *wt, int M, int N, int Np, int groups,
                    double *phi, double *phi_plus, double *F, cudaStream_t st);
// cells [c0,c0+nc) of in[groups][N] -> out[g + groups*(c-c0)]
void launch_transpose_gn(const double *in, double *out, int groups, int N, int c0, int nc, cudaStream_t st);
void launch_export_psi(const double2 *ends, double *psi, int M, int N, int Np, int groups, cudaStream_t st);
void launch_export_ends(const double2 *neg, const double2 *pos, double *out, int M, int N, int Np, int groups,
                        cudaStream_t st);
// phi: the batch's scalar intensity [groups][N] (group-major), as written by launch_moments for the same state
void launch_diagnostics(const double2 *ends, const double *phi, const double *mu, const double *wt, const double *sigma,
                        const double *de_ave, int M, int N, int Np, int G0, int groups, double dx, double T,
                        double *balance, double *left_ends, double *right_ends, cudaStream_t st);
// E[c] = (add ? E[c] : 0) + scale * sum_g phi[g][c] (ascending g, one thread per cell: deterministic) -- radiation energy of a time step
void launch_group_sum(const double *phi, int groups, int N, double scale, bool add, double *E, cudaStream_t st);
} // namespace rtk
