// var_kernel.cuh -- cell-dependent material fields rho(c), T(c), kappa_g(c): EXTENSION.
//
// The reference's slab is uniform -- rho, T and kappa_g are filled with constants, the spatial variation is a TODO at
// solver.cpp:150-157 and correction.h:91 -- so this path has no reference to compare with beyond uniform inputs
// (where it must, and does, reproduce the uniform path).  Its oracle is the cell-dependent restatement in
// oracle/rt_oracle.cpp, which is bit-identical to the reference-pinned uniform oracle for uniform arrays.
#pragma once
#include "rt_common.cuh"
#include "step_kernel.cuh"

namespace rtk
{
// Per-cell tables of the owned groups: what the steppers need of the material state.  Every table is kept TWICE, once per
// sweep direction: entry (sign, gl, j) at (sign*groups + gl)*pitch + j with j the storage index ALONG THE SWEEP (sign 1,
// mu > 0: j = cell ; sign 0, mu < 0: j = N-1-cell), pitch = N rounded up to 16 -- so every row reads its tile of a
// table as one aligned TMA box in its own order (var_tile_kernel.cu).  The forward copy (sign 1) serves everything else.
struct CellTables
{
   const double *sigma, *srcB, *cor1, *cor2, *cor3; // FORWARD copies: rho*kappa_g, rho*kappa_g*B_g, correction terms (correction.cpp:328-340)
   int pitch;
};

// tensor maps of the five tables (rows = 2*groups, SWIZZLE_128B boxes of one tile) for var_tile_kernel
struct alignas(64) MatMaps
{
   CUtensorMap m[5];
};

struct VarTileArgs
{
   SweepArgs s;        // time levels, look-back state, tickets (s.coef unused)
   const double *mu;   // M
   const double *bdry; // i + M*g (global group)
   int G0, Gl;         // first owned group, owned groups (= rows of one table copy)
   double dx, tau, dt, beta;
};
// fused BDF2 step with cell-dependent fields (var_step_kernel.cu): all four stage sweeps in one pass per tile
struct VarStepArgs
{
   StepArgs s;         // ends^n / ends^(n+1), FOUR look-back words per tile, tickets (s.coef / s.coef_step unused)
   const double *mu;   // M
   const double *bdry; // i + M*g (global group)
   int G0, Gl;
   double dx, tau, dt, beta; // tau = dt/2
};
// var_warp_kernel.cu: the fused step from warp tiles -- chain_warps consecutive sub-tiles of var_warp_subtile_cells() cells per
// CTA tile; maps / mats must have boxes of ONE sub-tile
bool var_warp_chain_supported(int chain_warps); // 2, 4, 8
int var_warp_subtile_cells();
cudaError_t launch_var_warp(bool corr, int chain_warps, const SweepMaps &maps, const MatMaps &mats, const VarStepArgs &va, int num_sms,
                            cudaStream_t st, int *grid_out);
bool var_step_shape_supported(int threads, int cpt); // 64 x 4, 128 x 4
// mat_slots: shared-memory slots of material tables, 2 or (64 x 4 only) 1 ; scan_steps: what launch_var_scan_steps found (5 = all)
cudaError_t launch_var_step(bool corr, int threads, int cpt, int mat_slots, int scan_steps, const SweepMaps &maps, const MatMaps &mats,
                            const VarStepArgs &va, int num_sms, cudaStream_t st, int *grid_out);
// *max_steps (device) = the largest number of warp-scan steps any row needs (0..5), from the smallest sigma of every group
void launch_var_scan_steps(const double *sigma_fwd, int pitch, int N, int groups, const double *mu, int M, double dx, double tau, double dt,
                           double beta, int cpt, int *max_steps, cudaStream_t st);

bool var_tile_shape_supported(int threads, int cpt); // 64 x 4, 64 x 8, 128 x 4
// persistent CTAs of `threads` x `cpt` cells; the caller advances its ticket base like after launch_sweep
cudaError_t launch_var_tile(int kind, bool corr, int threads, int cpt, const SweepMaps &maps, const MatMaps &mats, const VarTileArgs &va,
                            int num_sms, cudaStream_t st, int *grid_out);

struct CellTermsArgs
{
   int G, N, G0, groups;     // all groups enter the edge stencils; tables are written for the owned ones [G0, G0+groups)
   const double *e_edge;     // G+1
   const double *kappa_cell; // GROUP-MAJOR [g][c] (transposed from the caller's g + G*c at rt_create)
   const double *rho_cell;   // N
   const double *T_cell;     // N
   const double *B_cell;     // [g][c], jk units (planck kernels with ncell = N)
   const double *dBdT_cell;  // [g][c] ; only read when sdB is wanted
   double *sigma, *srcB, *cor1, *cor2, *cor3; // both sweep-ordered copies, see CellTables (base of the sign-0 copy)
   int pitch;                // N rounded up to 16
   double *sdB;              // [gl][pitch] forward only: sigma_g dB_g/dT, or nullptr (material-temperature update only)
   double *B_left, *dEB_left, *B_right, *dEB_right; // G each: material state of cell 0 / cell N-1 (boundary sources)
};
void launch_cell_terms(const CellTermsArgs &a, cudaStream_t st);      // tables of the owned groups
void launch_boundary_state(const CellTermsArgs &a, cudaStream_t st);  // B_left .. dEB_right, all G groups

// bdry(i,g) and, with use_mg_equilib, psi_source(i,g) from the boundary cells' state (mu>0: cell 0, mu<0: cell N-1)
void launch_boundary_table_var(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V, const double *mu,
                               const double *B_left, const double *dEB_left, const double *B_right, const double *dEB_right,
                               double *psi_source, double *bdry, cudaStream_t st);

// psi = ends(L) = ends(R) = B_g(T(c)) (solver.cpp:166-181 with the cell's temperature)
void launch_init_state_var(double2 *buf, const double *B_cell, int M, int N, int Np, int G, int G0, int groups, cudaStream_t st);

struct VarSweepArgs
{
   const double2 *x1[2]; // X1 for mu<0 / mu>0 rows
   const double2 *x2;    // CN, BDF
   const double2 *xc;    // BDF with the correction: source cell values
   double2 *out;
   CellTables t;
   const double *mu;   // M
   const double *bdry; // i + M*g (global group)
   int M, N, Np, G0, rows, reflective_left;
   double dx, tau, dt, beta;
};
// one CTA per chain, tiles in sequence (no look-back): BE / CN / BDF stage with cell-dependent stepper constants
cudaError_t launch_var_sweep(int kind, bool corr, const VarSweepArgs &a, cudaStream_t st, int *grid_out);

// ---- material-temperature update: EXTENSION, parity unpinned (the reference has none); the
// definition is oracle/rt_oracle.cpp rto_temperature_update.
struct ExchangeArgs
{
   int N, groups, pitch;     // owned groups; phi is [gl][N], the tables [gl][pitch] (forward copies)
   double W;                 // sum of the quadrature weights
   const double *sigma, *srcB, *sdB, *phi;
   double *partial;          // [chunk][3][N] scratch, chunk = EXCHANGE_GROUPS_PER_CHUNK groups
   double *sums;             // [3][N]: A, E, D summed over the owned groups (ascending)
};
constexpr int EXCHANGE_GROUPS_PER_CHUNK = 32;
inline int exchange_chunks(int groups) { return (groups + EXCHANGE_GROUPS_PER_CHUNK - 1) / EXCHANGE_GROUPS_PER_CHUNK; }
// absorbed / emitted power and the emission derivative per cell, summed over the owned groups; returns kernels launched
int launch_exchange_sums(const ExchangeArgs &a, cudaStream_t st);
// T <- T' from the (all-reduced) sums; T_prev <- T.  bdf2 = later steps of a BDF2 run
void launch_temperature_update(int N, bool bdf2, double dt, double cv, const double *rho_cell, const double *sums, double *T,
                               double *T_prev, int update_no, int *first_bad, cudaStream_t st);

// balance (solver.cpp:240-284) with per-cell sigma and T, group-end spectra (826-850)
void launch_diagnostics_var(const double2 *ends, const double *phi, const double *mu, const double *wt, const double *sigma_cell,
                            int pitch, const double *T_cell, const double *de_ave, int M, int N, int Np, int G0, int groups, double dx,
                            double *balance, double *left_ends, double *right_ends, cudaStream_t st);
} // namespace rtk
