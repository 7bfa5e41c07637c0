// step_kernel.cuh -- launch interface of the fused BDF2-step kernel (step_kernel.cu).
#pragma once
#include "rt_common.cuh"

namespace rtk
{
// Arguments of one fused BDF2-step launch (all four stages of solver.cpp:706-809 for every owned row of the batch).
struct StepArgs
{
   const double2 *x;     // ends^n, raw pointer (row-indexed with pitch Np): one-cell halo copies and the L2 prefetch
   int ix, iout;         // ends^n and ends^(n+1) as indices into SweepMaps::m (out of place: prev_ends stays intact)
   const double2 *coef[3]; // stepper constants of the batch's rows for BE(dt/2), CN(dt/2), BDF(dt/2, dt)
   const double *coef_step; // the same constants packed per row, STEP_COEF_DOUBLES each (warp-tile kernel)
   double2 *tile_status;   // FOUR {value, flag} words per tile: one look-back chain per stage
   unsigned long long *ticket;
   unsigned long long ticket_base;
   int ticket_batch;
   unsigned int epoch;
   int M, N, Np;
   int rows;
   int tiles_per_row;
   int reflective_left;
   int n_chains, tiles_per_chain;
   int l2_prefetch;
   // one-scan kernel (step4_kernel.cu)
   const double *coef4;  // packed row constants, STEP4_COEF_DOUBLES each: stage constants, T^(2^s), response R
   double *phi_part;     // PHI: partial sums of w_m psi per run of rows [(gl*2 + sign)*phi_slots + slot][phi_pitch], sweep order
   long long phi_pitch;  // cells per partial row: whole tiles (tiles_per_row * tile cells)
   int phi_slots;        // runs per (group, sign): (M/2) / ticket_batch
};

constexpr int STEP_COEF_DOUBLES = 64; // packed BE + CN + BDF constants of one row (step_warp_kernel.cu)
constexpr int STEP4_COEF_DOUBLES = 160; // the same + chunk transfer matrices + stage-3 response (step4_kernel.cu)

// threads = 32: one warp per tile (step_warp_kernel.cu, the default); 64 / 128: CTA-wide tiles (step_kernel.cu)
bool step_shape_supported(int threads, int cpt);
void launch_pack_step_coef(const double *be, const double *cn, const double *bdf, double *out, int rows, bool corr, cudaStream_t st);
bool step_chain_supported(int chain_warps); // warps (consecutive 32*cpt-cell sub-tiles) per CTA tile: 1, 2, 4 or 8
cudaError_t launch_step_warp(bool corr, int cpt, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st,
                             int *grid_out);
// The caller must advance its ticket base by ceil(n_chains*tiles_per_chain / ticket_batch) + *grid_out afterwards.
// chain_warps only matters for threads = 32.
cudaError_t launch_step(bool corr, int threads, int cpt, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms,
                        cudaStream_t st, int *grid_out);

// ---- step4_kernel.cu: the fused step with ONE 4-vector scan per sub-tile (default)
bool step4_chain_supported(int chain_warps); // 2, 4 or 8 warps (consecutive 256-cell sub-tiles) per CTA tile
void launch_pack_step4_coef(const double *be, const double *cn, const double *bdf, const double *wt, double *out, int rows, int M, bool corr,
                            cudaStream_t st);
// phi: accumulate w_m psi of the step's result per run of a.ticket_batch rows into a.phi_part (a.ticket_batch must divide M/2)
cudaError_t launch_step4(bool corr, bool phi, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st,
                         int *grid_out);
// phi[gl][c] from the partial sums of one step4 launch
void launch_phi_combine(const double *part, long long pitch, int slots, int N, int groups, double *phi, cudaStream_t st);
} // namespace rtk
