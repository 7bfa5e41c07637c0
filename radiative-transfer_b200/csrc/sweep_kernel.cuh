// sweep_kernel.cuh -- launch interface of the stage-sweep kernel (sweep_kernel.cu).
#pragma once
#include "rt_common.cuh"

namespace rtk
{
// kind: KIND_BE / KIND_CN / KIND_BDF ; corr: correction source live (use_correction && V != 0).
// Launches min(total tiles, num_sms * resident CTAs per SM) persistent CTAs on `st`.
// The caller must advance its ticket base by ceil(n_chains*tiles_per_chain / ticket_batch) + *grid_out afterwards.
// Tile shape: threads (64 / 128 / 256) x cells per thread (4 / 8); tile = threads * cpt cells.
bool sweep_shape_supported(int threads, int cpt);
// table[(gl*M + i)*COEF_ENTRY_DOUBLES ...] for the a.groups owned groups: stepper constants of stage `kind`
void launch_coef_table(int kind, bool corr, const CoefArgs &a, double *table, cudaStream_t st);
// tensor map of one time-level buffer [rows][Np] of double2 whose box is one tile of `tile_cells` cells
cudaError_t make_tile_map(CUtensorMap *map, void *base, int rows, int Np, int tile_cells);
// ring: shared-memory slots for the X1/result tile, 2 or 3 (0 = 3 unless 2 lets one more CTA reside per SM)
cudaError_t launch_sweep(int kind, bool corr, int threads, int cpt, int ring, const SweepMaps &maps, const SweepArgs &a,
                         int num_sms, cudaStream_t st, int *grid_out);
} // namespace rtk
