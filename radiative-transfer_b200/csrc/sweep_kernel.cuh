// sweep_kernel.cuh -- launch interface of the stage-sweep kernel (sweep_kernel.cu).
#pragma once
#include "rt_common.cuh"

namespace rtk
{
// kind: KIND_BE / KIND_CN / KIND_BDF ; corr: correction source live (use_correction && V != 0).
// Launches min(total tiles, num_sms * resident CTAs per SM) persistent CTAs on `st`.
// The caller must advance its ticket base by (n_chains*tiles_per_chain + *grid_out) afterwards.
// threads: CTA size 64 / 128 / 256 (tile = 8 * threads cells)
cudaError_t launch_sweep(int kind, bool corr, int threads, const SweepArgs &a, int num_sms, cudaStream_t st, int *grid_out);
} // namespace rtk
