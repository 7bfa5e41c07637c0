// rt_common.cuh -- shared definitions for the sm_100a kernels of librt_cuda.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

namespace rtk
{
// Physical constants, bit-identical to the reference's (include/Constants.h:9-23).
// PI is the reference's truncated value on purpose.
constexpr double H_KEV_SH = 4.141895e-10;
constexpr double K_JPK = 1.601558e-25;
constexpr double C_LIGHT = 299.79245800;
constexpr double PI_TRUNC = 3.1415926546;
constexpr double RAD_A = 1.3653104e-2;
constexpr double VALIDATION_TOL = 1.e-6;

// ---------------------------------------------------------------------------------------------
// Device data layout ("group-major, sweep-ordered rows")
//
//   One ROW per (local group gl, direction i):  row = gl*M + i.
//   A row holds the N cells of that (group, direction) in SWEEP ORDER -- storage index j is the
//   j-th cell the upwind sweep visits: cell = j for mu>0, cell = N-1-j for mu<0 -- padded to a
//   pitch of Np = round_up(N, 8) cells.
//   Each cell is one double2 (u, v) = (inflow-side node, outflow-side node):
//        mu>0: (u,v) = (psi_L, psi_R)        mu<0: (u,v) = (psi_R, psi_L)
//   With this ordering the three steppers are the SAME formulas for both signs of mu with |mu| as
//   streaming constant (mirror symmetry of the reference's mu<0 / mu>0 branches, solver.cpp:327-399).
//   The cell average psi = (u+v)/2 is never stored (solver.cpp:352: psi is exactly 0.5*(L+R)).
// ---------------------------------------------------------------------------------------------


enum StageKind : int { KIND_BE = 0, KIND_CN = 1, KIND_BDF = 2 };

constexpr int RT_MAX_DEVICES = 64;    // per-device launch configuration caches (shared-memory opt-in, occupancy)
// Launch plans (shared-memory opt-in, resident CTAs per SM) are cached PER DEVICE: cudaFuncAttributeMaxDynamicSharedMemorySize
// is a per-device attribute, and one process may own handles on several GPUs (rt_params::device, device_count).
struct DevicePlanCache
{
   std::atomic<int> v[RT_MAX_DEVICES]; // 0 = not planned on that device yet (static storage: zero-initialised)
};
inline cudaError_t device_plan_slot(DevicePlanCache &c, std::atomic<int> **slot)
{
   int dev = 0;
   const cudaError_t e = cudaGetDevice(&dev);
   if (e != cudaSuccess) return e;
   if (dev < 0 || dev >= RT_MAX_DEVICES) return cudaErrorInvalidDevice;
   *slot = &c.v[dev];
   return cudaSuccess;
}

constexpr int COEF_ENTRY_DOUBLES = 18; // sizeof(RowCoef) / 8 (sweep_kernel.cu)

// Inputs of the per-row stepper-constant table (one entry per owned (group, direction) and stage kind).
struct CoefArgs
{
   const double *sigma;  // rho*kappa_g                                   (indexed by GLOBAL group)
   const double *srcB;   // rho*kappa_g*B_g  (reference: rho*kappa*B, solver.cpp:337)
   const double *cor1, *cor2, *cor3;
   const double *mu;     // M
   const double *bdry;   // M*G inflow values psi_source(i,g) already masked by the BC switch (0 where vacuum)
   int M, G0, groups;    // G0 = global index of the first owned group
   double dx, tau, dt, beta;
};

// Tensor maps of the four rotating time-level buffers (SWIZZLE_128B tile boxes), passed as a __grid_constant__ kernel
// parameter; SweepArgs::ix1/ix2/ixc/iout index into it.
struct alignas(64) SweepMaps
{
   CUtensorMap m[4];
};

// Arguments of one stage-sweep launch.
struct SweepArgs
{
   // time levels; X1 per direction sign s (0: mu<0 rows, 1: mu>0 rows) because the BDF stage reads a different `half`
   // buffer for the two signs.  Raw pointers (row-indexed with pitch Np) for the L2 prefetch and the
   // one-cell halo copies; the tiles themselves move through the tensor maps.
   const double2 *x1[2]; // BE/CN: ends before the stage ; BDF: half_ends
   const double2 *x2;    // CN/BDF: prev_ends             ; BE: unused
   const double2 *xc;    // BDF with live correction: ends before the stage (psi for the source)
   int ix1[2], ix2, ixc, iout; // the same buffers (and the output buffer) as indices into SweepMaps::m
   const double2 *coef;  // stepper constants of this stage kind, entry `row` of the batch at coef + row*9 (144-byte entries)
   // look-back state
   double2 *tile_status;          // [chains*tiles_per_chain] {value, flag}: flag = epoch*2+1 -> value is the tile's zero-inflow
                                  // aggregate, epoch*2+2 -> value is the true outflow at the tile's last cell
   unsigned long long *ticket;    // monotonically increasing counter of ticket DRAWS (ticket_batch tickets per draw)
   unsigned long long ticket_base;
   int ticket_batch;
   unsigned int epoch;
   int M, N, Np;
   int rows;             // groups_in_batch * M
   int tiles_per_row;
   int reflective_left;  // 1: mu>0 rows take their inflow from the mirrored mu<0 row of the same stage
   int n_chains, tiles_per_chain;
   int l2_prefetch;      // 1: bulk-prefetch the tile two tickets ahead into L2
   int lookback_walk;    // 0: deterministic (wait for the predecessor's inclusive value) ; 1: classic aggregate walk-back
};
} // namespace rtk
