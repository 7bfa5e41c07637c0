// warp_tile.cuh -- what the kernels built from warp tiles share (step_warp_kernel.cu, step4_kernel.cu, var_warp_kernel.cu):
// the words through which the warps of a CTA hand the recurrence carry to each other and count their progress, in shared
// memory.  Volatile accesses: the producer is another warp of the same CTA, there is no barrier between them.
#pragma once
#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"

namespace rtk
{
// {value, sequence} carry words between the warps of a CTA: one 128-bit shared-memory access each way
__device__ __forceinline__ double2 lds_word(const double2 *p)
{
   double2 v;
   asm volatile("ld.volatile.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(smem_u32(p)) : "memory");
   return v;
}
__device__ __forceinline__ void sts_word(double2 *p, double value, int seq1)
{
   const double f = __longlong_as_double((long long)seq1);
   asm volatile("st.volatile.shared.v2.f64 [%0], {%1, %2};" ::"r"(smem_u32(p)), "d"(value), "d"(f) : "memory");
}
// 32-bit counters of the descriptor ring / the warps' progress, and the sequence half of a carry word
__device__ __forceinline__ int ld_flag(const int *p)
{
   int v;
   asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
   return v;
}
__device__ __forceinline__ void st_flag(int *p, int v)
{
   asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

// one tile descriptor of a warp-tile kernel's ring: 16 bytes
struct WarpTile
{
   int row;   // row of the batch, -1 = ticket past the end
   int chain; // look-back chain (row, or pair of mirrored rows with a reflective left boundary)
   int p;     // tile position in the chain
   int tf;    // tile in row * 4 + sgn * 2 + junction
};

// ticket t -> tile: tickets run over the chains of one tile position before the next (p-major), so a tile's predecessor
// has normally been finished long before it is looked at.  half_m = M/2, dm = chains per group (M/2 with a reflective
// left boundary, where a chain is the mu < 0 row followed by its mirrored mu > 0 row; M otherwise)
__device__ __forceinline__ WarpTile decode_warp_tile(const StepArgs &a, long long t, long long total_tiles, int half_m, int dm)
{
   WarpTile d;
   d.row = -1; d.chain = 0; d.p = 0; d.tf = 0;
   if (t < total_tiles)
   {
      int p, chain;
      if (total_tiles < 0x7fffffffll)
      {
         p = (int)((unsigned)t / (unsigned)a.n_chains);
         chain = (int)((unsigned)t - (unsigned)p * (unsigned)a.n_chains);
      }
      else
      {
         p = (int)(t / a.n_chains);
         chain = (int)(t - (long long)p * a.n_chains);
      }
      const int gl = chain / dm, pp = chain - gl * dm;
      int sgn, row, tr;
      if (a.reflective_left)
      {
         sgn = (p >= a.tiles_per_row) ? 1 : 0;
         row = gl * a.M + (sgn ? half_m + pp : half_m - 1 - pp);
         tr = sgn ? p - a.tiles_per_row : p;
      }
      else
      {
         sgn = (pp >= half_m) ? 1 : 0;
         row = chain;
         tr = p;
      }
      d.row = row; d.chain = chain; d.p = p;
      d.tf = tr * 4 + sgn * 2 + ((tr == 0 && p > 0) ? 1 : 0);
   }
   return d;
}
} // namespace rtk
