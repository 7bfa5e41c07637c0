// warp_tile.cuh -- what the kernels built from warp tiles share (step_warp_kernel.cu, step4_kernel.cu, var_warp_kernel.cu):
// the words through which the warps of a CTA hand the recurrence carry to each other and count their progress, in shared
// memory.  Volatile accesses: the producer is another warp of the same CTA, there is no barrier between them.
#pragma once
#include "rt_common.cuh"
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
} // namespace rtk
