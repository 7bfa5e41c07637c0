// rt_internal.h -- helpers rt_capi.cu lends to rt_multi.cu (not part of the C ABI, not exported from the library)
#pragma once
#include <vector>

#include "../../include/rt_cuda.h"

// sets the calling thread's rt_last_error() text and returns 1
__attribute__((visibility("hidden"))) int rt_internal_fail(const char *msg);
// one NCCL communicator per entry of `devices` (ncclCommInitAll, single process); comms[ndev] receives opaque handles
__attribute__((visibility("hidden"))) int rt_internal_nccl_init_all(void **comms, int ndev, const int *devices);
__attribute__((visibility("hidden"))) void rt_internal_nccl_destroy(void *comm);
// the rt_allreduce_fn that sums with ncclAllReduce on the communicator passed as `user`
__attribute__((visibility("hidden"))) int rt_internal_nccl_hook(void *device_buf, size_t count, void *cuda_stream, void *user);

// what correction_print.cu formats: the handle's scalars and its per-group tables (all G groups), read back from the device
struct rt_print_tables
{
   int G;
   double T, kappa_grey;
   std::vector<double> e_edge, kappa, B, dBdT, dEB, dsigEdE, dkapEB, kappa_edge, cor1, cor2, cor3;
};
__attribute__((visibility("hidden"))) int rt_internal_print_tables(rt_solver *s, rt_print_tables *t);
