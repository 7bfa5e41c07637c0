// tools/stream_mix.cu -- what HBM bandwidth does a plain streaming kernel reach for the read/write mixes of the
// stage sweeps?  (BE = 1 read + 1 write stream, CN = 2R + 1W, BDF = 3R + 1W written in place over the third input.)
// Diagnostic only: gives the practical ceiling per mix next to the copy peak of MEASURED_PEAKS.json.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/stream_mix tools/stream_mix.cu && tools/stream_mix
#include <cstdio>
#include <cuda_runtime.h>

template <int R, bool INPLACE>
__global__ void __launch_bounds__(256) mix(const double2 *__restrict__ a, const double2 *__restrict__ b, const double2 *c,
                                           double2 *out, size_t n)
{
   const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
   for (size_t base = ((size_t)blockIdx.x * blockDim.x) * 4 + threadIdx.x; base < n; base += stride)
   {
      double2 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
      {
         const size_t i = base + (size_t)k * blockDim.x;
         if (i < n)
         {
            double2 x = __ldcs(a + i);
            if (R >= 2) { const double2 y = __ldcs(b + i); x.x += y.x; x.y += y.y; }
            if (R >= 3) { const double2 z = __ldcs(c + i); x.x += z.x; x.y -= z.y; }
            v[k] = x;
         }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
      {
         const size_t i = base + (size_t)k * blockDim.x;
         if (i < n) __stcs((INPLACE ? const_cast<double2 *>(c) : out) + i, v[k]);
      }
   }
}

template <int R>
__global__ void __launch_bounds__(256) readonly(const double2 *__restrict__ a, const double2 *__restrict__ b,
                                                const double2 *__restrict__ c, double *sink, size_t n)
{
   const size_t stride = (size_t)gridDim.x * blockDim.x;
   double acc = 0.0;
   for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
   {
      const double2 x = __ldcs(a + i);
      acc += x.x + x.y;
      if (R >= 2) { const double2 y = __ldcs(b + i); acc += y.x; }
      if (R >= 3) { const double2 z = __ldcs(c + i); acc += z.y; }
   }
   if (acc == 123.456) *sink = acc;
}

int main()
{
   const size_t n = (size_t)1 << 28; // 4 GiB per array
   double2 *a, *b, *c, *o;
   double *sink;
   cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16); cudaMalloc(&c, n * 16); cudaMalloc(&o, n * 16); cudaMalloc(&sink, 8);
   cudaMemset(a, 0, n * 16); cudaMemset(b, 0, n * 16); cudaMemset(c, 0, n * 16); cudaMemset(o, 0, n * 16);
   cudaEvent_t e0, e1;
   cudaEventCreate(&e0); cudaEventCreate(&e1);
   int sms = 0;
   cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
   for (int ctas_per_sm = 4; ctas_per_sm <= 8; ctas_per_sm += 4)
   {
      const int grid = sms * ctas_per_sm;
      auto time = [&](const char *name, double bytes, auto launch) {
         float best = 1e30f;
         for (int it = 0; it < 6; ++it)
         {
            cudaEventRecord(e0);
            launch();
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (it > 0 && ms < best) best = ms;
         }
         printf("grid %5d  %-26s %8.1f GB/s\n", grid, name, bytes / best / 1e6);
      };
      time("1R+1W (BE, copy)", 32.0 * n, [&] { mix<1, false><<<grid, 256>>>(a, b, c, o, n); });
      time("2R+1W (CN)", 48.0 * n, [&] { mix<2, false><<<grid, 256>>>(a, b, c, o, n); });
      time("3R+1W (BDF, separate out)", 64.0 * n, [&] { mix<3, false><<<grid, 256>>>(a, b, c, o, n); });
      time("3R+1W in place (BDF)", 64.0 * n, [&] { mix<3, true><<<grid, 256>>>(a, b, c, o, n); });
      time("1R", 16.0 * n, [&] { readonly<1><<<grid, 256>>>(a, b, c, sink, n); });
      time("3R", 48.0 * n, [&] { readonly<3><<<grid, 256>>>(a, b, c, sink, n); });
   }
   cudaError_t e = cudaDeviceSynchronize();
   printf("status: %s\n", cudaGetErrorString(e));
   return e != cudaSuccess;
}
