// moments.cu -- state initialisation, angular moment reductions, diagnostics and the layout
// conversions at the C-ABI edge.  fp64, sm_100a.
//
// Reference being reproduced (results, not code):
//   initial condition psi = ends = B_g                      src/solver.cpp:166-181
//   compute_angle_integrated_intensity / positive / flux    src/solver.cpp:191-237
//   compute_balance                                         src/solver.cpp:240-284
//   compute_group_ends                                      src/solver.cpp:826-850
#include "rt_common.cuh"
#include "moments.cuh"

namespace rtk
{
// rows x Np cells of (B_g, B_g); pad cells are zero
__global__ void init_state_kernel(double2 *__restrict__ buf, const double *__restrict__ B, int M, int N, int Np, int G0,
                                  long long total)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= total) return;
   const int j = (int)(idx % Np);
   const int row = (int)(idx / Np);
   const double v = (j < N) ? B[G0 + row / M] : 0.0;
   buf[idx] = make_double2(v, v);
}

// phi, phi_plus, F for one batch, group-major output [gl][c].  One thread per (gl, c); the sum over
// directions runs in ascending m like the reference.  For a fixed m a warp reads 32 consecutive
// cells of one row (512 contiguous bytes, reversed for mu<0 rows).  The direction loop is blocked by
// UNR (8 when M is a multiple of 8, else 4) with the loads issued first, so every thread keeps UNR
// independent 16-byte loads in flight (the plain loop serialised them and reached only 2.7 TB/s).
template <int UNR>
__global__ void __launch_bounds__(256) moments_kernel(const double2 *__restrict__ ends, const double *__restrict__ mu,
                                                      const double *__restrict__ wt, int M, int N, int Np, int groups,
                                                      double *__restrict__ phi, double *__restrict__ phi_plus,
                                                      double *__restrict__ F)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= (long long)groups * N) return;
   const int c = (int)(idx % N), gl = (int)(idx / N);
   const int half_m = M >> 1;
   const double2 *base = ends + (size_t)gl * M * Np;
   const int jneg = N - 1 - c;
   double a = 0.0, b = 0.0, f = 0.0;
   int m = 0;
   for (; m + UNR <= M; m += UNR)
   {
      double2 e[UNR];
#pragma unroll
      for (int k = 0; k < UNR; ++k) e[k] = __ldcs(base + (size_t)(m + k) * Np + ((m + k) < half_m ? jneg : c));
#pragma unroll
      for (int k = 0; k < UNR; ++k)
      {
         const double psi = 0.5 * (e[k].x + e[k].y);
         const double w = wt[m + k];
         a += w * psi;
         if (m + k >= half_m) b += w * psi;
         f += mu[m + k] * w * psi;
      }
   }
   for (; m < M; ++m)
   {
      const double2 e = __ldcs(base + (size_t)m * Np + (m < half_m ? jneg : c));
      const double psi = 0.5 * (e.x + e.y);
      const double w = wt[m];
      a += w * psi;
      if (m >= half_m) b += w * psi;
      f += mu[m] * w * psi;
   }
   if (phi) phi[idx] = a;
   if (phi_plus) phi_plus[idx] = b;
   if (F) F[idx] = f;
}

// cells [c0, c0+nc) of a group-major array [groups][N] -> reference order g + groups*(c-c0), 32x32 tiles
__global__ void transpose_gn_kernel(const double *__restrict__ in, double *__restrict__ out, int groups, int N, int c0,
                                    int nc)
{
   __shared__ double tile[32][33];
   const int cb = blockIdx.x * 32, g0 = blockIdx.y * 32;
   for (int r = threadIdx.y; r < 32; r += blockDim.y)
   {
      const int g = g0 + r, c = cb + threadIdx.x;
      if (g < groups && c < nc) tile[r][threadIdx.x] = in[(size_t)g * N + c0 + c];
   }
   __syncthreads();
   for (int r = threadIdx.y; r < 32; r += blockDim.y)
   {
      const int c = cb + r, g = g0 + threadIdx.x;
      if (g < groups && c < nc) out[(size_t)g + (size_t)groups * c] = tile[threadIdx.x][r];
   }
}

// psi(i,g,c) at i + M*(g + groups*c)
__global__ void export_psi_kernel(const double2 *__restrict__ ends, double *__restrict__ psi, int M, int N, int Np,
                                  int groups)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= (long long)M * groups * N) return;
   const int i = (int)(idx % M);
   const long long r = idx / M;
   const int g = (int)(r % groups), c = (int)(r / groups);
   const int j = (i < (M >> 1)) ? (N - 1 - c) : c;
   const double2 e = ends[(size_t)(g * M + i) * Np + j];
   psi[idx] = 0.5 * (e.x + e.y);
}

// ends(i,g,c,s) at i + M*(g + groups*(c + N*s)); (u,v) -> (L,R)
__global__ void export_ends_kernel(const double2 *__restrict__ neg, const double2 *__restrict__ pos,
                                   double *__restrict__ out, int M, int N, int Np, int groups)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   const long long n3 = (long long)M * groups * N;
   if (idx >= n3) return;
   const int i = (int)(idx % M);
   const long long r = idx / M;
   const int g = (int)(r % groups), c = (int)(r / groups);
   const bool is_neg = i < (M >> 1);
   const int j = is_neg ? (N - 1 - c) : c;
   const double2 e = (is_neg ? neg : pos)[(size_t)(g * M + i) * Np + j];
   out[idx] = is_neg ? e.y : e.x;      // left node
   out[idx + n3] = is_neg ? e.x : e.y; // right node
}

__device__ __forceinline__ double block_sum(double v, double *red)
{
#pragma unroll
   for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
   const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
   if (lane == 0) red[warp] = v;
   __syncthreads();
   v = (threadIdx.x < nw) ? red[threadIdx.x] : 0.0;
   if (warp == 0)
   {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
   }
   __syncthreads();
   return v; // valid in warp 0
}

// One block per local group: balance (solver.cpp:240-284) and group-end spectra (solver.cpp:826-850).
// The sum over cells of sigma*phi*dx is a warp-shuffle block reduction; the sum over directions of the
// boundary currents runs in ascending m in one thread.
__global__ void diagnostics_kernel(const double2 *__restrict__ ends, const double *__restrict__ phi, const double *__restrict__ mu,
                                   const double *__restrict__ wt, const double *__restrict__ sigma,
                                   const double *__restrict__ de_ave, int M, int N, int Np, int G0, double dx, double T,
                                   double *__restrict__ balance, double *__restrict__ left_ends,
                                   double *__restrict__ right_ends)
{
   __shared__ double red[32];
   const int gl = blockIdx.x, g = G0 + gl, half_m = M >> 1;
   const double sig = sigma[g];
   // sum_c sigma*phi(g,c)*dx from the scalar intensity the moments kernel has just written (1/(2M) of the bytes of `ends`)
   double part = 0.0;
   const double *prow = phi + (size_t)gl * N;
   for (int c = threadIdx.x; c < N; c += blockDim.x) part += sig * prow[c] * dx;
   const double absorbed = block_sum(part, red);
   if (threadIdx.x == 0)
   {
      double j_half_minus = 0., j_half_plus = 0., jN_half_minus = 0., jN_half_plus = 0., l = 0., r = 0.;
      for (int m = 0; m < M; ++m)
      {
         const double2 *row = ends + (size_t)(gl * M + m) * Np;
         const double v_first = row[0].y, v_last = row[N - 1].y; // outflow nodes of the first / last swept cell
         const double mw = mu[m] * wt[m];
         if (m < half_m)
         {
            j_half_minus -= v_last * mw;   // ends(m,g,0,0):   cell 0 is swept last, left node = outflow
            jN_half_minus -= v_first * mw; // ends(m,g,N-1,0)
            l += v_last;
         }
         else
         {
            j_half_plus += v_first * mw;   // ends(m,g,0,1)
            jN_half_plus += v_last * mw;   // ends(m,g,N-1,1)
            r += v_last;
         }
      }
      const double ac = RAD_A * C_LIGHT;
      const double src = sig * ac * (T * T) * (T * T) * dx * (double)N;
      const double sources = j_half_plus + jN_half_minus + src;
      const double sinks = jN_half_plus + j_half_minus + absorbed;
      balance[gl] = fabs(sinks - sources) / sources;
      left_ends[gl] = l / (de_ave[g] * C_LIGHT);
      right_ends[gl] = r / (de_ave[g] * C_LIGHT);
   }
}

// ------------------------------------------------------------------------------------------ host side
static inline unsigned blocks_for(long long n, int bs) { return (unsigned)((n + bs - 1) / bs); }

void launch_init_state(double2 *buf, const double *B, int M, int N, int Np, int G0, int groups, cudaStream_t st)
{
   const long long total = (long long)groups * M * Np;
   init_state_kernel<<<blocks_for(total, 256), 256, 0, st>>>(buf, B, M, N, Np, G0, total);
}

void launch_moments(const double2 *ends, const double *mu, const double *wt, int M, int N, int Np, int groups,
                    double *phi, double *phi_plus, double *F, cudaStream_t st)
{
   const long long total = (long long)groups * N;
   if (M % 8 == 0) moments_kernel<8><<<blocks_for(total, 256), 256, 0, st>>>(ends, mu, wt, M, N, Np, groups, phi, phi_plus, F);
   else moments_kernel<4><<<blocks_for(total, 256), 256, 0, st>>>(ends, mu, wt, M, N, Np, groups, phi, phi_plus, F);
}

void launch_transpose_gn(const double *in, double *out, int groups, int N, int c0, int nc, cudaStream_t st)
{
   dim3 grid((nc + 31) / 32, (groups + 31) / 32), block(32, 8);
   transpose_gn_kernel<<<grid, block, 0, st>>>(in, out, groups, N, c0, nc);
}

void launch_export_psi(const double2 *ends, double *psi, int M, int N, int Np, int groups, cudaStream_t st)
{
   const long long total = (long long)M * groups * N;
   export_psi_kernel<<<blocks_for(total, 256), 256, 0, st>>>(ends, psi, M, N, Np, groups);
}

void launch_export_ends(const double2 *neg, const double2 *pos, double *out, int M, int N, int Np, int groups,
                        cudaStream_t st)
{
   const long long total = (long long)M * groups * N;
   export_ends_kernel<<<blocks_for(total, 256), 256, 0, st>>>(neg, pos, out, M, N, Np, groups);
}

void launch_diagnostics(const double2 *ends, const double *phi, const double *mu, const double *wt, const double *sigma,
                        const double *de_ave, int M, int N, int Np, int G0, int groups, double dx, double T,
                        double *balance, double *left_ends, double *right_ends, cudaStream_t st)
{
   diagnostics_kernel<<<groups, 1024, 0, st>>>(ends, phi, mu, wt, sigma, de_ave, M, N, Np, G0, dx, T, balance, left_ends,
                                              right_ends);
}
__global__ void group_sum_kernel(const double *__restrict__ phi, int groups, int N, double scale, int add, double *__restrict__ E)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= N) return;
   double acc = 0.0;
   for (int g = 0; g < groups; ++g) acc += phi[(size_t)g * N + c]; // coalesced over cells
   E[c] = add ? E[c] + scale * acc : scale * acc;
}

void launch_group_sum(const double *phi, int groups, int N, double scale, bool add, double *E, cudaStream_t st)
{
   group_sum_kernel<<<(N + 127) / 128, 128, 0, st>>>(phi, groups, N, scale, add ? 1 : 0, E);
}
} // namespace rtk
