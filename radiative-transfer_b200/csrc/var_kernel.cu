// var_kernel.cu -- cell-dependent material fields rho(c), T(c), kappa_g(c): EXTENSION, see var_kernel.cuh.
//
// What changes against the uniform slab: the stepper constants (alpha, gamma, the 2x2 inverse, source and time-level
// weights) depend on the cell, so the row recurrence is  v_j = alpha_j v_(j-1) + beta_j  with a cell-dependent factor:
// the scan carries PAIRS (A, W) = (product of alphas, local solution) and composes them as
// (A2, W2) o (A1, W1) = (A2 A1, A2 W1 + W2).  The per-cell constants are evaluated on the fly from five per-(group, cell)
// table entries with exactly the formulas of the uniform path (coef_from_material, sweep_device.cuh), the right-hand
// sides with exactly its stage_pass1.  First version: one CTA per chain walks its tiles in sequence (no look-back, no
// tickets, plain coalesced loads); with uniform fields it reproduces the uniform kernels to rounding.
#include "rt_common.cuh"
#include "sweep_device.cuh"
#include "var_kernel.cuh"

namespace rtk
{
// ------------------------------------------------------------------------------------------ per-cell group terms
__device__ __forceinline__ double pf_cell(double E, double T) // correction.cpp:11-22
{
   const double inv_h3c2 = 1.0 / ((H_KEV_SH * H_KEV_SH * H_KEV_SH) * (C_LIGHT * C_LIGHT));
   return K_JPK * (E * E * E) * inv_h3c2 / expm1(E / T);
}

// The reference's G-length Correction pipeline (correction.cpp:125-139 edge opacities, 162-249 edge stencils, 328-340
// correction terms; same operations as group_terms_kernel) with the cell's rho, T, kappa, for one (group, cell).  All
// per-cell tables are group-major ([g][c], cell fastest): neighbouring lanes are neighbouring cells of one group.
struct CellEval
{
   double sig, Bg, dEB, dsig, dkap;
};

__device__ __forceinline__ CellEval cell_terms_eval(const CellTermsArgs &a, int g, int c)
{
   const int G = a.G;
   const size_t N = (size_t)a.N;
   const double *e = a.e_edge;
   const double T = a.T_cell[c], rho = a.rho_cell[c];
   const double kg = a.kappa_cell[(size_t)g * N + c];
   // opacity at the lower / upper edge of the group (linear interpolation between the group centres)
   auto edge_opacity = [&](int ge, double k_below, double k_above) {
      const double ea = 0.5 * (e[ge] + e[ge + 1]), eam = 0.5 * (e[ge - 1] + e[ge]);
      const double wl = (ea - e[ge]) / (ea - eam), wr = (e[ge] - eam) / (ea - eam);
      return k_below * wl + k_above * wr;
   };
   const double ke_lo = (g == 0) ? kg : edge_opacity(g, a.kappa_cell[(size_t)(g - 1) * N + c], kg);
   const double ke_hi = (g == G - 1) ? kg : edge_opacity(g + 1, kg, a.kappa_cell[(size_t)(g + 1) * N + c]);
   const double epf_lo = (g == 0) ? 0.0 : e[g] * pf_cell(e[g], T);
   const double epf_hi = (g < G - 1 || g == 0) ? e[g + 1] * pf_cell(e[g + 1], T) : 0.0;
   const double de = e[g + 1] - e[g];
   CellEval r;
   if (g == 0)
   {
      r.dEB = epf_hi;
      r.dsig = ke_hi * e[1] / de;
      r.dkap = ke_hi * epf_hi;
   }
   else if (g < G - 1)
   {
      r.dEB = epf_hi - epf_lo;
      r.dsig = (ke_hi * e[g + 1] - ke_lo * e[g]) / de;
      r.dkap = ke_hi * epf_hi - ke_lo * epf_lo;
   }
   else
   {
      r.dEB = -epf_lo;
      r.dsig = 0.0; // set below
      r.dkap = -ke_lo * epf_lo;
   }
   if (g == G - 1) r.dsig = -ke_hi * e[G] / de; // for G == 1 this overwrites entry 0, like the reference
   r.sig = rho * kg;
   r.Bg = a.B_cell[(size_t)g * N + c];
   return r;
}

// one thread per (owned group, cell), cell fastest
__global__ void cell_terms_kernel(CellTermsArgs a)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= (long long)a.groups * a.N) return;
   const int gl = (int)(idx / a.N), c = (int)(idx - (long long)gl * a.N), g = a.G0 + gl;
   const CellEval r = cell_terms_eval(a, g, c);
   // both sweep-ordered copies (CellTables): sign 1 (mu > 0) at j = c, sign 0 (mu < 0) at j = N-1-c
   const size_t fwd = ((size_t)a.groups + gl) * a.pitch + c, rev = (size_t)gl * a.pitch + (a.N - 1 - c);
   const double srcB = r.sig * r.Bg, c2 = 3.0 * r.sig * r.Bg - r.dkap, c3 = r.dsig * (4.0 * r.Bg - r.dEB);
   a.sigma[fwd] = r.sig; a.sigma[rev] = r.sig;
   a.srcB[fwd] = srcB;   a.srcB[rev] = srcB;
   a.cor1[fwd] = r.dsig; a.cor1[rev] = r.dsig;
   a.cor2[fwd] = c2;     a.cor2[rev] = c2;
   a.cor3[fwd] = c3;     a.cor3[rev] = c3;
   if (a.sdB) a.sdB[(size_t)gl * a.pitch + c] = r.sig * a.dBdT_cell[(size_t)g * a.N + c];
}

// material state of the two boundary cells for ALL groups (the equilibrium boundary sources need them)
__global__ void boundary_state_kernel(CellTermsArgs a)
{
   const int idx = blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= 2 * a.G) return;
   const int g = idx >> 1, right = idx & 1;
   const CellEval r = cell_terms_eval(a, g, right ? a.N - 1 : 0);
   if (right) { a.B_right[g] = r.Bg; a.dEB_right[g] = r.dEB; }
   else { a.B_left[g] = r.Bg; a.dEB_left[g] = r.dEB; }
}

void launch_cell_terms(const CellTermsArgs &a, cudaStream_t st)
{
   const long long total = (long long)a.groups * a.N;
   cell_terms_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(a);
}

void launch_boundary_state(const CellTermsArgs &a, cudaStream_t st)
{
   boundary_state_kernel<<<(2 * a.G + 127) / 128, 128, 0, st>>>(a);
}

__global__ void boundary_table_var_kernel(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V,
                                          const double *__restrict__ mu, const double *__restrict__ B_left,
                                          const double *__restrict__ dEB_left, const double *__restrict__ B_right,
                                          const double *__restrict__ dEB_right, double *__restrict__ psi_source,
                                          double *__restrict__ bdry)
{
   const int idx = blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= M * G) return;
   const int i = idx % M, g = idx / M;
   const bool neg = mu[i] < 0.0;
   double src = psi_source[idx];
   if (use_mg_equilib)
   {
      // solver.cpp:297-314 with the state of the cell the boundary feeds into
      const double Bg = neg ? B_right[g] : B_left[g], dEBg = neg ? dEB_right[g] : dEB_left[g];
      double val = 4.0 * Bg - dEBg;
      val *= mu[i] * V / C_LIGHT;
      val += Bg;
      src = val;
      psi_source[idx] = val;
   }
   double b;
   if (neg) b = (bc_right == 1) ? src : 0.0;
   else b = (bc_left == 0 || bc_left == 1) ? src : 0.0; // vacuum falls through to source
   bdry[idx] = b;
}

void launch_boundary_table_var(int M, int G, int bc_left, int bc_right, int use_mg_equilib, double V, const double *mu,
                               const double *B_left, const double *dEB_left, const double *B_right, const double *dEB_right,
                               double *psi_source, double *bdry, cudaStream_t st)
{
   boundary_table_var_kernel<<<(M * G + 255) / 256, 256, 0, st>>>(M, G, bc_left, bc_right, use_mg_equilib, V, mu, B_left, dEB_left,
                                                                 B_right, dEB_right, psi_source, bdry);
}

__global__ void init_state_var_kernel(double2 *__restrict__ buf, const double *__restrict__ B_cell, int M, int N, int Np, int G,
                                      int G0, long long total)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= total) return;
   const int j = (int)(idx % Np);
   const int row = (int)(idx / Np);
   const int gl = row / M, i = row % M;
   double v = 0.0;
   if (j < N)
   {
      const int cell = (i < (M >> 1)) ? (N - 1 - j) : j; // rows are stored in sweep order
      v = B_cell[(size_t)(G0 + gl) * N + cell];
   }
   buf[idx] = make_double2(v, v);
}

void launch_init_state_var(double2 *buf, const double *B_cell, int M, int N, int Np, int G, int G0, int groups, cudaStream_t st)
{
   const long long total = (long long)groups * M * Np;
   init_state_var_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(buf, B_cell, M, N, Np, G, G0, total);
}

// ------------------------------------------------------------------------------------------ the stage sweep
constexpr int VT = 128, VC = 4, VTILE = VT * VC, VWARPS = VT / 32;

template <int KIND, bool CORR>
__global__ void __launch_bounds__(VT) var_sweep_kernel(const VarSweepArgs a)
{
   __shared__ double s_A[VWARPS], s_W[VWARPS]; // per-warp composition of the tile's cells
   __shared__ double s_carry;                  // outflow node of the last cell processed so far in the chain
   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const int M = a.M, N = a.N, half_m = M >> 1;
   const int chain = blockIdx.x;
   const int nseg = a.reflective_left ? 2 : 1;
   const double kbc = (KIND == KIND_BE) ? 1.0 : (KIND == KIND_CN ? 2.0 : 6.0);

   for (int seg = 0; seg < nseg; ++seg)
   {
      int row;
      if (a.reflective_left)
      {
         const int gl = chain / half_m, pp = chain - gl * half_m;
         row = gl * M + (seg ? half_m + pp : half_m - 1 - pp);
      }
      else row = chain;
      const int gl = row / M, i = row - gl * M, g = a.G0 + gl;
      const bool sgn = i >= half_m;
      const double mu = a.mu[i];
      const double2 *x1 = a.x1[sgn ? 1 : 0] + (size_t)row * a.Np;
      const double2 *x2 = a.x2 ? a.x2 + (size_t)row * a.Np : nullptr;
      const double2 *xc = a.xc ? a.xc + (size_t)row * a.Np : nullptr;
      double2 *out = a.out + (size_t)row * a.Np;
      const size_t trow = (size_t)gl * a.t.pitch;
      // inflow of the row's first cell: boundary value, or (reflective left) the mirrored row's last outflow node;
      // all three inflow carries equal it there (solver.cpp:695-697) -> factor kbc, no upstream old-level terms
      double tile_in = kbc * ((seg == 0) ? a.bdry[(size_t)i + (size_t)M * g] : s_carry);

      for (int tile_start = 0; tile_start < N; tile_start += VTILE)
      {
         const int j0 = tile_start + tid * VC;
         double alpha[VC], gamma[VC], beta[VC], U[VC];
         double A = 1.0, W = 0.0;
         double up1 = 0.0, up2 = 0.0;
         if (j0 > 0 && j0 < N)
         {
            if (KIND == KIND_BDF) up1 = x1[j0 - 1].y;
            if (KIND != KIND_BE) up2 = x2[j0 - 1].y;
         }
#pragma unroll
         for (int jj = 0; jj < VC; ++jj)
         {
            const int j = j0 + jj;
            alpha[jj] = 1.0; gamma[jj] = 0.0; beta[jj] = 0.0; U[jj] = 0.0; // identity element past the end of the row
            if (j < N)
            {
               const int cell = sgn ? j : N - 1 - j; // rows are stored in sweep order
               const size_t o = trow + cell;
               const RowCoef c = fold_coef<KIND, CORR>(coef_from_material(KIND, CORR, mu, a.t.sigma[o], a.t.srcB[o],
                                                                          CORR ? a.t.cor1[o] : 0.0, CORR ? a.t.cor2[o] : 0.0,
                                                                          CORR ? a.t.cor3[o] : 0.0, a.dx, a.tau, a.dt, a.beta));
               double2 X1[1], X2[1];
               double xs[1], wl[1], Uc[1];
               X1[0] = x1[j];
               X2[0] = (KIND != KIND_BE) ? x2[j] : X1[0];
               if (KIND == KIND_BDF)
               {
                  xs[0] = 0.0;
                  if (CORR) { const double2 s2 = xc[j]; xs[0] = s2.x + s2.y; }
               }
               else xs[0] = X1[0].x + X1[0].y;
               stage_pass1<KIND, CORR, 1>(c, X1, X2, xs, up1, up2, j == 0, wl, Uc);
               alpha[jj] = c.alpha; gamma[jj] = c.gamma; beta[jj] = wl[0]; U[jj] = Uc[0];
               W = fma(c.alpha, W, wl[0]);
               A *= c.alpha;
               if (KIND == KIND_BDF) up1 = X1[0].y;
               if (KIND != KIND_BE) up2 = X2[0].y;
            }
         }
         // ---- block scan of the (A, W) pairs: (A2, W2) o (A1, W1) = (A2 A1, A2 W1 + W2)
         double Ai = A, Wi = W;
#pragma unroll
         for (int d = 1; d < 32; d <<= 1)
         {
            const double Au = __shfl_up_sync(0xffffffffu, Ai, d), Wu = __shfl_up_sync(0xffffffffu, Wi, d);
            if (lane >= d)
            {
               Wi = fma(Ai, Wu, Wi);
               Ai *= Au;
            }
         }
         double Ae = __shfl_up_sync(0xffffffffu, Ai, 1), We = __shfl_up_sync(0xffffffffu, Wi, 1);
         if (lane == 0) { Ae = 1.0; We = 0.0; }
         if (lane == 31) { s_A[warp] = Ai; s_W[warp] = Wi; }
         __syncthreads();
         double Ap = 1.0, Wp = 0.0; // composition of the warps before this one
         for (int w2 = 0; w2 < warp; ++w2)
         {
            Wp = fma(s_A[w2], Wp, s_W[w2]);
            Ap *= s_A[w2];
         }
         const double Wx = fma(Ae, Wp, We), Ax = Ae * Ap; // everything upstream of this thread inside the tile
         // ---- pass 2: true values from the true inflow
         double v = fma(Ax, tile_in, Wx);
         const int last = min(N, tile_start + VTILE) - 1; // last cell of the row inside this tile
#pragma unroll
         for (int jj = 0; jj < VC; ++jj)
         {
            const int j = j0 + jj;
            if (j < N)
            {
               const double vn = fma(alpha[jj], v, beta[jj]);
               out[j] = make_double2(fma(gamma[jj], v, U[jj]), vn);
               v = vn;
            }
         }
         if (last >= j0 && last < j0 + VC) s_carry = v;
         __syncthreads();
         tile_in = s_carry;
      }
      // s_carry now holds the outflow node of the row's last cell: the mirrored row's inflow (reflective left boundary)
   }
}

template <int KIND>
static cudaError_t launch_var_kind(bool corr, const VarSweepArgs &a, int grid, cudaStream_t st)
{
   if (corr) var_sweep_kernel<KIND, true><<<grid, VT, 0, st>>>(a);
   else var_sweep_kernel<KIND, false><<<grid, VT, 0, st>>>(a);
   return cudaGetLastError();
}

cudaError_t launch_var_sweep(int kind, bool corr, const VarSweepArgs &a, cudaStream_t st, int *grid_out)
{
   const int grid = a.reflective_left ? a.rows / 2 : a.rows;
   if (grid_out) *grid_out = grid;
   if (kind == KIND_BE) return launch_var_kind<KIND_BE>(corr, a, grid, st);
   if (kind == KIND_CN) return launch_var_kind<KIND_CN>(corr, a, grid, st);
   if (kind == KIND_BDF) return launch_var_kind<KIND_BDF>(corr, a, grid, st);
   return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------------------------------ material-temperature update
// Group sums per cell, the one place where energy groups couple.  Thread per cell (the tables are [gl][c]: a warp reads
// 256 contiguous bytes per table and group), groups cut into chunks over blockIdx.y so that N-sized problems still fill
// the machine; the chunks are then added in ascending order (deterministic, no atomics).
__global__ void exchange_partial_kernel(ExchangeArgs a)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= a.N) return;
   const int g0 = blockIdx.y * EXCHANGE_GROUPS_PER_CHUNK, g1 = min(a.groups, g0 + EXCHANGE_GROUPS_PER_CHUNK);
   double A = 0.0, E = 0.0, D = 0.0;
   for (int gl = g0; gl < g1; ++gl)
   {
      const size_t o = (size_t)gl * a.pitch + c;
      A += a.sigma[o] * a.phi[(size_t)gl * a.N + c];
      E += a.W * a.srcB[o];
      D += a.W * a.sdB[o];
   }
   double *p = a.partial + (size_t)blockIdx.y * 3 * a.N;
   p[c] = A; p[a.N + c] = E; p[2 * (size_t)a.N + c] = D;
}

__global__ void exchange_reduce_kernel(int N, int chunks, const double *__restrict__ partial, double *__restrict__ sums)
{
   const int idx = blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= 3 * N) return;
   double v = 0.0;
   for (int k = 0; k < chunks; ++k) v += partial[(size_t)k * 3 * N + idx];
   sums[idx] = v;
}

int launch_exchange_sums(const ExchangeArgs &a, cudaStream_t st)
{
   const int chunks = exchange_chunks(a.groups);
   exchange_partial_kernel<<<dim3((a.N + 127) / 128, chunks), 128, 0, st>>>(a);
   exchange_reduce_kernel<<<(3 * a.N + 255) / 256, 256, 0, st>>>(a.N, chunks, a.partial, a.sums);
   return 2;
}

// first_bad: 0 until some cell's new temperature is not a positive finite number, then the (1-based) number of the update
// that produced it -- the explicitly coupled update is only conditionally stable (stiff slabs at large dt run away, in the
// oracle's restatement exactly as here), and a solve must not carry on silently with T <= 0 or NaN
__global__ void temperature_update_kernel(int N, int bdf2, double dt, double cv, const double *__restrict__ rho_cell,
                                          const double *__restrict__ sums, double *__restrict__ T, double *__restrict__ T_prev,
                                          int update_no, int *__restrict__ first_bad)
{
   const int c = blockIdx.x * blockDim.x + threadIdx.x;
   if (c >= N) return;
   const double A = sums[c], E = sums[N + c], D = sums[2 * (size_t)N + c];
   const double rcv = rho_cell[c] * cv, t = T[c];
   double tn;
   if (!bdf2) tn = t + dt * (A - E) / (rcv + dt * D);
   else tn = (rcv * (2. * t - 0.5 * T_prev[c]) + dt * (A - E + D * t)) / (1.5 * rcv + dt * D);
   T_prev[c] = t;
   T[c] = tn;
   if (!(tn > 0.0) || !isfinite(tn)) atomicCAS(first_bad, 0, update_no);
}

void launch_temperature_update(int N, bool bdf2, double dt, double cv, const double *rho_cell, const double *sums, double *T,
                               double *T_prev, int update_no, int *first_bad, cudaStream_t st)
{
   temperature_update_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, bdf2 ? 1 : 0, dt, cv, rho_cell, sums, T, T_prev, update_no, first_bad);
}

// ------------------------------------------------------------------------------------------ diagnostics
__device__ __forceinline__ double block_sum_var(double v, double *red)
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

// One block per local group: diagnostics_kernel (moments.cu) with sigma and T taken per cell
__global__ void diagnostics_var_kernel(const double2 *__restrict__ ends, const double *__restrict__ phi, const double *__restrict__ mu,
                                       const double *__restrict__ wt, const double *__restrict__ sigma_cell, int pitch,
                                       const double *__restrict__ T_cell, const double *__restrict__ de_ave, int M, int N, int Np,
                                       int G0, double dx, double *__restrict__ balance, double *__restrict__ left_ends,
                                       double *__restrict__ right_ends)
{
   __shared__ double red[32];
   const int gl = blockIdx.x, g = G0 + gl, half_m = M >> 1;
   const double ac = RAD_A * C_LIGHT;
   double part = 0.0, emit = 0.0;
   const double *prow = phi + (size_t)gl * N, *srow = sigma_cell + (size_t)gl * pitch;
   for (int c = threadIdx.x; c < N; c += blockDim.x)
   {
      const double T = T_cell[c];
      part += srow[c] * prow[c] * dx;
      emit += srow[c] * ac * (T * T) * (T * T) * dx;
   }
   const double absorbed = block_sum_var(part, red);
   const double src = block_sum_var(emit, red);
   if (threadIdx.x == 0)
   {
      double j_half_minus = 0., j_half_plus = 0., jN_half_minus = 0., jN_half_plus = 0., l = 0., r = 0.;
      for (int m = 0; m < M; ++m)
      {
         const double2 *row = ends + (size_t)(gl * M + m) * Np;
         const double v_first = row[0].y, v_last = row[N - 1].y;
         const double mw = mu[m] * wt[m];
         if (m < half_m)
         {
            j_half_minus -= v_last * mw;
            jN_half_minus -= v_first * mw;
            l += v_last;
         }
         else
         {
            j_half_plus += v_first * mw;
            jN_half_plus += v_last * mw;
            r += v_last;
         }
      }
      const double sources = j_half_plus + jN_half_minus + src;
      const double sinks = jN_half_plus + j_half_minus + absorbed;
      balance[gl] = fabs(sinks - sources) / sources;
      left_ends[gl] = l / (de_ave[g] * C_LIGHT);
      right_ends[gl] = r / (de_ave[g] * C_LIGHT);
   }
}

void launch_diagnostics_var(const double2 *ends, const double *phi, const double *mu, const double *wt, const double *sigma_cell,
                            int pitch, const double *T_cell, const double *de_ave, int M, int N, int Np, int G0, int groups, double dx,
                            double *balance, double *left_ends, double *right_ends, cudaStream_t st)
{
   diagnostics_var_kernel<<<groups, 1024, 0, st>>>(ends, phi, mu, wt, sigma_cell, pitch, T_cell, de_ave, M, N, Np, G0, dx, balance,
                                                  left_ends, right_ends);
}
} // namespace rtk
