// rt_capi.cu -- host side of librt_cuda.so: the C ABI declared in include/rt_cuda.h.
//
// Owns the device tables, the four rotating time-level buffers, the look-back state and the stage
// schedule of rt::Solver::solve (reference src/solver.cpp:590-823).  No CPU compute path exists:
// every numerical result comes from the kernels in group_terms.cu, sweep_kernel.cu and moments.cu.
#include "../../include/rt_cuda.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "group_terms.cuh"
#include "moments.cuh"
#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "var_kernel.cuh"
#include "sweep_kernel.cuh"

using namespace rtk;

static thread_local std::string g_err;

static int fail(const char *fmt, ...)
{
   char buf[1024];
   va_list ap;
   va_start(ap, fmt);
   vsnprintf(buf, sizeof buf, fmt, ap);
   va_end(ap);
   g_err = buf;
   return 1;
}

#define CK(call)                                                                                              \
   do                                                                                                         \
   {                                                                                                          \
      cudaError_t e_ = (call);                                                                                \
      if (e_ != cudaSuccess) return fail("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
   } while (0)

struct EventPair
{
   cudaEvent_t a, b;
   int kind; // 0..2 sweep kinds, 3 moments, 4 fused BDF2 step
};

struct rt_solver
{
   rt_params p;
   std::vector<double> mu, wt, e_edge, kappa, psi_source;
   int num_sms = 0;
   cudaStream_t st = nullptr, copy_st = nullptr;
   // G-length tables
   double *d_mu = nullptr, *d_wt = nullptr, *d_e_edge = nullptr, *d_kappa = nullptr, *d_T = nullptr;
   double *d_B = nullptr, *d_dBdT = nullptr, *d_kappa_edge = nullptr, *d_epf = nullptr, *d_dEB = nullptr;
   double *d_dsig = nullptr, *d_dkap = nullptr, *d_cor1 = nullptr, *d_cor2 = nullptr, *d_cor3 = nullptr;
   double *d_sigma = nullptr, *d_srcB = nullptr, *d_de_ave = nullptr, *d_sums = nullptr;
   double *d_psi_source = nullptr, *d_bdry = nullptr;
   double *d_coef[3] = {nullptr, nullptr, nullptr}; // per stage kind: Gl*M entries of COEF_ENTRY_DOUBLES doubles
   double *d_coef_step = nullptr;                   // the three kinds packed per row (STEP_COEF_DOUBLES each)
   double *d_coef4 = nullptr;                       // the same + chunk transfer matrices and responses (STEP4_COEF_DOUBLES each, step4_kernel.cu)
   double sums[3] = {0, 0, 0};
   // cell-dependent material fields (extension, var_kernel.cu): inputs, Planck per cell, per-cell tables [gl][c]
   bool var = false;
   std::vector<double> rho_cell, T_cell, kappa_cell; // kappa_cell: GROUP-MAJOR rows [tab_g0, tab_g0 + tab_n) only
   // Per-cell tables are sharded with the groups: kappa of the owned groups and one neighbour on either side (the edge
   // opacities interpolate between group centres, correction.cpp:125-139), B and dB/dT of the owned groups only.  The
   // kernels index them with the GLOBAL group number through base pointers (kap_base() etc.).  The equilibrium boundary
   // sources need ALL groups, but only at the two boundary cells: compact [G][2] copies (bd_*).
   int tab_g0 = 0, tab_n = 0;
   std::vector<double> bd_kappa;                                   // [G][2]: kappa_g at cell 0 / cell N-1
   double *d_bd_T = nullptr, *d_bd_rho = nullptr, *d_bd_kappa = nullptr, *d_bd_B = nullptr, *d_bd_dB = nullptr;
   int var_scan_steps = 5; // warp-scan steps the fused cell-dependent step has to run (launch_var_scan_steps; RT_CUDA_VAR_SCAN_STEPS=5 forces all)
   double *kap_base() const { return d_kappa_c - (ptrdiff_t)tab_g0 * p.N; }
   double *B_base() const { return d_B_c - (ptrdiff_t)p.group_begin * p.N; }
   double *dB_base() const { return d_dBdT_c - (ptrdiff_t)p.group_begin * p.N; }
   double *d_rho_c = nullptr, *d_T_c = nullptr, *d_kappa_c = nullptr, *d_B_c = nullptr, *d_dBdT_c = nullptr;
   double *d_vt[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; // sigma, srcB, cor1, cor2, cor3: [sign][gl][vt_pitch] (var_kernel.cuh)
   int vt_pitch = 0;               // N rounded up to 16
   double *vt_fwd(int k) const { return d_vt[k] + (size_t)Gl * vt_pitch; } // the forward (mu > 0) copy
   MatMaps mat_maps;
   int var_step_mat_slots = 1;     // RT_CUDA_VAR_STEP_MATBUF=1|2 (fused cell-dependent step; 1 = one more resident CTA, measured +10 %)
   int var_threads = 64, var_cpt = 4; // tile shape of var_tile_kernel: RT_CUDA_VAR_THREADS=64|128 (0 = the chain-walk kernel), RT_CUDA_VAR_CPT=4|8
   // fused cell-dependent step from warp tiles (var_warp_kernel.cu): RT_CUDA_VAR_WARP=1|0, RT_CUDA_VAR_CHAIN=2|4|8 warps per CTA tile
   bool var_warp = false;
   int var_chain = 4;
   SweepMaps maps_vwarp;
   MatMaps mat_maps_vwarp;
   double *d_edge_state[4] = {nullptr, nullptr, nullptr, nullptr};  // B_left, dEB_left, B_right, dEB_right (G each)
   // material-temperature update (extension; rt_params::cv > 0)
   bool tupdate = false, T_moved = false;
   double wsum = 0.0;              // sum of the quadrature weights, ascending
   double *d_sdB = nullptr;        // [gl][c] sigma_g dB_g/dT
   double *d_T_prev = nullptr, *d_xsums = nullptr, *d_xpart = nullptr; // N ; 3N (A, E, D) ; chunks*3N
   double *d_ppart = nullptr;      // 2N: per-cell partial sums of the raw Planck integrals (sharded evaluation)
   long long tsteps_done = 0;      // temperature updates since the state was initialised
   int *d_T_bad = nullptr;         // 0, or the number of the first update that left some cell with T <= 0 or a non-finite T
   rt_allreduce_fn allreduce = nullptr;
   void *allreduce_user = nullptr;
   void *nccl_comm = nullptr;      // communicator created by rt_set_nccl (owned by the handle) ; rt_multi owns its own
   // batching
   int Gl = 0, Gb = 0, n_batches = 0, Np = 0;
   size_t buf_elems = 0; // double2 elements per time-level buffer
   double2 *E[4] = {nullptr, nullptr, nullptr, nullptr};
   int plan_levels = 4;  // levels the batch size was planned for
   int levels = 0;       // time-level buffers allocated so far: 2 for the fused step / BE / CN time stepping, the other two on demand
   SweepMaps maps[3]; // per stage kind (tile shapes may differ): tensor maps of E[0..3]
   SweepMaps maps_var; // tile shape of var_tile_kernel
   int i_ends = 0, i_prev = -1, i_half_neg = -1, i_half_pos = -1;
   int batch_start = 0, batch_groups = 0; // local group range of the batch currently on the device
   int stages_done = 0;
   // look-back state
   double2 *d_tile_status = nullptr;
   unsigned long long *d_ticket = nullptr;
   unsigned long long ticket_base = 0;
   unsigned epoch = 0;
   int last_grid = 0, last_block = 0;
   // tile shape per stage kind {BE, CN, BDF}: CTA size, cells per thread, X1 ring depth (0 = automatic)
   // (measured on C4, gpurun exp8/exp9: 64 x 8 is best or equal for all three kinds)
   int threads[3] = {64, 64, 64}, cpt[3] = {8, 8, 8}, ring[3] = {0, 0, 0};
   int tile_cells[3] = {512, 512, 512}, tiles_per_row[3] = {0, 0, 0}, max_tiles_per_row = 0;
   // results retained across batches (group-major [gl][c]) and per-group diagnostics
   double *d_phi = nullptr, *d_phip = nullptr, *d_F = nullptr;
   double *d_balance = nullptr, *d_left = nullptr, *d_right = nullptr;
   bool results_valid = false;
   double *d_stage = nullptr; // staging for layout conversion
   size_t stage_elems = 0;
   // per-step output
   double *step_phi_host = nullptr;
   int step_group_major = 1;
   double *d_step_phi[2] = {nullptr, nullptr};
   cudaEvent_t step_ev[2] = {nullptr, nullptr};
   int step_parity = 0;
   // scalar intensity of every step from the fused step kernel itself (step4_kernel.cu, PHI): partial sums per run of rows
   double *d_phi_part = nullptr;
   long long phi_pitch = 0;   // cells per partial row (whole tiles)
   int phi_run = 1;           // rows per run = tickets per draw: the largest divisor of M/2 that is <= 8
   bool step_phi_fused = false; // the step just completed has left its partial sums in d_phi_part
   // per-step history (rt_set_step_history): radiation energy and material temperature of every time step, kept on the
   // device ([hist_cap][N] each) and delivered when the call that ran the steps returns
   double *hist_E_host = nullptr, *hist_T_host = nullptr;
   double *d_hist_E = nullptr, *d_hist_T = nullptr;
   int hist_cap = 0, hist_lo = 0, hist_hi = 0; // rows [hist_lo, hist_hi) were written since the last delivery
   // timing
   rt_timing tm;
   std::vector<EventPair> ev_pool;
   size_t ev_used = 0;
   bool timing_on = true;
   // fused BDF2-step kernel (step_kernel.cu): RT_CUDA_FUSED=0 runs every step stage by stage
   bool fused = true, fused_var = false, half_stale = false;
   // fused uniform step: 0 = automatic (default): the four-scan kernel (step_warp_kernel.cu, ~2 % faster on C4) for plain steps,
   // the one-scan kernel (step4_kernel.cu) when the step also has to deliver phi (it adds up w_m psi in the same pass and
   // saves the separate moments pass over the state) ; RT_CUDA_STEP_KERNEL=1 | 4 forces one of them
   int step_kernel = 0;
   bool step4_ok = true;       // the tile shape in use is one the one-scan kernel supports
   int step_threads = 32, step_cpt = 8, step_chain = 4, step_box_cells = 256, step_tile_cells = 1024, step_tiles_per_row = 0;
   SweepMaps maps_step;
   int ticket_batch = 8;       // tickets per atomic draw (RT_CUDA_TICKET_BATCH)
   bool l2_prefetch = true;    // RT_CUDA_L2_PREFETCH=0 switches the two-tiles-ahead L2 prefetch off
   bool step_l2_prefetch = false; // fused step: every warp already has a tile in flight; RT_CUDA_STEP_L2_PREFETCH=1 turns it on
   bool lookback_walk = false; // RT_CUDA_LOOKBACK=walk: aggregate walk-back (faster with very few rows, not bitwise reproducible)
};

// ---------------------------------------------------------------------------------------------
extern "C" const char *rt_last_error(void) { return g_err.c_str(); }
extern "C" int rt_abi_version(void) { return RT_ABI_VERSION; }

// 12-point Gauss-Legendre rule computed like the reference's Planck::gauss_quad_setup
// (src/Planck.cpp:231-337): long double Newton iteration with the derivative scaled by 13/12 (the loop
// counter after the recurrence), weights renormalised to sum to 2.  Rounded to fp64 for the device.
static void gauss12_host(double *pt, double *w)
{
   const int order = 12;
   long double lp[order], lw[order], wsum = 0;
   for (int i = 0; i < (order + 1) / 2; i++)
   {
      long double mu = std::cos((double)3.1415926546 * (i + 0.75) / (order + 0.5)), old_mu, pj = 0, pjm1 = 0, pjm2 = 0, dp = 0;
      bool conv = false;
      while (!conv)
      {
         pjm1 = 0;
         pj = 1;
         int j;
         for (j = 1; j <= order; j++)
         {
            pjm2 = pjm1;
            pjm1 = pj;
            pj = ((2 * j - 1) * mu * pjm1 - (j - 1) * pjm2) / j;
         }
         dp = j * (mu * pj - pjm1) / (mu * mu - 1);
         old_mu = mu;
         mu = old_mu - pj / dp;
         if (fabsl(mu - old_mu) < 2.220446049250313e-16L) conv = true;
      }
      lp[i] = -mu;
      lp[order - 1 - i] = mu;
      lw[i] = 1 / ((1 - mu * mu) * dp * dp);
      lw[order - 1 - i] = lw[i];
      wsum += lw[i] + lw[order - 1 - i];
   }
   for (int i = 0; i < order; i++)
   {
      pt[i] = (double)lp[i];
      w[i] = (double)(lw[i] * 2 / wsum);
   }
}

template <typename T>
static int dalloc(T **p, size_t n)
{
   CK(cudaMalloc((void **)p, std::max<size_t>(n, 1) * sizeof(T)));
   return 0;
}

// The large state buffers come from the device's stream-ordered memory pool with the release threshold lifted: a
// solver created after another one was destroyed (same process) gets the same physical memory back at once instead
// of paying the driver's unmap / remap of >100 GB (measured: 0.03 s ... 0.5 s per rt_create otherwise).
// RT_CUDA_POOL_KEEP=0 returns the memory to the driver on rt_destroy.
static bool pool_keep()
{
   const char *e = getenv("RT_CUDA_POOL_KEEP");
   return !(e && e[0] == '0');
}

static int pool_setup(int device)
{
   cudaMemPool_t pool;
   CK(cudaDeviceGetDefaultMemPool(&pool, device));
   unsigned long long thr = pool_keep() ? ~0ull : 0ull;
   CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
   return 0;
}

// bytes the pool holds but nobody uses: available to this process although cudaMemGetInfo does not count them as free
static size_t pool_idle_bytes(int device)
{
   cudaMemPool_t pool;
   unsigned long long reserved = 0, used = 0;
   if (cudaDeviceGetDefaultMemPool(&pool, device) != cudaSuccess) return 0;
   if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) != cudaSuccess) return 0;
   if (cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) != cudaSuccess) return 0;
   return reserved > used ? (size_t)(reserved - used) : 0;
}

template <typename T>
static int pool_alloc(rt_solver *s, T **p, size_t n)
{
   CK(cudaMallocAsync((void **)p, std::max<size_t>(n, 1) * sizeof(T), s->st));
   return 0;
}

static int upload(rt_solver *s, double *dst, const double *src, size_t n)
{
   CK(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyHostToDevice, s->st));
   s->tm.h2d_bytes += (long long)(n * sizeof(double));
   return 0;
}

static int download(rt_solver *s, double *dst, const double *src, size_t n)
{
   CK(cudaMemcpyAsync(dst, src, n * sizeof(double), cudaMemcpyDeviceToHost, s->st));
   CK(cudaStreamSynchronize(s->st));
   s->tm.d2h_bytes += (long long)(n * sizeof(double));
   return 0;
}

static int flush_events(rt_solver *s)
{
   if (s->ev_used == 0) return 0;
   CK(cudaStreamSynchronize(s->st));
   for (size_t k = 0; k < s->ev_used; ++k)
   {
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, s->ev_pool[k].a, s->ev_pool[k].b));
      if (s->ev_pool[k].kind < 3) s->tm.sweep_ms[s->ev_pool[k].kind] += ms;
      else if (s->ev_pool[k].kind == 4) s->tm.step_ms += ms;
      else if (s->ev_pool[k].kind == 5) s->tm.tupdate_ms += ms;
      else if (s->ev_pool[k].kind == 6) s->tm.tables_ms += ms;
      else s->tm.moments_ms += ms;
   }
   s->ev_used = 0;
   return 0;
}

static int event_begin(rt_solver *s, int kind, EventPair **out)
{
   *out = nullptr;
   if (!s->timing_on) return 0;
   if (s->ev_used == s->ev_pool.size())
   {
      if (s->ev_pool.size() >= 2048) { if (flush_events(s)) return 1; }
      else
      {
         EventPair ep;
         CK(cudaEventCreate(&ep.a));
         CK(cudaEventCreate(&ep.b));
         ep.kind = 0;
         s->ev_pool.push_back(ep);
      }
   }
   EventPair *ep = &s->ev_pool[s->ev_used++];
   ep->kind = kind;
   CK(cudaEventRecord(ep->a, s->st));
   *out = ep;
   return 0;
}

static int event_end(rt_solver *s, EventPair *ep)
{
   if (ep) CK(cudaEventRecord(ep->b, s->st));
   return 0;
}

// Planck integrals, edge stencils, correction terms, validation sums, boundary table -- all on the device.
// extension: Planck integrals per (cell, group) and the Correction pipeline per cell at the current T(c); with
// `with_boundary` also the boundary sources from the boundary cells (they overwrite the uniform boundary table and are
// built ONCE: a temperature update leaves the incoming boundary intensities alone, like the oracle)
static int compute_var_tables(rt_solver *s, bool with_boundary, bool sharded = false)
{
   const rt_params &p = s->p;
   if (sharded)
   {
      // a group shard inside a temperature-coupled solve: integrate the owned groups only; the last group's grey
      // remainder needs the sum over all groups -> second (2N) all-reduce of the step
      s->tm.kernel_launches += launch_planck_shard_begin(p.G, p.N, p.group_begin, s->Gl, s->d_e_edge, s->d_T_c, s->B_base(), s->dB_base(),
                                                         s->d_ppart, s->st);
      CK(cudaGetLastError());
      const int rc = s->allreduce(s->d_ppart, (size_t)2 * p.N, (void *)s->st, s->allreduce_user);
      if (rc) return fail("rt: the all-reduce hook failed with code %d in the Planck remainder after stage %d", rc, s->stages_done);
      if (p.group_begin + s->Gl == p.G)
         s->tm.kernel_launches += launch_planck_shard_finish(p.G, p.N, s->d_T_c, s->d_ppart, s->B_base(), s->dB_base(), s->st);
   }
   else
   {
   // the last group keeps "the previous value" when the grey remainder is not positive (Planck.cpp:73-76); the oracle
   // starts every evaluation from zeroed vectors
   if (p.group_begin + s->Gl == p.G)
   {
      CK(cudaMemsetAsync(s->B_base() + (size_t)(p.G - 1) * p.N, 0, (size_t)p.N * sizeof(double), s->st));
      CK(cudaMemsetAsync(s->dB_base() + (size_t)(p.G - 1) * p.N, 0, (size_t)p.N * sizeof(double), s->st));
   }
   // the owned groups only; the owner of the last group integrates the others on the fly for the remainder (no table of them)
   s->tm.kernel_launches += launch_planck_range(p.G, p.N, p.group_begin, s->Gl, s->d_e_edge, s->d_T_c, s->B_base(), s->dB_base(), s->st);
   }
   CellTermsArgs ca;
   ca.G = p.G; ca.N = p.N; ca.G0 = p.group_begin; ca.groups = s->Gl;
   ca.e_edge = s->d_e_edge; ca.kappa_cell = s->kap_base(); ca.rho_cell = s->d_rho_c; ca.T_cell = s->d_T_c; ca.B_cell = s->B_base();
   ca.dBdT_cell = s->dB_base();
   ca.sigma = s->d_vt[0]; ca.srcB = s->d_vt[1]; ca.cor1 = s->d_vt[2]; ca.cor2 = s->d_vt[3]; ca.cor3 = s->d_vt[4];
   ca.pitch = s->vt_pitch;
   ca.sdB = s->d_sdB;
   ca.B_left = s->d_edge_state[0]; ca.dEB_left = s->d_edge_state[1]; ca.B_right = s->d_edge_state[2]; ca.dEB_right = s->d_edge_state[3];
   launch_cell_terms(ca, s->st);
   s->tm.kernel_launches += 1;
   if (with_boundary && s->fused_var && p.ts_method == 3)
   {
      // once per solver (sigma does not change during a solve): can the later steps of the warp scans matter for any row?
      const char *se = getenv("RT_CUDA_VAR_SCAN_STEPS");
      s->var_scan_steps = 5;
      if (!(se && atoi(se) == 5))
      {
         int *d_n = nullptr, n = 5;
         CK(cudaMalloc(&d_n, sizeof(int)));
         launch_var_scan_steps(s->vt_fwd(0), s->vt_pitch, p.N, s->Gl, s->d_mu, p.M, p.dx, p.dt / 2., p.dt, p.V / C_LIGHT, 4, d_n, s->st);
         s->tm.kernel_launches += 1;
         const cudaError_t e = cudaMemcpyAsync(&n, d_n, sizeof(int), cudaMemcpyDeviceToHost, s->st);
         if (e == cudaSuccess) cudaStreamSynchronize(s->st);
         cudaFree(d_n);
         if (e != cudaSuccess) return fail("rt: scan-step evaluation failed: %s", cudaGetErrorString(e));
         s->var_scan_steps = n;
      }
   }
   if (with_boundary)
   {
      // all G groups at the two boundary cells: Planck integrals at their two temperatures, then the same per-cell pipeline
      // on the compact [G][2] tables (column 0 = cell 0, column 1 = cell N-1)
      CK(cudaMemsetAsync(s->d_bd_B + (size_t)(p.G - 1) * 2, 0, 2 * sizeof(double), s->st));
      CK(cudaMemsetAsync(s->d_bd_dB + (size_t)(p.G - 1) * 2, 0, 2 * sizeof(double), s->st));
      s->tm.kernel_launches += launch_planck_range(p.G, 2, 0, p.G, s->d_e_edge, s->d_bd_T, s->d_bd_B, s->d_bd_dB, s->st);
      CellTermsArgs cb = ca;
      cb.N = 2; cb.kappa_cell = s->d_bd_kappa; cb.rho_cell = s->d_bd_rho; cb.T_cell = s->d_bd_T; cb.B_cell = s->d_bd_B; cb.dBdT_cell = s->d_bd_dB;
      launch_boundary_state(cb, s->st);
      s->tm.kernel_launches += 1;
      if (upload(s, s->d_psi_source, s->psi_source.data(), (size_t)p.M * p.G)) return 1; // the caller's sources, before the BC switch
      launch_boundary_table_var(p.M, p.G, p.bc_left, p.bc_right, p.use_mg_equilib, p.V, s->d_mu, s->d_edge_state[0], s->d_edge_state[1],
                                s->d_edge_state[2], s->d_edge_state[3], s->d_psi_source, s->d_bdry, s->st);
      s->tm.kernel_launches += 1;
   }
   CK(cudaGetLastError());
   return 0;
}

static int compute_group_tables(rt_solver *s)
{
   const rt_params &p = s->p;
   s->tm.kernel_launches += launch_planck(p.G, 1, s->d_e_edge, s->d_T, s->d_B, s->d_dBdT, s->st);
   GroupTermsArgs a;
   a.G = p.G; a.T = p.T; a.rho = p.rho;
   a.e_edge = s->d_e_edge; a.kappa = s->d_kappa; a.B = s->d_B; a.dBdT = s->d_dBdT;
   a.kappa_edge = s->d_kappa_edge; a.epf = s->d_epf; a.dEB = s->d_dEB; a.dsigEdE = s->d_dsig; a.dkapEB = s->d_dkap;
   a.cor1 = s->d_cor1; a.cor2 = s->d_cor2; a.cor3 = s->d_cor3; a.sigma = s->d_sigma; a.srcB = s->d_srcB;
   a.de_ave = s->d_de_ave; a.sums = s->d_sums;
   s->tm.kernel_launches += launch_group_terms(a, s->st);
   s->tm.kernel_launches += launch_boundary_table(p.M, p.G, p.bc_left, p.bc_right, p.use_mg_equilib, p.V, s->d_mu, s->d_B,
                                                  s->d_dEB, s->d_psi_source, s->d_bdry, s->st);
   CK(cudaGetLastError());
   if (download(s, s->sums, s->d_sums, 3)) return 1;
   // stepper constants per (stage kind, owned group, direction); they never change during a solve
   CoefArgs c;
   c.sigma = s->d_sigma; c.srcB = s->d_srcB; c.cor1 = s->d_cor1; c.cor2 = s->d_cor2; c.cor3 = s->d_cor3;
   c.mu = s->d_mu; c.bdry = s->d_bdry;
   c.M = p.M; c.G0 = p.group_begin; c.groups = s->Gl;
   c.dx = p.dx; c.dt = p.dt; c.beta = p.V / C_LIGHT;
   c.tau = (p.ts_method == 3) ? p.dt / 2. : p.dt; // solver.cpp:708-806: BDF2 stages all advance dt/2
   const bool corr = p.use_correction && p.V != 0.0;
   for (int kind = 0; kind < 3; ++kind)
   {
      launch_coef_table(kind, corr, c, s->d_coef[kind], s->st);
      s->tm.kernel_launches++;
   }
   launch_pack_step_coef(s->d_coef[0], s->d_coef[1], s->d_coef[2], s->d_coef_step, s->Gl * p.M, corr, s->st);
   launch_pack_step4_coef(s->d_coef[0], s->d_coef[1], s->d_coef[2], s->d_wt, s->d_coef4, s->Gl * p.M, p.M, corr, s->st);
   s->tm.kernel_launches += 2;
   CK(cudaGetLastError());
   if (s->var && compute_var_tables(s, true)) return 1;
   return 0;
}

static bool sums_ok(const rt_solver *s)
{
   const double ac = RAD_A * C_LIGHT, T = s->p.T;
   const double acT4 = ac * std::pow(T, 4), dacT4 = 4.0 * ac * std::pow(T, 3);
   if (std::fabs(acT4 - s->sums[0]) > VALIDATION_TOL || std::fabs(dacT4 - s->sums[1]) > VALIDATION_TOL) return false;
   if (std::fabs(s->sums[2] - s->p.kappa_grey * acT4) > VALIDATION_TOL) return false;
   return true;
}

// psi = ends = B_g for the batch starting at local group `start` (solver.cpp:166-181)
static int load_batch(rt_solver *s, int start)
{
   s->batch_start = start;
   s->batch_groups = std::min(s->Gb, s->Gl - start);
   s->i_ends = 0;
   s->i_prev = s->i_half_neg = s->i_half_pos = -1;
   s->half_stale = false;
   s->stages_done = 0;
   if (s->tupdate)
   {
      // a new run starts from the caller's T(c): rebuild the per-cell tables if an update has moved it
      if (s->T_moved)
      {
         if (upload(s, s->d_T_c, s->T_cell.data(), s->p.N)) return 1;
         if (compute_var_tables(s, false)) return 1;
         s->T_moved = false;
      }
      s->tsteps_done = 0;
      CK(cudaMemsetAsync(s->d_T_bad, 0, sizeof(int), s->st));
   }
   if (s->var) launch_init_state_var(s->E[0], s->B_base(), s->p.M, s->p.N, s->Np, s->p.G, s->p.group_begin + start, s->batch_groups, s->st);
   else launch_init_state(s->E[0], s->d_B, s->p.M, s->p.N, s->Np, s->p.group_begin + start, s->batch_groups, s->st);
   s->tm.kernel_launches += 1;
   CK(cudaGetLastError());
   return 0;
}

static void release_nccl(rt_solver *s); // (NCCL section at the end of the file)

// time-level buffers are allocated on demand: two for the fused step, four when a step runs stage by stage
static int ensure_levels(rt_solver *s, int n)
{
   for (int k = s->levels; k < n; ++k)
   {
      if (pool_alloc(s, &s->E[k], s->buf_elems))
         return fail("%s (time-level buffer %d of %d: %zu bytes -- half_ends / a stage-by-stage step need four levels; create the solver with "
                     "RT_CUDA_LEVELS=4 or fewer groups per batch)", g_err.c_str(), k + 1, n, s->buf_elems * sizeof(double2));
      const int rows = s->Gb * s->p.M;
      for (int kind = 0; kind < 3; ++kind) CK(make_tile_map(&s->maps[kind].m[k], s->E[k], rows, s->Np, s->tile_cells[kind]));
      CK(make_tile_map(&s->maps_step.m[k], s->E[k], rows, s->Np, s->step_box_cells));
      if (s->var && s->var_threads > 0) CK(make_tile_map(&s->maps_var.m[k], s->E[k], rows, s->Np, s->var_threads * s->var_cpt));
      if (s->var && s->var_warp) CK(make_tile_map(&s->maps_vwarp.m[k], s->E[k], rows, s->Np, var_warp_subtile_cells()));
      s->levels = k + 1;
   }
   return 0;
}

extern "C" int rt_destroy(rt_solver *s)
{
   if (!s) return 0;
   cudaSetDevice(s->p.device);
   if (s->st) cudaStreamSynchronize(s->st);
   double *tabs[] = {s->d_mu, s->d_wt, s->d_e_edge, s->d_kappa, s->d_T, s->d_B, s->d_dBdT, s->d_kappa_edge, s->d_epf,
                     s->d_dEB, s->d_dsig, s->d_dkap, s->d_cor1, s->d_cor2, s->d_cor3, s->d_sigma, s->d_srcB, s->d_de_ave,
                     s->d_sums, s->d_psi_source, s->d_bdry, s->d_phi, s->d_phip, s->d_F,
                     s->d_balance, s->d_left, s->d_right, s->d_stage, s->d_step_phi[0], s->d_step_phi[1],
                     s->d_coef[0], s->d_coef[1], s->d_coef[2], s->d_coef_step, s->d_coef4, s->d_phi_part, s->d_hist_E, s->d_hist_T};
   for (double *t : tabs) if (t) cudaFree(t);
   double *vtabs[] = {s->d_rho_c, s->d_T_c, s->d_kappa_c, s->d_B_c, s->d_dBdT_c, s->d_vt[0], s->d_vt[1], s->d_vt[2], s->d_vt[3],
                      s->d_vt[4], s->d_edge_state[0], s->d_edge_state[1], s->d_edge_state[2], s->d_edge_state[3],
                      s->d_sdB, s->d_T_prev, s->d_xsums, s->d_xpart, s->d_ppart, s->d_bd_T, s->d_bd_rho, s->d_bd_kappa, s->d_bd_B, s->d_bd_dB};
   for (double *t : vtabs) if (t) cudaFree(t);
   for (int k = 0; k < 4; ++k) if (s->E[k]) cudaFreeAsync(s->E[k], s->st);
   if (s->st) cudaStreamSynchronize(s->st);
   if (!pool_keep())
   {
      cudaMemPool_t pool;
      if (cudaDeviceGetDefaultMemPool(&pool, s->p.device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
   }
   if (s->d_T_bad) cudaFree(s->d_T_bad);
   if (s->d_tile_status) cudaFree(s->d_tile_status);
   if (s->d_ticket) cudaFree(s->d_ticket);
   for (auto &ep : s->ev_pool) { cudaEventDestroy(ep.a); cudaEventDestroy(ep.b); }
   for (int k = 0; k < 2; ++k) if (s->step_ev[k]) cudaEventDestroy(s->step_ev[k]);
   release_nccl(s);
   if (s->copy_st) cudaStreamDestroy(s->copy_st);
   if (s->st) cudaStreamDestroy(s->st);
   delete s;
   return 0;
}

// ---------------------------------------------------------------------------------------------
// creation
static int validate_params(const rt_params *p)
{
   if (!p) return fail("rt_create: null params");
   if (p->M < 2 || (p->M & 1)) return fail("rt_create: M must be even and >= 2 (got %d)", p->M);
   if (p->N < 1 || p->G < 1) return fail("rt_create: N and G must be >= 1");
   if (p->group_begin < 0 || p->group_count < 1 || p->group_begin + p->group_count > p->G)
      return fail("rt_create: group shard [%d,%d) outside [0,%d)", p->group_begin, p->group_begin + p->group_count, p->G);
   if (p->ts_method < 1 || p->ts_method > 3) return fail("Incorrect timestepping method provided.");
   if (p->bc_left < 0 || p->bc_left > 2 || p->bc_right < 0 || p->bc_right > 2) return fail("Incorrect boundary conditions provided.");
   if (!p->mu || !p->wt || !p->e_edge || !p->kappa) return fail("rt_create: mu, wt, e_edge and kappa are required");
   for (int i = 0; i < p->M; ++i)
      if (p->mu[i] == 0.0 || (i < p->M / 2) != (p->mu[i] < 0.0)) return fail("mu should not be 0 and must be ascending with M/2 negative nodes");
   if (p->max_timesteps < 0) return fail("rt_create: max_timesteps < 0");
   if (p->cv < 0.0) return fail("rt_create: cv must be >= 0 (0 = no material-temperature update)");
   if ((p->rho_cell || p->T_cell || p->kappa_cell || p->cv > 0.0) && p->validate)
      return fail("rt_create: the validation sums (correction.cpp:39-63) are defined for the uniform slab only; set validate = 0 with cell-dependent fields");
   return 0;
}

extern "C" int rt_create(const rt_params *p, rt_solver **out)
{
   if (!out) return fail("rt_create: null out");
   *out = nullptr;
   if (validate_params(p)) return 1;
   int ndev = 0;
   if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
      return fail("rt_create: no CUDA device available (librt_cuda has no CPU fallback)");
   if (p->device < 0 || p->device >= ndev) return fail("rt_create: device %d not in [0,%d)", p->device, ndev);
   CK(cudaSetDevice(p->device));
   cudaDeviceProp prop;
   CK(cudaGetDeviceProperties(&prop, p->device));
   if (prop.major < 10) return fail("rt_create: device %d is sm_%d%d; librt_cuda is built for sm_100a only", p->device, prop.major, prop.minor);

   rt_solver *s = new rt_solver();
   std::memset(&s->tm, 0, sizeof s->tm);
   s->p = *p;
   const int M = p->M, N = p->N, G = p->G;
   s->mu.assign(p->mu, p->mu + M);
   s->wt.assign(p->wt, p->wt + M);
   s->e_edge.assign(p->e_edge, p->e_edge + G + 1);
   s->kappa.assign(p->kappa, p->kappa + G);
   s->psi_source.assign((size_t)M * G, 0.0); // zero when absent
   if (p->psi_source) std::copy(p->psi_source, p->psi_source + (size_t)M * G, s->psi_source.begin());
   s->p.mu = s->mu.data(); s->p.wt = s->wt.data(); s->p.e_edge = s->e_edge.data(); s->p.kappa = s->kappa.data();
   s->p.psi_source = s->psi_source.data();
   s->tupdate = p->cv > 0.0;
   s->var = p->rho_cell || p->T_cell || p->kappa_cell || s->tupdate;
   if (s->var)
   {
      s->rho_cell.assign(N, p->rho);
      s->T_cell.assign(N, p->T);
      s->tab_g0 = std::max(0, p->group_begin - 1);
      s->tab_n = std::min(G, p->group_begin + p->group_count + 1) - s->tab_g0;
      s->kappa_cell.resize((size_t)s->tab_n * N);
      s->bd_kappa.resize((size_t)G * 2);
      if (p->rho_cell) std::copy(p->rho_cell, p->rho_cell + N, s->rho_cell.begin());
      if (p->T_cell) std::copy(p->T_cell, p->T_cell + N, s->T_cell.begin());
      auto kap = [&](int g, int c) { return p->kappa_cell ? p->kappa_cell[(size_t)g + (size_t)G * c] : p->kappa[g]; };
      for (int c = 0; c < N; ++c)
         for (int t = 0; t < s->tab_n; ++t)
            s->kappa_cell[(size_t)t * N + c] = kap(s->tab_g0 + t, c); // group-major on the device
      for (int g = 0; g < G; ++g) { s->bd_kappa[2 * (size_t)g] = kap(g, 0); s->bd_kappa[2 * (size_t)g + 1] = kap(g, N - 1); }
      s->p.rho_cell = s->rho_cell.data(); s->p.T_cell = s->T_cell.data(); s->p.kappa_cell = nullptr; // (the copy is transposed)
   }
   s->num_sms = prop.multiProcessorCount;
   s->Gl = p->group_count;
   s->Np = (N + 7) & ~7;
   // RT_CUDA_SWEEP_THREADS / _CPT / _RING: one value for all stage kinds or "be,cn,bdf"
   auto env3 = [](const char *name, int out[3]) {
      const char *e = getenv(name);
      if (!e) return;
      int v[3];
      const int got = sscanf(e, "%d,%d,%d", &v[0], &v[1], &v[2]);
      if (got == 1) v[1] = v[2] = v[0];
      if (got == 1 || got == 3) for (int k = 0; k < 3; ++k) out[k] = v[k];
   };
   int want_t[3] = {s->threads[0], s->threads[1], s->threads[2]}, want_c[3] = {s->cpt[0], s->cpt[1], s->cpt[2]};
   env3("RT_CUDA_SWEEP_THREADS", want_t);
   env3("RT_CUDA_SWEEP_CPT", want_c);
   env3("RT_CUDA_SWEEP_RING", s->ring);
   for (int k = 0; k < 3; ++k)
   {
      if (sweep_shape_supported(want_t[k], want_c[k])) { s->threads[k] = want_t[k]; s->cpt[k] = want_c[k]; }
      if (s->ring[k] != 2 && s->ring[k] != 3) s->ring[k] = 0;
      s->tile_cells[k] = s->threads[k] * s->cpt[k];
      s->tiles_per_row[k] = (N + s->tile_cells[k] - 1) / s->tile_cells[k];
      s->max_tiles_per_row = std::max(s->max_tiles_per_row, s->tiles_per_row[k]);
   }
   {
      int st3[3] = {s->step_threads, 0, 0}, sc3[3] = {s->step_cpt, 0, 0};
      env3("RT_CUDA_STEP_THREADS", st3);
      env3("RT_CUDA_STEP_CPT", sc3);
      if (step_shape_supported(st3[0], sc3[0])) { s->step_threads = st3[0]; s->step_cpt = sc3[0]; }
      const char *ce = getenv("RT_CUDA_STEP_CHAIN");
      if (ce && step_chain_supported(atoi(ce))) s->step_chain = atoi(ce);
      if (s->step_threads != 32) s->step_chain = 1;
      s->step_box_cells = s->step_threads * s->step_cpt;          // one TMA box: a CTA tile, or one warp's sub-tile
      s->step_tile_cells = s->step_box_cells * s->step_chain;     // cells per ticket
      s->step_tiles_per_row = (N + s->step_tile_cells - 1) / s->step_tile_cells;
      s->max_tiles_per_row = std::max(s->max_tiles_per_row, 4 * s->step_tiles_per_row); // 4 look-back words per tile
      if (s->var) s->max_tiles_per_row = std::max(s->max_tiles_per_row, 4 * ((N + 255) / 256)); // fused cell-dependent step, 256-cell tiles
      if (const char *ke = getenv("RT_CUDA_STEP_KERNEL")) s->step_kernel = (atoi(ke) == 1) ? 1 : (atoi(ke) == 4 ? 4 : 0);
      s->step4_ok = s->step_threads == 32 && s->step_cpt == 8 && step4_chain_supported(s->step_chain);
      if (!s->step4_ok) s->step_kernel = 1;
      for (int r = 1; r <= 8; ++r)
         if ((M / 2) % r == 0) s->phi_run = r;
      const char *fe = getenv("RT_CUDA_FUSED");
      s->fused = !(fe && fe[0] == '0') && !s->var;
      // cell-dependent fields: the fused step of var_step_kernel.cu (RT_CUDA_FUSED=0 or RT_CUDA_VAR_FUSED=0: stage by stage)
      const char *vfe = getenv("RT_CUDA_VAR_FUSED");
      s->fused_var = s->var && !(fe && fe[0] == '0') && !(vfe && vfe[0] == '0');
   }
   const char *tenv = getenv("RT_CUDA_TIMING");
   s->timing_on = !(tenv && tenv[0] == '0');
   if (const char *be = getenv("RT_CUDA_TICKET_BATCH"))
   {
      const int b = atoi(be);
      if (b >= 1 && b <= 64) s->ticket_batch = b;
   }
   const char *penv = getenv("RT_CUDA_L2_PREFETCH");
   s->l2_prefetch = !(penv && penv[0] == '0');
   { const char *se = getenv("RT_CUDA_STEP_L2_PREFETCH"); s->step_l2_prefetch = se && se[0] == '1'; }
   const char *lenv = getenv("RT_CUDA_LOOKBACK");
   s->lookback_walk = lenv && std::strcmp(lenv, "walk") == 0;

#define CKD(x) do { if (x) { rt_destroy(s); return 1; } } while (0)
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fail("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); rt_destroy(s); return 1; } } while (0)
   CKC(cudaStreamCreateWithFlags(&s->st, cudaStreamNonBlocking));
   CKC(cudaStreamCreateWithFlags(&s->copy_st, cudaStreamNonBlocking));
   CKD(dalloc(&s->d_mu, M)); CKD(dalloc(&s->d_wt, M)); CKD(dalloc(&s->d_e_edge, G + 1)); CKD(dalloc(&s->d_kappa, G));
   CKD(dalloc(&s->d_T, 1)); CKD(dalloc(&s->d_B, G)); CKD(dalloc(&s->d_dBdT, G)); CKD(dalloc(&s->d_kappa_edge, G + 1));
   CKD(dalloc(&s->d_epf, G + 1)); CKD(dalloc(&s->d_dEB, G)); CKD(dalloc(&s->d_dsig, G)); CKD(dalloc(&s->d_dkap, G));
   CKD(dalloc(&s->d_cor1, G)); CKD(dalloc(&s->d_cor2, G)); CKD(dalloc(&s->d_cor3, G)); CKD(dalloc(&s->d_sigma, G));
   CKD(dalloc(&s->d_srcB, G)); CKD(dalloc(&s->d_de_ave, G)); CKD(dalloc(&s->d_sums, 3));
   CKD(dalloc(&s->d_psi_source, (size_t)M * G)); CKD(dalloc(&s->d_bdry, (size_t)M * G));
   CKD(dalloc(&s->d_balance, s->Gl)); CKD(dalloc(&s->d_left, s->Gl)); CKD(dalloc(&s->d_right, s->Gl));
   for (int k = 0; k < 3; ++k) CKD(dalloc(&s->d_coef[k], (size_t)s->Gl * M * COEF_ENTRY_DOUBLES));
   CKD(dalloc(&s->d_coef_step, (size_t)s->Gl * M * STEP_COEF_DOUBLES));
   CKD(dalloc(&s->d_coef4, (size_t)s->Gl * M * STEP4_COEF_DOUBLES));
   CKD(dalloc(&s->d_ticket, 1));
   CKC(cudaMemsetAsync(s->d_ticket, 0, sizeof(unsigned long long), s->st));
   CKC(cudaMemsetAsync(s->d_B, 0, G * sizeof(double), s->st));
   CKC(cudaMemsetAsync(s->d_dBdT, 0, G * sizeof(double), s->st));
   CKD(upload(s, s->d_mu, s->mu.data(), M)); CKD(upload(s, s->d_wt, s->wt.data(), M));
   CKD(upload(s, s->d_e_edge, s->e_edge.data(), G + 1)); CKD(upload(s, s->d_kappa, s->kappa.data(), G));
   CKD(upload(s, s->d_T, &s->p.T, 1)); CKD(upload(s, s->d_psi_source, s->psi_source.data(), (size_t)M * G));
   {
      double pt[12], w[12];
      gauss12_host(pt, w);
      upload_gauss_rule(pt, w);
   }
   if (s->var)
   {
      CKD(dalloc(&s->d_rho_c, N)); CKD(dalloc(&s->d_T_c, N)); CKD(dalloc(&s->d_kappa_c, (size_t)s->tab_n * N));
      CKD(dalloc(&s->d_B_c, (size_t)s->Gl * N)); CKD(dalloc(&s->d_dBdT_c, (size_t)s->Gl * N));
      CKD(dalloc(&s->d_bd_T, 2)); CKD(dalloc(&s->d_bd_rho, 2)); CKD(dalloc(&s->d_bd_kappa, (size_t)G * 2));
      CKD(dalloc(&s->d_bd_B, (size_t)G * 2)); CKD(dalloc(&s->d_bd_dB, (size_t)G * 2));
      s->vt_pitch = (N + 15) & ~15;
      for (int k = 0; k < 5; ++k)
      {
         const size_t n = (size_t)2 * s->Gl * s->vt_pitch;
         CKD(dalloc(&s->d_vt[k], n));
         CKC(cudaMemsetAsync(s->d_vt[k], 0, n * sizeof(double), s->st)); // the padding is read by TMA boxes (and masked)
      }
      if (const char *ve = getenv("RT_CUDA_VAR_THREADS"))
      {
         const int t = atoi(ve);
         if (t == 0 || var_tile_shape_supported(t, 4)) s->var_threads = t;
      }
      if (const char *ce = getenv("RT_CUDA_VAR_CPT"))
         if (s->var_threads > 0 && var_tile_shape_supported(s->var_threads, atoi(ce))) s->var_cpt = atoi(ce);
      if (!(s->var_threads > 0 && var_step_shape_supported(s->var_threads, s->var_cpt))) s->fused_var = false;
      if (const char *me = getenv("RT_CUDA_VAR_STEP_MATBUF"))
         if (atoi(me) == 1 || atoi(me) == 2) s->var_step_mat_slots = atoi(me);
      if (const char *we = getenv("RT_CUDA_VAR_WARP")) s->var_warp = atoi(we) != 0;
      if (const char *ce = getenv("RT_CUDA_VAR_CHAIN"))
         if (var_warp_chain_supported(atoi(ce))) s->var_chain = atoi(ce);
      if (!s->fused_var || p->ts_method != 3) s->var_warp = false;
      for (int k = 0; k < 4; ++k) CKD(dalloc(&s->d_edge_state[k], G));
      if (s->tupdate)
      {
         CKD(dalloc(&s->d_sdB, (size_t)s->Gl * s->vt_pitch)); CKD(dalloc(&s->d_T_prev, N)); CKD(dalloc(&s->d_xsums, (size_t)3 * N)); CKD(dalloc(&s->d_ppart, (size_t)2 * N));
         CKD(dalloc(&s->d_xpart, (size_t)exchange_chunks(s->Gl) * 3 * N));
         CKD(dalloc(&s->d_T_bad, 1));
         CKC(cudaMemsetAsync(s->d_T_bad, 0, sizeof(int), s->st));
         for (int m = 0; m < M; ++m) s->wsum += s->wt[m];
      }
      CKC(cudaMemsetAsync(s->d_B_c, 0, (size_t)s->Gl * N * sizeof(double), s->st));
      CKC(cudaMemsetAsync(s->d_dBdT_c, 0, (size_t)s->Gl * N * sizeof(double), s->st));
      CKD(upload(s, s->d_rho_c, s->rho_cell.data(), N)); CKD(upload(s, s->d_T_c, s->T_cell.data(), N));
      CKD(upload(s, s->d_kappa_c, s->kappa_cell.data(), (size_t)s->tab_n * N));
      std::vector<double>().swap(s->kappa_cell); // the device copy is the only one a run needs
      const double bT[2] = {s->T_cell[0], s->T_cell[N - 1]}, brho[2] = {s->rho_cell[0], s->rho_cell[N - 1]};
      CKD(upload(s, s->d_bd_T, bT, 2)); CKD(upload(s, s->d_bd_rho, brho, 2));
      CKD(upload(s, s->d_bd_kappa, s->bd_kappa.data(), (size_t)G * 2));
   }

   // ---- memory plan: time-level buffers of (u,v) per unknown + look-back state + result / per-step output arrays.
   // The fused BDF2 step reads ends^n and writes ends^(n+1): TWO levels (BE / CN time stepping likewise); the other two
   // (half_ends, the stage-2 result) exist only when a step runs stage by stage -- RT_CUDA_FUSED=0, a partial step through
   // rt_run_stages, rt_get_ends(which = 2) -- and are allocated on demand (ensure_levels).  C4 (2^20 x S16 x 256 groups)
   // is 137 GB at two levels: one resident batch on a B200 instead of two.
   size_t free_b = 0, total_b = 0;
   CKC(cudaMemGetInfo(&free_b, &total_b));
   CKD(pool_setup(p->device));
   free_b += pool_idle_bytes(p->device);
   int plan_levels = (p->ts_method != 3 || s->fused || s->fused_var) ? 2 : 4;
   if (const char *le = getenv("RT_CUDA_LEVELS"))
      if (atoi(le) == 4) plan_levels = 4;
   s->plan_levels = plan_levels;
   s->phi_pitch = (long long)s->step_tiles_per_row * s->step_tile_cells;
   const size_t phi_slots = (size_t)(M / 2) / s->phi_run;
   // per resident group: state, look-back words, per-step output staging (2 x phi) and its partial sums
   const size_t per_group = (size_t)plan_levels * M * s->Np * sizeof(double2) + (size_t)M * s->max_tiles_per_row * 20 +
                            (size_t)2 * N * sizeof(double) + 2 * phi_slots * (size_t)s->phi_pitch * sizeof(double);
   const size_t keep = (size_t)3 * s->Gl * N * sizeof(double); // phi / phi_plus / F of every owned group
   const size_t reserve = (size_t)1 << 30; // layout-conversion staging, allocator slack
   int gb = p->groups_per_batch;
   if (gb <= 0)
   {
      const size_t usable = free_b > reserve + keep ? free_b - reserve - keep : free_b / 2;
      const size_t fit = usable / per_group;
      gb = (int)std::min<size_t>(fit, (size_t)s->Gl);
      if (gb < 1) { fail("rt_create: not even one group (%zu bytes) fits in %zu free bytes", per_group, free_b); rt_destroy(s); return 1; }
      // balance the batches: same number of batches, equal sizes
      const int nb = (s->Gl + gb - 1) / gb;
      gb = (s->Gl + nb - 1) / nb;
   }
   gb = std::min(gb, s->Gl);
   if (s->var && gb < s->Gl)
   {
      fail("rt_create: cell-dependent fields need all %d owned groups resident (only %d fit / were asked for)", s->Gl, gb);
      rt_destroy(s);
      return 1;
   }
   s->Gb = gb;
   s->n_batches = (s->Gl + gb - 1) / gb;
   s->buf_elems = (size_t)gb * M * s->Np;
   CKD(ensure_levels(s, plan_levels));
   if (s->var && s->var_threads > 0)
      for (int k = 0; k < 5; ++k) // a table row as pitch/2 16-byte "cells": the same 128-byte-swizzled box as a time-level tile
         CKC(make_tile_map(&s->mat_maps.m[k], s->d_vt[k], 2 * s->Gl, s->vt_pitch / 2, s->var_threads * s->var_cpt / 2));
   if (s->var && s->var_warp)
      for (int k = 0; k < 5; ++k) CKC(make_tile_map(&s->mat_maps_vwarp.m[k], s->d_vt[k], 2 * s->Gl, s->vt_pitch / 2, var_warp_subtile_cells() / 2));
   const size_t ntiles = (size_t)gb * M * s->max_tiles_per_row;
   CKD(dalloc(&s->d_tile_status, ntiles));
   CKC(cudaMemsetAsync(s->d_tile_status, 0, ntiles * sizeof(double2), s->st));
   if (s->n_batches > 1)
   {
      CKD(dalloc(&s->d_phi, (size_t)s->Gl * N)); CKD(dalloc(&s->d_phip, (size_t)s->Gl * N)); CKD(dalloc(&s->d_F, (size_t)s->Gl * N));
   }
   CKD(compute_group_tables(s));
   CKD(load_batch(s, 0));
   CKC(cudaStreamSynchronize(s->st));
#undef CKD
#undef CKC
   *out = s;
   return 0;
}

extern "C" int rt_get_plan(const rt_solver *s, rt_plan *plan)
{
   if (!s || !plan) return fail("rt_get_plan: null argument");
   plan->groups_per_batch = s->Gb;
   plan->n_batches = s->n_batches;
   plan->resident = s->n_batches == 1;
   plan->tile_cells = s->tile_cells[KIND_BE];
   plan->tiles_per_row = s->tiles_per_row[KIND_BE];
   plan->sweep_grid = s->last_grid;
   plan->sweep_block = s->last_block ? s->last_block : s->threads[KIND_BE];
   plan->fused = ((s->fused || s->fused_var) && s->p.ts_method == 3) ? 1 : 0;
   plan->step_tile_cells = s->var ? (s->var_warp ? s->var_chain * var_warp_subtile_cells() : s->var_threads * s->var_cpt) : s->step_tile_cells;
   plan->step_block = s->var ? (s->var_warp ? 32 * s->var_chain : s->var_threads) : s->step_threads * s->step_chain;
   plan->step_chain = s->var ? 1 : s->step_chain;
   plan->bytes_per_group = (size_t)s->plan_levels * s->p.M * s->Np * sizeof(double2);
   plan->state_bytes = (size_t)s->levels * s->buf_elems * sizeof(double2);
   plan->time_levels = s->levels;
   plan->var_scan_steps = s->var ? s->var_scan_steps : 5;
   plan->step_kernel = (s->step_kernel == 4 || (s->step_kernel == 0 && ((s->step_phi_host && s->step_group_major) || s->hist_E_host) && s->d_phi_part)) ? 4 : 1;
   cudaSetDevice(s->p.device);
   cudaMemGetInfo(&plan->free_bytes, &plan->total_bytes);
   return 0;
}

// ---------------------------------------------------------------------------------------------
// stage schedule (solver.cpp:606-820)
static int pick_free(int a, int b, int c, int d)
{
   for (int k = 0; k < 4; ++k)
      if (k != a && k != b && k != c && k != d) return k;
   return -1;
}

static int total_stages(const rt_solver *s) { return s->p.max_timesteps * (s->p.ts_method == 3 ? 4 : 1); }

// launch the per-stage sweep kernel: X1 = E[i1n] / E[i1p] (mu<0 / mu>0 rows), X2 = E[i2], XC = E[ic] -> E[out]
static int launch_stage(rt_solver *s, int kind, int i1n, int i1p, int i2, int ic, int out, bool counted)
{
   const rt_params &p = s->p;
   const bool corr = p.use_correction && p.V != 0.0; // V == 0 makes total_correction exactly zero
   SweepArgs a;
   a.x1[0] = s->E[i1n]; a.x1[1] = s->E[i1p];
   a.x2 = i2 >= 0 ? s->E[i2] : nullptr;
   a.xc = ic >= 0 ? s->E[ic] : nullptr;
   a.ix1[0] = i1n; a.ix1[1] = i1p; a.ix2 = i2 >= 0 ? i2 : 0; a.ixc = ic >= 0 ? ic : 0; a.iout = out;
   a.coef = reinterpret_cast<const double2 *>(s->d_coef[kind] + (size_t)s->batch_start * p.M * COEF_ENTRY_DOUBLES);
   a.tile_status = s->d_tile_status;
   a.ticket = s->d_ticket; a.ticket_base = s->ticket_base; a.ticket_batch = s->ticket_batch;
   a.epoch = ++s->epoch;
   a.M = p.M; a.N = p.N; a.Np = s->Np;
   a.rows = s->batch_groups * p.M;
   a.tiles_per_row = s->tiles_per_row[kind];
   a.reflective_left = (p.bc_left == 2) ? 1 : 0;
   a.n_chains = a.reflective_left ? a.rows / 2 : a.rows;
   a.tiles_per_chain = a.reflective_left ? 2 * a.tiles_per_row : a.tiles_per_row;
   a.lookback_walk = s->lookback_walk ? 1 : 0;
   a.l2_prefetch = s->l2_prefetch ? 1 : 0;

   EventPair *ep = nullptr;
   if (counted && event_begin(s, kind, &ep)) return 1;
   int grid = 0;
   if (s->var)
   {
      // cell-dependent material fields (extension)
      const double tau = (p.ts_method == 3) ? p.dt / 2. : p.dt;
      if (s->var_threads > 0)
      {
         // var_tile_kernel.cu: the uniform sweep's machinery (tickets, TMA tiles, look-back) with per-cell constants
         const int tile = s->var_threads * s->var_cpt;
         VarTileArgs va;
         va.s = a;
         va.s.tiles_per_row = (p.N + tile - 1) / tile;
         va.s.tiles_per_chain = a.reflective_left ? 2 * va.s.tiles_per_row : va.s.tiles_per_row;
         va.s.lookback_walk = 0;
         va.mu = s->d_mu; va.bdry = s->d_bdry; va.G0 = p.group_begin; va.Gl = s->Gl;
         va.dx = p.dx; va.tau = tau; va.dt = p.dt; va.beta = p.V / C_LIGHT;
         CK(launch_var_tile(kind, corr, s->var_threads, s->var_cpt, s->maps_var, s->mat_maps, va, s->num_sms, s->st, &grid));
         const unsigned long long total_tiles = (unsigned long long)va.s.n_chains * va.s.tiles_per_chain;
         s->ticket_base += (total_tiles + s->ticket_batch - 1) / s->ticket_batch + (unsigned long long)grid;
         s->last_block = s->var_threads;
      }
      else
      {
         // var_kernel.cu, the first version kept as a cross-check: one CTA per chain walks its tiles in sequence
         VarSweepArgs v;
         v.x1[0] = s->E[i1n]; v.x1[1] = s->E[i1p]; v.x2 = a.x2; v.xc = a.xc; v.out = s->E[out];
         v.t.sigma = s->vt_fwd(0); v.t.srcB = s->vt_fwd(1); v.t.cor1 = s->vt_fwd(2); v.t.cor2 = s->vt_fwd(3); v.t.cor3 = s->vt_fwd(4);
         v.t.pitch = s->vt_pitch;
         v.mu = s->d_mu; v.bdry = s->d_bdry;
         v.M = p.M; v.N = p.N; v.Np = s->Np; v.G0 = p.group_begin; v.rows = a.rows; v.reflective_left = a.reflective_left;
         v.dx = p.dx; v.tau = tau; v.dt = p.dt; v.beta = p.V / C_LIGHT;
         CK(launch_var_sweep(kind, corr, v, s->st, &grid));
         s->last_block = 128;
      }
      if (counted && event_end(s, ep)) return 1;
      s->last_grid = grid;
      s->tm.kernel_launches++;
      if (counted)
      {
         s->tm.sweep_launches[kind]++;
         s->tm.stages++;
         s->tm.unknown_stages += (long long)a.rows * p.N;
      }
      return 0;
   }
   CK(launch_sweep(kind, corr, s->threads[kind], s->cpt[kind], s->ring[kind], s->maps[kind], a, s->num_sms, s->st, &grid));
   if (counted && event_end(s, ep)) return 1;
   const unsigned long long total_tiles = (unsigned long long)a.n_chains * a.tiles_per_chain;
   s->ticket_base += (total_tiles + s->ticket_batch - 1) / s->ticket_batch + (unsigned long long)grid;
   s->last_grid = grid;
   s->last_block = s->threads[kind];
   s->tm.kernel_launches++;
   if (counted)
   {
      s->tm.sweep_launches[kind]++;
      s->tm.stages++;
      s->tm.unknown_stages += (long long)a.rows * p.N;
   }
   return 0;
}

static int materialize_half(rt_solver *s);

// one stage sweep of the batch currently on the device
static int run_one_stage(rt_solver *s)
{
   const rt_params &p = s->p;
   const int it = s->stages_done;
   if (ensure_levels(s, p.ts_method == 3 ? 4 : 2)) return 1;
   int kind, out;
   int i1n, i1p, i2 = -1, ic = -1; // buffer indices: X1 for mu<0 / mu>0 rows, X2, XC
   if (p.ts_method == 1)
   {
      kind = KIND_BE;
      out = pick_free(s->i_ends, -1, -1, -1);
      i1n = i1p = s->i_ends;
      s->i_prev = s->i_ends; // prev_ends = ends (solver.cpp:624), by aliasing instead of copying
   }
   else if (p.ts_method == 2)
   {
      kind = KIND_CN;
      out = pick_free(s->i_ends, -1, -1, -1);
      i1n = i1p = i2 = s->i_ends;
      s->i_prev = s->i_ends;
   }
   else
   {
      switch (it % 4)
      {
      case 0: // BE predictor; prev_ends = ends
         // after a fused step half_ends exists only on demand; the reference keeps step n's half_ends until stage 1 of step
         // n+1 overwrites it, so it has to be materialised now, while prev_ends^n (its input) is still intact
         if (s->half_stale && materialize_half(s)) return 1;
         kind = KIND_BE;
         out = pick_free(s->i_ends, s->i_half_neg, s->i_half_pos, -1);
         i1n = i1p = s->i_ends;
         s->i_prev = s->i_ends;
         break;
      case 1: // CN corrector; afterwards half_ends = CN result for mu<0 rows, stage-0 result for mu>0 rows
         kind = KIND_CN;
         out = pick_free(s->i_ends, s->i_prev, -1, -1);
         i1n = i1p = s->i_ends;
         i2 = s->i_prev;
         s->i_half_pos = s->i_ends;
         s->i_half_neg = out;
         s->half_stale = false;
         break;
      case 2: // second BE predictor
         kind = KIND_BE;
         out = pick_free(s->i_ends, s->i_prev, s->i_half_neg, s->i_half_pos);
         i1n = i1p = s->i_ends;
         break;
      default: // BDF corrector, written in place over the stage-2 result (only read for the correction source, same cell)
         kind = KIND_BDF;
         out = s->i_ends;
         i1n = s->i_half_neg;
         i1p = s->i_half_pos;
         i2 = s->i_prev;
         ic = s->i_ends;
         break;
      }
   }
   if (out < 0) return fail("internal: no free time-level buffer");
   if (launch_stage(s, kind, i1n, i1p, i2, ic, out, true)) return 1;
   s->i_ends = out;
   s->stages_done++;
   s->step_phi_fused = false;
   s->results_valid = false;
   return 0;
}

// One whole BDF2 step (4 stage sweeps) in a single pass over the state (step_kernel.cu), out of place: afterwards
// prev_ends is the old buffer; half_ends is not stored (rt_get_ends re-materialises it on demand).
static int run_fused_step_var(rt_solver *s);

static int run_fused_step(rt_solver *s)
{
   if (s->var) return run_fused_step_var(s);
   const rt_params &p = s->p;
   if (ensure_levels(s, 2)) return 1;
   const int in = s->i_ends, out = pick_free(in, -1, -1, -1);
   // the per-step scalar intensity rides along in the one-scan kernel: partial sums of w_m psi per run of rows
   const bool want_phi = ((s->step_phi_host && s->step_group_major) || s->hist_E_host) && s->d_phi_part;
   const bool v4 = s->step_kernel == 4 || (s->step_kernel == 0 && want_phi);
   const bool phi = v4 && want_phi;
   StepArgs a;
   a.x = s->E[in];
   a.ix = in; a.iout = out;
   for (int k = 0; k < 3; ++k)
      a.coef[k] = reinterpret_cast<const double2 *>(s->d_coef[k] + (size_t)s->batch_start * p.M * COEF_ENTRY_DOUBLES);
   a.coef_step = s->d_coef_step + (size_t)s->batch_start * p.M * STEP_COEF_DOUBLES;
   a.coef4 = s->d_coef4 + (size_t)s->batch_start * p.M * STEP4_COEF_DOUBLES;
   a.phi_part = phi ? s->d_phi_part : nullptr;
   a.phi_pitch = s->phi_pitch;
   a.phi_slots = (p.M / 2) / s->phi_run;
   a.tile_status = s->d_tile_status;
   // tickets per draw: the warp-tile kernels decode a batch with 8 lanes at most; with PHI a draw is one run of rows
   const int tb = phi ? s->phi_run : std::min(s->ticket_batch, 8);
   a.ticket = s->d_ticket; a.ticket_base = s->ticket_base; a.ticket_batch = tb;
   a.epoch = ++s->epoch;
   a.M = p.M; a.N = p.N; a.Np = s->Np;
   a.rows = s->batch_groups * p.M;
   a.tiles_per_row = s->step_tiles_per_row;
   a.reflective_left = (p.bc_left == 2) ? 1 : 0;
   a.n_chains = a.reflective_left ? a.rows / 2 : a.rows;
   a.tiles_per_chain = a.reflective_left ? 2 * a.tiles_per_row : a.tiles_per_row;
   a.l2_prefetch = s->step_l2_prefetch ? 1 : 0;
   const bool corr = p.use_correction && p.V != 0.0;
   EventPair *ep;
   if (event_begin(s, 4, &ep)) return 1;
   int grid = 0;
   if (v4) CK(launch_step4(corr, phi, s->step_chain, s->maps_step, a, s->num_sms, s->st, &grid));
   else CK(launch_step(corr, s->step_threads, s->step_cpt, s->step_chain, s->maps_step, a, s->num_sms, s->st, &grid));
   if (event_end(s, ep)) return 1;
   const unsigned long long total_tiles = (unsigned long long)a.n_chains * a.tiles_per_chain;
   s->ticket_base += (total_tiles + tb - 1) / tb + (unsigned long long)grid;
   s->last_grid = grid;
   s->last_block = s->step_threads * s->step_chain;
   s->i_prev = in;
   s->i_ends = out;
   s->i_half_neg = s->i_half_pos = -1;
   s->half_stale = true;
   s->step_phi_fused = phi;
   s->stages_done += 4;
   s->tm.step_launches++;
   s->tm.kernel_launches++;
   s->tm.stages += 4;
   s->tm.unknown_stages += 4ll * a.rows * p.N;
   s->results_valid = false;
   return 0;
}

// the same with cell-dependent material fields (var_step_kernel.cu)
static int run_fused_step_var(rt_solver *s)
{
   const rt_params &p = s->p;
   const int in = s->i_ends, out = pick_free(in, -1, -1, -1);
   const int tile = s->var_warp ? s->var_chain * var_warp_subtile_cells() : s->var_threads * s->var_cpt;
   VarStepArgs va;
   StepArgs &a = va.s;
   a.x = s->E[in];
   a.ix = in; a.iout = out;
   for (int k = 0; k < 3; ++k) a.coef[k] = nullptr;
   a.coef_step = nullptr;
   a.tile_status = s->d_tile_status;
   a.ticket = s->d_ticket; a.ticket_base = s->ticket_base; a.ticket_batch = s->ticket_batch;
   a.epoch = ++s->epoch;
   a.M = p.M; a.N = p.N; a.Np = s->Np;
   a.rows = s->batch_groups * p.M;
   a.tiles_per_row = (p.N + tile - 1) / tile;
   a.reflective_left = (p.bc_left == 2) ? 1 : 0;
   a.n_chains = a.reflective_left ? a.rows / 2 : a.rows;
   a.tiles_per_chain = a.reflective_left ? 2 * a.tiles_per_row : a.tiles_per_row;
   a.l2_prefetch = 0;
   va.mu = s->d_mu; va.bdry = s->d_bdry; va.G0 = p.group_begin; va.Gl = s->Gl;
   va.dx = p.dx; va.tau = p.dt / 2.; va.dt = p.dt; va.beta = p.V / C_LIGHT;
   const bool corr = p.use_correction && p.V != 0.0;
   EventPair *ep;
   if (event_begin(s, 4, &ep)) return 1;
   int grid = 0;
   if (s->var_warp) CK(launch_var_warp(corr, s->var_chain, s->maps_vwarp, s->mat_maps_vwarp, va, s->num_sms, s->st, &grid));
   else CK(launch_var_step(corr, s->var_threads, s->var_cpt, s->var_threads == 64 ? s->var_step_mat_slots : 2, s->var_scan_steps, s->maps_var, s->mat_maps, va, s->num_sms, s->st, &grid));
   if (event_end(s, ep)) return 1;
   const unsigned long long total_tiles = (unsigned long long)a.n_chains * a.tiles_per_chain;
   s->ticket_base += (total_tiles + s->ticket_batch - 1) / s->ticket_batch + (unsigned long long)grid;
   s->last_grid = grid;
   s->last_block = s->var_warp ? 32 * s->var_chain : s->var_threads;
   s->i_prev = in;
   s->i_ends = out;
   s->i_half_neg = s->i_half_pos = -1;
   s->half_stale = true;
   s->stages_done += 4;
   s->tm.step_launches++;
   s->tm.kernel_launches++;
   s->tm.stages += 4;
   s->tm.unknown_stages += 4ll * a.rows * p.N;
   s->results_valid = false;
   return 0;
}

// half_ends after a fused step: re-run the step's stage 0 (BE) and stage 1 (CN) from prev_ends with the per-stage kernel
// into the two free buffers -- half = stage-1 result for mu<0 rows, stage-0 result for mu>0 rows
static int materialize_half(rt_solver *s)
{
   if (!s->half_stale) return 0;
   if (ensure_levels(s, 4)) return 1;
   const int A = s->i_prev, C = pick_free(s->i_ends, A, -1, -1), D = pick_free(s->i_ends, A, C, -1);
   if (A < 0 || C < 0 || D < 0) return fail("internal: cannot re-materialise half_ends");
   if (launch_stage(s, KIND_BE, A, A, -1, -1, C, false)) return 1;
   if (launch_stage(s, KIND_CN, C, C, A, -1, D, false)) return 1;
   s->i_half_pos = C;
   s->i_half_neg = D;
   s->half_stale = false;
   return 0;
}

static int step_output(rt_solver *s);
static int step_history(rt_solver *s);
static int deliver_history(rt_solver *s);
static int finish_batch(rt_solver *s);
static int ensure_result_arrays(rt_solver *s);

// Material-temperature update after a completed time step (extension; definition: rt_cuda.h, rt_params::cv;
// oracle: rto_temperature_update).  phi of the new state -> group sums A, E, D over the owned groups -> all-reduce hook
// (the one exchange step of the path: groups couple here and only here) -> T' -> Planck integrals and per-cell tables at T'.
static int temperature_update(rt_solver *s)
{
   const rt_params &p = s->p;
   if (s->Gl < p.G && !s->allreduce)
      return fail("rt: this handle owns groups [%d,%d) of %d and the material-temperature update (cv > 0) sums absorption and emission "
                  "over ALL groups: set the exchange with rt_set_allreduce (or create the handle with device_count > 1) -- refusing to "
                  "update T from a partial sum", p.group_begin, p.group_begin + s->Gl, p.G);
   if (ensure_result_arrays(s)) return 1;
   EventPair *ep;
   if (event_begin(s, 5, &ep)) return 1;
   launch_moments(s->E[s->i_ends], s->d_mu, s->d_wt, p.M, p.N, s->Np, s->batch_groups, s->d_phi, nullptr, nullptr, s->st);
   ExchangeArgs x;
   x.N = p.N; x.groups = s->Gl; x.pitch = s->vt_pitch; x.W = s->wsum;
   x.sigma = s->vt_fwd(0); x.srcB = s->vt_fwd(1); x.sdB = s->d_sdB; x.phi = s->d_phi;
   x.partial = s->d_xpart; x.sums = s->d_xsums;
   s->tm.kernel_launches += 1 + launch_exchange_sums(x, s->st);
   CK(cudaGetLastError());
   if (s->allreduce)
   {
      const int rc = s->allreduce(s->d_xsums, (size_t)3 * p.N, (void *)s->st, s->allreduce_user);
      if (rc) return fail("rt: the all-reduce hook failed with code %d in the temperature update after stage %d", rc, s->stages_done);
   }
   launch_temperature_update(p.N, p.ts_method == 3 && s->tsteps_done > 0, p.dt, p.cv, s->d_rho_c, s->d_xsums, s->d_T_c, s->d_T_prev,
                             (int)s->tsteps_done + 1, s->d_T_bad, s->st);
   s->tm.kernel_launches += 1;
   if (event_end(s, ep)) return 1;
   if (event_begin(s, 6, &ep)) return 1;
   // on a group shard (hook set) the Planck integrals are sharded too; a handle that owns every group integrates them all
   if (compute_var_tables(s, false, s->allreduce != nullptr && s->Gl < p.G)) return 1;
   if (event_end(s, ep)) return 1;
   s->tsteps_done++;
   s->tm.tupdate_count++;
   s->T_moved = true;
   return 0;
}

// after the stream has been synchronised: did a temperature update leave the physical range?
static int check_temperature(rt_solver *s)
{
   if (!s->tupdate || !s->d_T_bad) return 0;
   int bad = 0;
   CK(cudaMemcpy(&bad, s->d_T_bad, sizeof(int), cudaMemcpyDeviceToHost));
   if (bad)
      return fail("rt: the material temperature left the physical range (T <= 0 or not finite in some cell) in temperature update %d: "
                  "the explicitly coupled update (rt_params::cv) is only conditionally stable -- reduce dt or raise cv "
                  "(dt = %g, cv = %g; the CPU oracle diverges on the same input)", bad, s->p.dt, s->p.cv);
   return 0;
}

static int run_stages(rt_solver *s, int n)
{
   const int per_step = s->p.ts_method == 3 ? 4 : 1;
   for (int k = 0; k < n;)
   {
      // whole BDF2 steps go through the fused step kernel; anything else (BE / CN time stepping, a partial step
      // requested through rt_run_stages) stage by stage
      if ((s->fused || s->fused_var) && s->p.ts_method == 3 && s->stages_done % 4 == 0 && n - k >= 4)
      {
         if (run_fused_step(s)) return 1;
         k += 4;
      }
      else
      {
         if (run_one_stage(s)) return 1;
         k += 1;
      }
      if (s->tupdate && s->stages_done % per_step == 0)
         if (temperature_update(s)) return 1;
      if (s->step_phi_host && s->stages_done % per_step == 0)
         if (step_output(s)) return 1;
      if ((s->hist_E_host || s->hist_T_host) && s->stages_done % per_step == 0)
         if (step_history(s)) return 1;
   }
   return 0;
}

extern "C" int rt_run_stages(rt_solver *s, int n)
{
   if (!s) return fail("rt_run_stages: null handle");
   if (s->n_batches != 1) return fail("rt_run_stages: only available when all owned groups are resident (use rt_solve)");
   CK(cudaSetDevice(s->p.device));
   if (run_stages(s, n)) return 1;
   if (flush_events(s)) return 1;
   if (deliver_history(s)) return 1;
   CK(cudaStreamSynchronize(s->st));
   if (s->step_phi_host) CK(cudaStreamSynchronize(s->copy_st));
   return check_temperature(s);
}

extern "C" int rt_solve(rt_solver *s)
{
   if (!s) return fail("rt_solve: null handle");
   CK(cudaSetDevice(s->p.device));
   if (s->p.validate && !sums_ok(s))
      return fail("Invalid Correction Terms: sum B = %.17g, sum dBdT = %.17g, sum kappa*B = %.17g", s->sums[0], s->sums[1], s->sums[2]);
   if (s->n_batches == 1)
   {
      if (run_stages(s, total_stages(s) - s->stages_done)) return 1;
   }
   else
   {
      for (int b = 0; b < s->n_batches; ++b)
      {
         if (load_batch(s, b * s->Gb)) return 1;
         if (run_stages(s, total_stages(s))) return 1;
         if (finish_batch(s)) return 1;
      }
      s->results_valid = true;
   }
   if (flush_events(s)) return 1;
   if (deliver_history(s)) return 1;
   CK(cudaStreamSynchronize(s->st));
   if (s->step_phi_host) CK(cudaStreamSynchronize(s->copy_st));
   return check_temperature(s);
}

extern "C" int rt_set_allreduce(rt_solver *s, rt_allreduce_fn fn, void *user)
{
   if (!s) return fail("rt_set_allreduce: null handle");
   s->allreduce = fn;
   s->allreduce_user = user;
   return 0;
}

extern "C" int rt_get_temperature(rt_solver *s, double *T, double *sums3)
{
   if (!s || !T) return fail("rt_get_temperature: null argument");
   CK(cudaSetDevice(s->p.device));
   if (!s->var)
   {
      for (int c = 0; c < s->p.N; ++c) T[c] = s->p.T;
      if (sums3) std::fill(sums3, sums3 + (size_t)3 * s->p.N, 0.0);
      return 0;
   }
   if (download(s, T, s->d_T_c, s->p.N)) return 1;
   if (sums3)
   {
      if (!s->tupdate || s->tsteps_done == 0) std::fill(sums3, sums3 + (size_t)3 * s->p.N, 0.0);
      else if (download(s, sums3, s->d_xsums, (size_t)3 * s->p.N)) return 1;
   }
   return 0;
}

extern "C" int rt_stages_done(const rt_solver *s, int *n) { if (!s || !n) return fail("null argument"); *n = s->stages_done; return 0; }
extern "C" int rt_total_stages(const rt_solver *s, int *n) { if (!s || !n) return fail("null argument"); *n = total_stages(s); return 0; }

extern "C" int rt_reset(rt_solver *s)
{
   if (!s) return fail("rt_reset: null handle");
   CK(cudaSetDevice(s->p.device));
   if (load_batch(s, 0)) return 1;
   s->results_valid = false;
   CK(cudaStreamSynchronize(s->st));
   return 0;
}

// ---------------------------------------------------------------------------------------------
// results
static int ensure_result_arrays(rt_solver *s)
{
   const size_t n = (size_t)s->Gl * s->p.N;
   if (!s->d_phi) { if (dalloc(&s->d_phi, n)) return 1; }
   if (!s->d_phip) { if (dalloc(&s->d_phip, n)) return 1; }
   if (!s->d_F) { if (dalloc(&s->d_F, n)) return 1; }
   return 0;
}

// moments and diagnostics of the batch on the device into the retained result arrays
static int finish_batch(rt_solver *s)
{
   const rt_params &p = s->p;
   if (ensure_result_arrays(s)) return 1;
   const size_t off = (size_t)s->batch_start * p.N;
   EventPair *ep;
   if (event_begin(s, 3, &ep)) return 1;
   launch_moments(s->E[s->i_ends], s->d_mu, s->d_wt, p.M, p.N, s->Np, s->batch_groups, s->d_phi + off, s->d_phip + off,
                  s->d_F + off, s->st);
   if (s->var)
      launch_diagnostics_var(s->E[s->i_ends], s->d_phi + off, s->d_mu, s->d_wt, s->vt_fwd(0), s->vt_pitch, s->d_T_c, s->d_de_ave, p.M, p.N, s->Np,
                             p.group_begin + s->batch_start, s->batch_groups, p.dx, s->d_balance + s->batch_start,
                             s->d_left + s->batch_start, s->d_right + s->batch_start, s->st);
   else
      launch_diagnostics(s->E[s->i_ends], s->d_phi + off, s->d_mu, s->d_wt, s->d_sigma, s->d_de_ave, p.M, p.N, s->Np,
                         p.group_begin + s->batch_start, s->batch_groups, p.dx, p.T, s->d_balance + s->batch_start,
                         s->d_left + s->batch_start, s->d_right + s->batch_start, s->st);
   if (event_end(s, ep)) return 1;
   s->tm.kernel_launches += 2;
   CK(cudaGetLastError());
   return 0;
}

static int ensure_results(rt_solver *s)
{
   if (s->results_valid) return 0;
   if (s->n_batches != 1) return fail("results are only available after rt_solve in batched mode");
   if (finish_batch(s)) return 1;
   s->results_valid = true;
   return 0;
}

// group-major device array [Gl][N] -> host array in reference order g + Gl*c, through a bounded staging buffer
static int download_reference_order(rt_solver *s, const double *d_src, double *host)
{
   const int Gl = s->Gl, N = s->p.N;
   const size_t budget = (size_t)32 << 20; // doubles (256 MB)
   int chunk = (int)std::min<size_t>((size_t)N, std::max<size_t>(32, budget / (size_t)Gl));
   const size_t need = (size_t)chunk * Gl;
   if (s->stage_elems < need)
   {
      if (s->d_stage) CK(cudaFree(s->d_stage));
      s->d_stage = nullptr;
      if (dalloc(&s->d_stage, need)) return 1;
      s->stage_elems = need;
   }
   for (int c0 = 0; c0 < N; c0 += chunk)
   {
      const int nc = std::min(chunk, N - c0);
      launch_transpose_gn(d_src, s->d_stage, Gl, N, c0, nc, s->st);
      s->tm.kernel_launches++;
      CK(cudaGetLastError());
      if (download(s, host + (size_t)Gl * c0, s->d_stage, (size_t)Gl * nc)) return 1;
   }
   return 0;
}

extern "C" int rt_moments(rt_solver *s, double *phi, double *phi_plus, double *F)
{
   if (!s) return fail("rt_moments: null handle");
   CK(cudaSetDevice(s->p.device));
   if (ensure_results(s)) return 1;
   if (phi && download_reference_order(s, s->d_phi, phi)) return 1;
   if (phi_plus && download_reference_order(s, s->d_phip, phi_plus)) return 1;
   if (F && download_reference_order(s, s->d_F, F)) return 1;
   return flush_events(s);
}

extern "C" int rt_moments_group_major(rt_solver *s, double *phi, double *phi_plus, double *F)
{
   if (!s) return fail("rt_moments_group_major: null handle");
   CK(cudaSetDevice(s->p.device));
   if (ensure_results(s)) return 1;
   const size_t n = (size_t)s->Gl * s->p.N;
   if (phi && download(s, phi, s->d_phi, n)) return 1;
   if (phi_plus && download(s, phi_plus, s->d_phip, n)) return 1;
   if (F && download(s, F, s->d_F, n)) return 1;
   return flush_events(s);
}

extern "C" int rt_balance(rt_solver *s, double *balance)
{
   if (!s || !balance) return fail("rt_balance: null argument");
   CK(cudaSetDevice(s->p.device));
   if (ensure_results(s)) return 1;
   return download(s, balance, s->d_balance, s->Gl);
}

extern "C" int rt_group_ends(rt_solver *s, double *left, double *right)
{
   if (!s) return fail("rt_group_ends: null handle");
   CK(cudaSetDevice(s->p.device));
   if (ensure_results(s)) return 1;
   if (left && download(s, left, s->d_left, s->Gl)) return 1;
   if (right && download(s, right, s->d_right, s->Gl)) return 1;
   return 0;
}

static int need_resident(const rt_solver *s, const char *what)
{
   if (s->n_batches != 1) return fail("%s: only available when all owned groups are resident on the device", what);
   return 0;
}

extern "C" int rt_get_psi(rt_solver *s, double *psi)
{
   if (!s || !psi) return fail("rt_get_psi: null argument");
   if (need_resident(s, "rt_get_psi")) return 1;
   CK(cudaSetDevice(s->p.device));
   const size_t n = (size_t)s->p.M * s->Gl * s->p.N;
   double *d_tmp = nullptr;
   if (dalloc(&d_tmp, n)) return 1;
   launch_export_psi(s->E[s->i_ends], d_tmp, s->p.M, s->p.N, s->Np, s->Gl, s->st);
   s->tm.kernel_launches++;
   int rc = download(s, psi, d_tmp, n);
   cudaFree(d_tmp);
   return rc;
}

extern "C" int rt_get_ends(rt_solver *s, int which, double *ends)
{
   if (!s || !ends) return fail("rt_get_ends: null argument");
   if (need_resident(s, "rt_get_ends")) return 1;
   CK(cudaSetDevice(s->p.device));
   const size_t n = (size_t)2 * s->p.M * s->Gl * s->p.N;
   int ineg, ipos;
   if (which == 0) ineg = ipos = s->i_ends;
   else if (which == 1) ineg = ipos = s->i_prev;
   else if (which == 2)
   {
      if (materialize_half(s)) return 1;
      ineg = s->i_half_neg; ipos = s->i_half_pos;
   }
   else return fail("rt_get_ends: which must be 0, 1 or 2");
   if (ineg < 0 || ipos < 0) { std::fill(ends, ends + n, 0.0); return 0; } // never assigned yet: zero (like a fresh resize in the oracle)
   double *d_tmp = nullptr;
   if (dalloc(&d_tmp, n)) return 1;
   launch_export_ends(s->E[ineg], s->E[ipos], d_tmp, s->p.M, s->p.N, s->Np, s->Gl, s->st);
   s->tm.kernel_launches++;
   int rc = download(s, ends, d_tmp, n);
   cudaFree(d_tmp);
   return rc;
}

extern "C" int rt_get_group_terms(rt_solver *s, double *B, double *dBdT, double *dEB, double *dsigEdE, double *dkapEB,
                                  double *kappa_edge)
{
   if (!s) return fail("rt_get_group_terms: null handle");
   CK(cudaSetDevice(s->p.device));
   const int G = s->p.G;
   if (B && download(s, B, s->d_B, G)) return 1;
   if (dBdT && download(s, dBdT, s->d_dBdT, G)) return 1;
   if (dEB && download(s, dEB, s->d_dEB, G)) return 1;
   if (dsigEdE && download(s, dsigEdE, s->d_dsig, G)) return 1;
   if (dkapEB && download(s, dkapEB, s->d_dkap, G)) return 1;
   if (kappa_edge && download(s, kappa_edge, s->d_kappa_edge, G + 1)) return 1;
   return 0;
}

extern "C" int rt_get_psi_source(rt_solver *s, double *psi_source)
{
   if (!s || !psi_source) return fail("rt_get_psi_source: null argument");
   CK(cudaSetDevice(s->p.device));
   return download(s, psi_source, s->d_psi_source, (size_t)s->p.M * s->p.G);
}

extern "C" int rt_validate(rt_solver *s, double *sums3, int *ok)
{
   if (!s) return fail("rt_validate: null handle");
   if (sums3) { sums3[0] = s->sums[0]; sums3[1] = s->sums[1]; sums3[2] = s->sums[2]; }
   if (ok) *ok = sums_ok(s) ? 1 : 0;
   return 0;
}

// the one-scan step kernel adds up w_m psi itself (RT_CUDA_STEP_PHI=0: separate moments pass over the state)
static int ensure_phi_part(rt_solver *s)
{
   const char *pe = getenv("RT_CUDA_STEP_PHI");
   if (s->fused && s->step_kernel != 1 && s->step4_ok && s->p.ts_method == 3 && !(pe && pe[0] == '0') && !s->d_phi_part)
   {
      const size_t np = (size_t)s->Gb * 2 * ((size_t)(s->p.M / 2) / s->phi_run) * (size_t)s->phi_pitch;
      if (dalloc(&s->d_phi_part, np)) return 1;
   }
   return 0;
}

// ---------------------------------------------------------------------------------------------
// per-step output: phi of the step just completed -> host, overlapped with the next step's sweeps
extern "C" int rt_set_step_output(rt_solver *s, double *phi_host, int group_major)
{
   if (!s) return fail("rt_set_step_output: null handle");
   CK(cudaSetDevice(s->p.device));
   if (phi_host && !group_major && s->n_batches != 1)
      return fail("rt_set_step_output: reference-order output needs all owned groups resident; use group_major=1");
   s->step_phi_host = phi_host;
   s->step_group_major = group_major;
   if (phi_host)
   {
      const size_t n = (size_t)s->Gb * s->p.N;
      for (int k = 0; k < 2; ++k)
      {
         if (!s->d_step_phi[k]) { if (dalloc(&s->d_step_phi[k], n)) return 1; }
         if (!s->step_ev[k]) CK(cudaEventCreateWithFlags(&s->step_ev[k], cudaEventDisableTiming));
      }
      if (group_major && ensure_phi_part(s)) return 1;
   }
   return 0;
}

// ---------------------------------------------------------------------------------------------
// per-step history: the two per-step quantities the north_star names besides the scalar intensity.
//   energy[k][c]      = (1/c) sum_g phi_g(c) after time step k: radiation energy density in the reference's normalisation
//                       of phi (sum_m w_m psi; equilibrium: a T^4), summed over the groups THIS handle owns
//   temperature[k][c] = T(c) after the update of time step k (cv > 0 only)
// Both stay on the device while the steps run (group batches add their groups' share to the same row) and are copied to
// the caller's arrays when rt_run_stages / rt_solve returns: no per-step synchronisation, steps * N * 8 bytes in all.
extern "C" int rt_set_step_history(rt_solver *s, double *energy, double *temperature, int capacity_steps)
{
   if (!s) return fail("rt_set_step_history: null handle");
   CK(cudaSetDevice(s->p.device));
   if ((energy || temperature) && capacity_steps < 1) return fail("rt_set_step_history: capacity_steps = %d", capacity_steps);
   if (temperature && !s->tupdate) return fail("rt_set_step_history: a temperature history needs the temperature update (rt_params::cv > 0)");
   const size_t n = (size_t)std::max(capacity_steps, 0) * s->p.N;
   if (n * sizeof(double) > ((size_t)4 << 30))
      return fail("rt_set_step_history: %d steps x %d cells is more than 4 GiB per array; run the steps in several calls", capacity_steps, s->p.N);
   if (s->d_hist_E) { CK(cudaFree(s->d_hist_E)); s->d_hist_E = nullptr; }
   if (s->d_hist_T) { CK(cudaFree(s->d_hist_T)); s->d_hist_T = nullptr; }
   s->hist_E_host = energy;
   s->hist_T_host = temperature;
   s->hist_cap = (energy || temperature) ? capacity_steps : 0;
   s->hist_lo = s->hist_hi = 0;
   if (energy)
   {
      if (dalloc(&s->d_hist_E, n)) return 1;
      if (!s->d_step_phi[0]) { if (dalloc(&s->d_step_phi[0], (size_t)s->Gb * s->p.N)) return 1; }
      if (ensure_phi_part(s)) return 1;
   }
   if (temperature) { if (dalloc(&s->d_hist_T, n)) return 1; }
   return 0;
}

// after a completed time step: this batch's share of the step's radiation energy, and T(c)
static int step_history(rt_solver *s)
{
   const rt_params &p = s->p;
   const int per_step = p.ts_method == 3 ? 4 : 1;
   const int k = s->stages_done / per_step - 1; // time step just completed, counted from the initial state
   if (k < 0 || k >= s->hist_cap) { s->step_phi_fused = false; return 0; } // beyond the caller's capacity: not recorded
   if (s->hist_E_host)
   {
      // phi of the step: left in a staging buffer by step_output a moment ago, or formed here
      const double *phi = nullptr;
      if (s->step_phi_host && s->step_group_major) phi = s->d_step_phi[s->step_parity ^ 1];
      else
      {
         EventPair *ep;
         if (event_begin(s, 3, &ep)) return 1;
         if (s->step_phi_fused)
            launch_phi_combine(s->d_phi_part, s->phi_pitch, (p.M / 2) / s->phi_run, p.N, s->batch_groups, s->d_step_phi[0], s->st);
         else
            launch_moments(s->E[s->i_ends], s->d_mu, s->d_wt, p.M, p.N, s->Np, s->batch_groups, s->d_step_phi[0], nullptr, nullptr, s->st);
         s->step_phi_fused = false;
         if (event_end(s, ep)) return 1;
         s->tm.kernel_launches++;
         phi = s->d_step_phi[0];
      }
      EventPair *ep;
      if (event_begin(s, 3, &ep)) return 1;
      // the first batch of groups starts the row, the others add to it (ascending batches: deterministic)
      launch_group_sum(phi, s->batch_groups, p.N, 1.0 / C_LIGHT, s->batch_start > 0, s->d_hist_E + (size_t)k * p.N, s->st);
      if (event_end(s, ep)) return 1;
      s->tm.kernel_launches++;
   }
   if (s->hist_T_host)
      CK(cudaMemcpyAsync(s->d_hist_T + (size_t)k * p.N, s->d_T_c, (size_t)p.N * sizeof(double), cudaMemcpyDeviceToDevice, s->st));
   if (s->hist_hi <= s->hist_lo) { s->hist_lo = k; s->hist_hi = k + 1; }
   else { s->hist_lo = std::min(s->hist_lo, k); s->hist_hi = std::max(s->hist_hi, k + 1); }
   return 0;
}

static int deliver_history(rt_solver *s)
{
   if (s->hist_hi <= s->hist_lo) return 0;
   const size_t off = (size_t)s->hist_lo * s->p.N, n = (size_t)(s->hist_hi - s->hist_lo) * s->p.N;
   if (s->hist_E_host && download(s, s->hist_E_host + off, s->d_hist_E + off, n)) return 1;
   if (s->hist_T_host && download(s, s->hist_T_host + off, s->d_hist_T + off, n)) return 1;
   s->hist_lo = s->hist_hi = 0;
   return 0;
}

static int step_output(rt_solver *s)
{
   const rt_params &p = s->p;
   const size_t n = (size_t)s->batch_groups * p.N;
   if (s->step_group_major)
   {
      const int k = s->step_parity;
      s->step_parity ^= 1;
      // the copy that last used this staging buffer must have finished before it is overwritten
      CK(cudaStreamWaitEvent(s->st, s->step_ev[k], 0));
      EventPair *ep;
      if (event_begin(s, 3, &ep)) return 1;
      if (s->step_phi_fused)
         launch_phi_combine(s->d_phi_part, s->phi_pitch, (p.M / 2) / s->phi_run, p.N, s->batch_groups, s->d_step_phi[k], s->st);
      else
         launch_moments(s->E[s->i_ends], s->d_mu, s->d_wt, p.M, p.N, s->Np, s->batch_groups, s->d_step_phi[k], nullptr, nullptr, s->st);
      s->step_phi_fused = false;
      if (event_end(s, ep)) return 1;
      s->tm.kernel_launches++;
      cudaEvent_t done;
      CK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
      CK(cudaEventRecord(done, s->st));
      CK(cudaStreamWaitEvent(s->copy_st, done, 0));
      CK(cudaEventDestroy(done));
      CK(cudaMemcpyAsync(s->step_phi_host + (size_t)s->batch_start * p.N, s->d_step_phi[k], n * sizeof(double),
                         cudaMemcpyDeviceToHost, s->copy_st));
      CK(cudaEventRecord(s->step_ev[k], s->copy_st));
   }
   else
   {
      launch_moments(s->E[s->i_ends], s->d_mu, s->d_wt, p.M, p.N, s->Np, s->batch_groups, s->d_step_phi[0], nullptr, nullptr, s->st);
      s->tm.kernel_launches++;
      if (download_reference_order(s, s->d_step_phi[0], s->step_phi_host)) return 1;
      return 0;
   }
   s->tm.d2h_bytes += (long long)(n * sizeof(double));
   return 0;
}

extern "C" int rt_get_timing(const rt_solver *s, rt_timing *t)
{
   if (!s || !t) return fail("rt_get_timing: null argument");
   *t = s->tm;
   return 0;
}

extern "C" int rt_reset_timing(rt_solver *s)
{
   if (!s) return fail("rt_reset_timing: null handle");
   if (flush_events(s)) return 1;
   std::memset(&s->tm, 0, sizeof s->tm);
   return 0;
}

// ---------------------------------------------------------------------------------------------
// standalone Planck evaluation for ncell temperatures (the per-cell extension of SURVEY 8(f).2; with
// ncell = 1 it is exactly what rt_create runs)
extern "C" int rt_planck_groups(int device, int G, const double *e_edge, int ncell, const double *T, double *B, double *dBdT,
                                double *kernel_ms)
{
   if (!e_edge || !T || !B || !dBdT || G < 1 || ncell < 1) return fail("rt_planck_groups: bad argument");
   int ndev = 0;
   if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("rt_planck_groups: no CUDA device available");
   CK(cudaSetDevice(device));
   double *d_e = nullptr, *d_T = nullptr, *d_B = nullptr, *d_dB = nullptr, *d_tmp = nullptr;
   const size_t n = (size_t)ncell * G;
   int rc = 1;
   cudaEvent_t a = nullptr, b = nullptr;
   do
   {
      if (dalloc(&d_e, G + 1) || dalloc(&d_T, ncell) || dalloc(&d_B, n) || dalloc(&d_dB, n) || dalloc(&d_tmp, n)) break;
      if (cudaMemcpy(d_e, e_edge, (G + 1) * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) { fail("rt_planck_groups: H2D failed"); break; }
      if (cudaMemcpy(d_T, T, ncell * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) { fail("rt_planck_groups: H2D failed"); break; }
      cudaMemset(d_B, 0, n * sizeof(double));
      cudaMemset(d_dB, 0, n * sizeof(double));
      double pt[12], w[12];
      gauss12_host(pt, w);
      upload_gauss_rule(pt, w);
      cudaEventCreate(&a);
      cudaEventCreate(&b);
      cudaEventRecord(a, 0);
      launch_planck(G, ncell, d_e, d_T, d_B, d_dB, 0);
      cudaEventRecord(b, 0);
      if (cudaDeviceSynchronize() != cudaSuccess) { fail("rt_planck_groups: kernel failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
      float ms = 0.f;
      cudaEventElapsedTime(&ms, a, b);
      if (kernel_ms) *kernel_ms = ms;
      // the device tables are group-major [G][ncell]; the caller gets [ncell][G] (layout conversion at the ABI edge only)
      bool ok = true;
      for (int k = 0; k < 2 && ok; ++k)
      {
         launch_transpose_gn(k ? d_dB : d_B, d_tmp, G, ncell, 0, ncell, 0);
         ok = cudaMemcpy(k ? dBdT : B, d_tmp, n * sizeof(double), cudaMemcpyDeviceToHost) == cudaSuccess;
      }
      if (!ok) { fail("rt_planck_groups: D2H failed"); break; }
      rc = 0;
   } while (0);
   if (a) cudaEventDestroy(a);
   if (b) cudaEventDestroy(b);
   cudaFree(d_e); cudaFree(d_T); cudaFree(d_B); cudaFree(d_dB); cudaFree(d_tmp);
   return rc;
}

// ---------------------------------------------------------------------------------------------
// Correction::generate_multigroup_opacities (correction.cpp:66-97) on the device: Planck integrals at T, then the
// Planck-weighted opacities normalised to kappa_grey * a c T^4.  Stand-alone like rt_planck_groups; its result is what a
// caller would pass as rt_params::kappa.
extern "C" int rt_multigroup_opacities(int device, int G, const double *e_edge, double T, double kappa_grey, double *kappa,
                                       double *emis_spec, double *ukappa, double *ckappa)
{
   if (!e_edge || G < 1 || !(T > 0.0)) return fail("rt_multigroup_opacities: bad argument");
   int ndev = 0;
   if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("rt_multigroup_opacities: no CUDA device available");
   CK(cudaSetDevice(device));
   double *d = nullptr; // e_edge (G+1) | T (1) | B | dBdT | kappa | emis | ukappa | ckappa
   const size_t n = (size_t)(G + 1) + 1 + 6 * (size_t)G;
   if (dalloc(&d, n)) return 1;
   double *d_e = d, *d_T = d + G + 1, *d_B = d_T + 1, *d_dB = d_B + G, *d_k = d_dB + G, *d_em = d_k + G, *d_uk = d_em + G,
          *d_ck = d_uk + G;
   int rc = 1;
   do
   {
      if (cudaMemset(d, 0, n * sizeof(double)) != cudaSuccess) { fail("rt_multigroup_opacities: memset failed"); break; }
      if (cudaMemcpy(d_e, e_edge, (G + 1) * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess ||
          cudaMemcpy(d_T, &T, sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) { fail("rt_multigroup_opacities: H2D failed"); break; }
      double pt[12], w[12];
      gauss12_host(pt, w);
      upload_gauss_rule(pt, w);
      launch_planck(G, 1, d_e, d_T, d_B, d_dB, 0);
      launch_multigroup_opacities(G, T, kappa_grey, d_e, d_B, d_k, d_em, d_uk, d_ck, 0);
      if (cudaDeviceSynchronize() != cudaSuccess) { fail("rt_multigroup_opacities: kernel failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
      struct { double *h; const double *dv; } outs[] = {{kappa, d_k}, {emis_spec, d_em}, {ukappa, d_uk}, {ckappa, d_ck}};
      bool ok = true;
      for (auto &o : outs)
         if (o.h && cudaMemcpy(o.h, o.dv, G * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) ok = false;
      if (!ok) { fail("rt_multigroup_opacities: D2H failed"); break; }
      rc = 0;
   } while (0);
   cudaFree(d);
   return rc;
}

// ---------------------------------------------------------------------------------------------
// measurement helpers
// For every batch of owned groups: re-initialise, run `warmup_steps` untimed time steps, then time exactly
// `timed_steps` time steps with CUDA events on the launching stream (all launches and the gaps between them).
// *device_ms = sum over batches -- the device time of `timed_steps` steps of ALL owned groups.
// rt_timing is reset at the start and afterwards covers the timed regions only.
extern "C" int rt_bench_steps(rt_solver *s, int warmup_steps, int timed_steps, double *device_ms)
{
   if (!s || !device_ms) return fail("rt_bench_steps: null argument");
   CK(cudaSetDevice(s->p.device));
   const int per_step = s->p.ts_method == 3 ? 4 : 1;
   if (rt_reset_timing(s)) return 1;
   cudaEvent_t e0, e1;
   CK(cudaEventCreate(&e0));
   CK(cudaEventCreate(&e1));
   double total = 0.0;
   const bool timing_was_on = s->timing_on;
   for (int b = 0; b < s->n_batches; ++b)
   {
      if (load_batch(s, b * s->Gb)) return 1;
      s->timing_on = false;
      const rt_timing keep = s->tm;
      if (run_stages(s, warmup_steps * per_step)) return 1;
      CK(cudaStreamSynchronize(s->st));
      CK(cudaStreamSynchronize(s->copy_st));
      s->tm = keep; // warm-up work is not counted
      s->timing_on = timing_was_on;
      CK(cudaEventRecord(e0, s->st));
      if (run_stages(s, timed_steps * per_step)) return 1;
      CK(cudaEventRecord(e1, s->st));
      CK(cudaStreamSynchronize(s->st));
      CK(cudaStreamSynchronize(s->copy_st));
      float ms = 0.f;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      total += ms;
      if (flush_events(s)) return 1;
   }
   CK(cudaEventDestroy(e0));
   CK(cudaEventDestroy(e1));
   s->results_valid = false;
   *device_ms = total;
   return check_temperature(s);
}

// pinned host memory for the caller-owned result buffers of the end-to-end path
extern "C" int rt_host_alloc(void **p, size_t bytes)
{
   if (!p) return fail("rt_host_alloc: null argument");
   CK(cudaMallocHost(p, bytes));
   return 0;
}
extern "C" int rt_host_free(void *p)
{
   if (p) CK(cudaFreeHost(p));
   return 0;
}

// ---------------------------------------------------------------------------------------------
// rt_planck_groups_noalloc -- the Planck group integrals of rt_planck_groups WITHOUT any allocation.
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_planck_groups allocates five device
// buffers per call (copies of e_edge and T, the kernel's group-major tables B and dB/dT, the staging of the layout
// conversion); here they live in ONE file-scope device array of a fixed, modest size that every GPU which loads the
// library holds until the process exits.
//
// The array's NAME is a host-side stand-in in host code, not a device address: what the kernels get is the pointer that
// cudaGetSymbolAddress returns, looked up once per device with that device current (noalloc_scratch below).
// Every call shares the array and finds the previous call's contents in it: the two tables are cleared at the start of
// each call (the last group keeps "the previous value" when the grey remainder is not positive, Planck.cpp:73-76, so
// they must start at zero).  Calls that overlap on one GPU -- two host threads, or this entry point next to itself on
// another stream -- RACE on the array: serialise them.  The call itself is synchronous (it ends with the result copies).
constexpr size_t NOALLOC_ALIGN = 32; // doubles: every sub-array starts on a 256-byte boundary
static __device__ __align__(256) double g_noalloc_scratch[RT_NOALLOC_MAX_DOUBLES];

static size_t noalloc_round(size_t n) { return (n + NOALLOC_ALIGN - 1) / NOALLOC_ALIGN * NOALLOC_ALIGN; }

// device address of the scratch array on the CURRENT device (cached per device; the lookup itself allocates nothing)
static int noalloc_scratch(int device, double **out)
{
   static std::atomic<double *> cache[RT_MAX_DEVICES];
   if (device < 0 || device >= RT_MAX_DEVICES) return fail("rt_planck_groups_noalloc: device %d outside [0,%d)", device, RT_MAX_DEVICES);
   double *p = cache[device].load(std::memory_order_acquire);
   if (!p)
   {
      const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p)
         return fail("rt_planck_groups_noalloc: cudaGetSymbolAddress(scratch) on device %d -> %s", device, cudaGetErrorString(e));
      cache[device].store(p, std::memory_order_release);
   }
   *out = p;
   return 0;
}

extern "C" int rt_planck_groups_noalloc(int device, int G, const double *e_edge, int ncell, const double *T, double *B, double *dBdT)
{
   if (!e_edge || !T || !B || !dBdT || G < 1 || ncell < 1) return fail("rt_planck_groups_noalloc: bad argument");
   // layout of the scratch: e_edge (G+1) | T (ncell) | B [G][ncell] | dB/dT [G][ncell] | staging [ncell][G]
   const size_t n = (size_t)G * (size_t)ncell;
   const size_t o_e = 0, o_T = o_e + noalloc_round((size_t)G + 1), o_B = o_T + noalloc_round((size_t)ncell), o_dB = o_B + noalloc_round(n),
                o_tmp = o_dB + noalloc_round(n), need = o_tmp + noalloc_round(n);
   if (need > (size_t)RT_NOALLOC_MAX_DOUBLES)
      return fail("rt_planck_groups_noalloc: G = %d groups x ncell = %d temperatures need %zu doubles of scratch, the static array holds "
                  "RT_NOALLOC_MAX_DOUBLES = %d (3*G*ncell + G + 1 + ncell, each part rounded up to 32): use rt_planck_groups, or fewer "
                  "temperatures per call", G, ncell, need, RT_NOALLOC_MAX_DOUBLES);
   int ndev = 0;
   if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("rt_planck_groups_noalloc: no CUDA device available");
   if (device < 0 || device >= ndev) return fail("rt_planck_groups_noalloc: device %d not in [0,%d)", device, ndev);
   CK(cudaSetDevice(device));
   double *s = nullptr;
   if (noalloc_scratch(device, &s)) return 1;
   double *d_e = s + o_e, *d_T = s + o_T, *d_B = s + o_B, *d_dB = s + o_dB, *d_tmp = s + o_tmp;
   CK(cudaMemcpyAsync(d_e, e_edge, ((size_t)G + 1) * sizeof(double), cudaMemcpyHostToDevice, 0));
   CK(cudaMemcpyAsync(d_T, T, (size_t)ncell * sizeof(double), cudaMemcpyHostToDevice, 0));
   // the array keeps the last call's contents: what must start at zero is cleared every time
   CK(cudaMemsetAsync(d_B, 0, n * sizeof(double), 0));
   CK(cudaMemsetAsync(d_dB, 0, n * sizeof(double), 0));
   double pt[12], w[12];
   gauss12_host(pt, w);
   upload_gauss_rule(pt, w);
   launch_planck(G, ncell, d_e, d_T, d_B, d_dB, 0);
   CK(cudaGetLastError());
   // the device tables are group-major [G][ncell]; the caller gets [ncell][G] like rt_planck_groups (conversion at the ABI edge)
   launch_transpose_gn(d_B, d_tmp, G, ncell, 0, ncell, 0);
   CK(cudaGetLastError());
   CK(cudaMemcpy(B, d_tmp, n * sizeof(double), cudaMemcpyDeviceToHost));
   launch_transpose_gn(d_dB, d_tmp, G, ncell, 0, ncell, 0);
   CK(cudaGetLastError());
   CK(cudaMemcpy(dBdT, d_tmp, n * sizeof(double), cudaMemcpyDeviceToHost));
   return 0;
}

// ---------------------------------------------------------------------------------------------
// Native NCCL for the material-temperature update's exchange (the ONE collective of the path).
// libnccl.so.2 is loaded with dlopen the first time it is asked for: a single-GPU host needs nothing new, and a process
// that already holds NCCL (torch) gets that very copy.  The types NCCL's ABI needs are declared here (nccl.h:37-38,
// 260-286, 392): no header or link-time dependency.  The all-reduce itself goes through the handle's existing hook
// (rt_set_allreduce), so temperature_update() has one code path whoever provides the sum.
#include <dlfcn.h>

namespace
{
struct NcclId { char internal[128]; };
typedef void *NcclComm;
struct NcclApi
{
   int (*GetUniqueId)(NcclId *) = nullptr;
   int (*CommInitRank)(NcclComm *, int, NcclId, int) = nullptr;
   int (*CommInitAll)(NcclComm *, int, const int *) = nullptr;
   int (*CommDestroy)(NcclComm) = nullptr;
   int (*AllReduce)(const void *, void *, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
   const char *(*GetErrorString)(int) = nullptr;
   void *dl = nullptr;
};
constexpr int NCCL_DOUBLE = 8, NCCL_SUM = 0;

// nullptr (and the error message set) when NCCL cannot be loaded
NcclApi *nccl_api()
{
   static NcclApi api;
   static std::atomic<int> state{0}; // 0 untried, 1 ready, -1 failed
   static std::string why;
   if (state.load(std::memory_order_acquire) == 0)
   {
      const char *name = getenv("RT_CUDA_NCCL_LIB");
      void *dl = dlopen(name ? name : "libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (!dl && !name) dl = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
      bool ok = dl != nullptr;
      if (!ok) why = std::string("dlopen(libnccl.so.2): ") + (dlerror() ? dlerror() : "not found");
      auto sym = [&](const char *n) -> void * {
         void *p = ok ? dlsym(dl, n) : nullptr;
         if (ok && !p) { ok = false; why = std::string("libnccl has no symbol ") + n; }
         return p;
      };
      api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
      api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
      api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
      api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
      api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
      api.dl = dl;
      state.store(ok ? 1 : -1, std::memory_order_release);
   }
   if (state.load(std::memory_order_acquire) != 1)
   {
      fail("rt: NCCL is not available (%s); set RT_CUDA_NCCL_LIB to the library, or pass your own sum with rt_set_allreduce", why.c_str());
      return nullptr;
   }
   return &api;
}

// the hook installed by rt_set_nccl / rt_multi: in-place sum on the library's own device buffer, on the library's stream
int nccl_allreduce_hook(void *buf, size_t count, void *stream, void *user)
{
   NcclApi *api = nccl_api();
   if (!api) return 1;
   const int rc = api->AllReduce(buf, buf, count, NCCL_DOUBLE, NCCL_SUM, (NcclComm)user, (cudaStream_t)stream);
   if (rc != 0) fprintf(stderr, "rt: ncclAllReduce failed: %s\n", api->GetErrorString(rc));
   return rc;
}
} // namespace

static void release_nccl(rt_solver *s)
{
   if (!s->nccl_comm) return;
   if (NcclApi *api = nccl_api()) api->CommDestroy((NcclComm)s->nccl_comm);
   s->nccl_comm = nullptr;
}

extern "C" int rt_nccl_unique_id(void *id128)
{
   if (!id128) return fail("rt_nccl_unique_id: null argument");
   NcclApi *api = nccl_api();
   if (!api) return 1;
   NcclId id;
   const int rc = api->GetUniqueId(&id);
   if (rc != 0) return fail("rt_nccl_unique_id: ncclGetUniqueId -> %s", api->GetErrorString(rc));
   std::memcpy(id128, id.internal, sizeof id.internal);
   return 0;
}

extern "C" int rt_set_nccl(rt_solver *s, const void *id128, int nranks, int rank)
{
   if (!s || !id128) return fail("rt_set_nccl: null argument");
   if (nranks < 1 || rank < 0 || rank >= nranks) return fail("rt_set_nccl: rank %d outside [0,%d)", rank, nranks);
   NcclApi *api = nccl_api();
   if (!api) return 1;
   CK(cudaSetDevice(s->p.device));
   release_nccl(s);
   NcclId id;
   std::memcpy(id.internal, id128, sizeof id.internal);
   NcclComm comm = nullptr;
   const int rc = api->CommInitRank(&comm, nranks, id, rank);
   if (rc != 0 || !comm) return fail("rt_set_nccl: ncclCommInitRank(rank %d of %d, device %d) -> %s", rank, nranks, s->p.device, api->GetErrorString(rc));
   s->nccl_comm = comm;
   s->allreduce = nccl_allreduce_hook;
   s->allreduce_user = comm;
   return 0;
}

// ---- helpers for rt_multi.cu (rt_internal.h)
#include "rt_internal.h"
int rt_internal_fail(const char *msg) { return fail("%s", msg); }
int rt_internal_nccl_init_all(void **comms, int ndev, const int *devices)
{
   NcclApi *api = nccl_api();
   if (!api) return 1;
   const int rc = api->CommInitAll((NcclComm *)comms, ndev, devices);
   if (rc != 0) return fail("rt: ncclCommInitAll over %d devices -> %s", ndev, api->GetErrorString(rc));
   return 0;
}
void rt_internal_nccl_destroy(void *comm)
{
   if (!comm) return;
   if (NcclApi *api = nccl_api()) api->CommDestroy((NcclComm)comm);
}
int rt_internal_nccl_hook(void *device_buf, size_t count, void *cuda_stream, void *user) { return nccl_allreduce_hook(device_buf, count, cuda_stream, user); }

int rt_internal_print_tables(rt_solver *s, rt_print_tables *t)
{
   CK(cudaSetDevice(s->p.device));
   const int G = s->p.G;
   t->G = G;
   t->T = s->p.T;
   t->kappa_grey = s->p.kappa_grey;
   t->e_edge = s->e_edge;
   t->kappa = s->kappa;
   struct { std::vector<double> *dst; const double *src; size_t n; } parts[] = {
      {&t->B, s->d_B, (size_t)G},         {&t->dBdT, s->d_dBdT, (size_t)G},   {&t->dEB, s->d_dEB, (size_t)G},
      {&t->dsigEdE, s->d_dsig, (size_t)G}, {&t->dkapEB, s->d_dkap, (size_t)G}, {&t->kappa_edge, s->d_kappa_edge, (size_t)G + 1},
      {&t->cor1, s->d_cor1, (size_t)G},   {&t->cor2, s->d_cor2, (size_t)G},   {&t->cor3, s->d_cor3, (size_t)G}};
   for (auto &p : parts)
   {
      p.dst->assign(p.n, 0.0);
      if (download(s, p.dst->data(), p.src, p.n)) return 1;
   }
   return 0;
}
