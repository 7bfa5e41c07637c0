// var_step_kernel.cu -- one whole BDF2 time step (BE, CN, BE, BDF; solver.cpp:606-820 with `_it % 4` = 0..3) in ONE pass
// over the state for CELL-DEPENDENT material fields: the fused step of step_kernel.cu (same stage algebra, same four
// look-back chains per tile, same half_ends handling) with the per-cell stepper constants, TMA-staged material tables
// and pair scan of var_tile_kernel.cu.  EXTENSION; oracle: the cell-dependent restatement.
//
//   stage 0  BE (tau=dt/2)   x1 = X (ends^n)                                   -> S0   (shared memory)
//   stage 1  CN              x1 = S0, x2 = X, upstream X          -> S1   (shared memory)
//   stage 2  BE              x1 = S1                                            -> S2   (only u+v kept, in registers)
//   stage 3  BDF             x1 = half = S1 (mu<0 rows) / S0 (mu>0 rows), x2 = X, source from S2 -> ends^(n+1)
//
// Per unknown and step 16 B in + 16 B out (+ the material boxes, 40 B per (group, cell) and sweep direction, shared by the
// M/2 rows of a sign through L2) instead of 176 B: the per-stage kernel is within 0.7-0.8 of the HBM ceiling, this one is
// bound by instruction issue / fp64 latency only.  The boxes of a tile serve all four stages (BE constants are formed
// twice: recomputing ~25 flops per cell is cheaper than 16 more live registers).
// Roles: thread 0 looks back and publishes (four words per tile), the first lane of the last warp draws tickets and issues
// every TMA load and store.
#include <cstdlib>
#include <type_traits>

#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"
#include "sweep_kernel.cuh"
#include "var_device.cuh"
#include "var_kernel.cuh"

namespace rtk
{
#ifndef VAR_STEP_REGS
#define VAR_STEP_REGS 168 // register budget per thread the resident-CTA target is derived from (A/B: -DVAR_STEP_REGS=128)
#endif
// MB = slots of material tables: 2 (loaded with X) or 1 (the tables are read once, at the start of a tile, so the next
// tile's boxes are issued after the tile's first barrier: almost a full tile of lead time and one more resident CTA)
template <bool CORR, int THREADS, int CPT, int MB>
constexpr int var_step_smem_bytes()
{
   return 4 * THREADS * CPT * 16 + MB * (CORR ? 5 : 2) * THREADS * CPT * 8 + 1024; // 2 X + S0 + S1 + MB x tables + slack
}
template <bool CORR, int THREADS, int CPT, int MB>
constexpr int var_step_min_blocks()
{
   const int by_smem = (227 * 1024) / (var_step_smem_bytes<CORR, THREADS, CPT, MB>() + 1024), by_regs = 65536 / (THREADS * VAR_STEP_REGS);
   return by_smem < by_regs ? by_smem : by_regs;
}

// NSTEPS: warp-scan steps that are run (5 = all; fewer when launch_var_scan_steps says that the later ones cannot matter for
// ANY row of the solver: a compile-time constant, so the full kernel carries nothing for it)
template <bool CORR, int THREADS, int CPT, int MB, int NSTEPS>
__global__ void __launch_bounds__(THREADS, var_step_min_blocks<CORR, THREADS, CPT, MB>())
var_step_kernel(const __grid_constant__ SweepMaps maps, const __grid_constant__ MatMaps mats, const VarStepArgs va)
{
   using TS = TileShape<THREADS, CPT>;
   constexpr int TILE = TS::TILE, TILE_BYTES = TS::TILE_BYTES, WARPS = TS::WARPS;
   constexpr int NT = CORR ? 5 : 2, MAT_BYTES = TILE * 8;
   constexpr int PROD = THREADS - 32;
   static_assert(CPT == 4 || CPT == 8, "table entries are read as 16-byte pairs of the swizzled image");

   const StepArgs &a = va.s;
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   auto x_slot = [&](int b) { return tiles + b * TILE_BYTES; };
   unsigned char *const t0 = tiles + 2 * TILE_BYTES;
   unsigned char *const t1 = tiles + 3 * TILE_BYTES;
   auto mat_slot = [&](int b) { return tiles + 4 * TILE_BYTES + (MB == 2 ? b : 0) * NT * MAT_BYTES; };

   __shared__ __align__(16) double2 s_halo[2];
   __shared__ __align__(8) unsigned long long s_full[2];
   __shared__ TileDesc s_desc[3];
   __shared__ long long s_ticket[3];
   __shared__ double s_wA[WARPS], s_wW[WARPS];
   __shared__ double s_ctile;

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;

   // ---- PROD only: TMA / bulk loads of one tile (X, the material boxes, the upstream cell) into pipeline buffer b
   auto issue_loads = [&](const TileDesc &ti, int b) {
      const bool halo = ti.tile_start > 0;
      mbar_expect_tx(&s_full[b], (unsigned)(TILE_BYTES + NT * MAT_BYTES) + (halo ? 16u : 0u));
      tma_load_3d(x_slot(b), &maps.m[a.ix], 0, ti.tile_start >> 3, ti.row, &s_full[b]);
      if (halo) bulk_load(&s_halo[b], a.x + ti.row_off - 1, 16u, &s_full[b]);
   };
   // PROD only: the tile's boxes of the material tables (counted in the expect_tx of issue_loads)
   auto issue_mat = [&](const TileDesc &ti, int b) {
      const int trow = ti.sgn * va.Gl + ti.gl; // the table copy in this row's sweep order
#pragma unroll
      for (int t = 0; t < NT; ++t) tma_load_3d(mat_slot(b) + t * MAT_BYTES, &mats.m[t], 0, ti.tile_start >> 4, trow, &s_full[b]);
   };

   TicketQueue tq;
   TicketCursor cursor;
   cursor.ticket = -2;
   long long ahead = 0;
   if (tid == PROD)
   {
      mbar_init(&s_full[0], 1);
      mbar_init(&s_full[1], 1);
      fence_proxy_async();
      const long long q0 = tq.take(a, total_tiles), q1 = tq.take(a, total_tiles);
      s_ticket[0] = q0;
      s_ticket[1] = q1;
      if (q0 < total_tiles)
      {
         s_desc[0] = decode_ticket(a, cursor, q0, TILE);
         issue_loads(s_desc[0], 0);
         issue_mat(s_desc[0], 0);
      }
      if (q1 < total_tiles) s_desc[1] = decode_ticket(a, cursor, q1, TILE);
      ahead = tq.take(a, total_tiles);
   }
   __syncthreads();
   long long cur = s_ticket[0];
   int buf = 0, slot = 0;
   unsigned phase = 0;

   while (cur < total_tiles)
   {
      const int slot_next = (slot == 2) ? 0 : slot + 1, slot_ahead = (slot_next == 2) ? 0 : slot_next + 1;
      const long long nxt = s_ticket[slot_next];
      const TileDesc ti = s_desc[slot];
      double2 *const status = a.tile_status + 4 * ti.status_idx; // 4 words: one look-back chain per stage
      double2 early[4];
      double pred_v0 = 0.0, pred_v1 = 0.0, bdry0 = 0.0; // thread 0 only
      if (tid == PROD)
      {
         bulk_wait_read<0>(); // the previous tile's store has read S0 / S1 (pass 2 of stage 0 overwrites them two barriers on)
         if (nxt < total_tiles)
         {
            const TileDesc dn = s_desc[slot_next];
            issue_loads(dn, buf ^ 1); // buffer buf^1 was last read in the previous tile
            if (MB == 2) issue_mat(dn, buf ^ 1);
         }
         s_ticket[slot_ahead] = ahead;
         if (ahead < total_tiles) s_desc[slot_ahead] = decode_ticket(a, cursor, ahead, TILE);
         ahead = tq.take(a, total_tiles);
      }
      if (tid == 0)
      {
#pragma unroll
         for (int k = 0; k < 4; ++k) early[k] = (ti.p > 0) ? ld_status(status - 4 + k) : make_double2(0.0, 0.0);
         if (ti.p == 0) bdry0 = va.bdry[(size_t)ti.dir + (size_t)a.M * (va.G0 + ti.gl)];
      }

      mbar_wait(&s_full[buf], (phase >> buf) & 1u);
      phase ^= 1u << buf;

      const unsigned char *const sx = x_slot(buf);
      const unsigned char *const sm = mat_slot(buf);
      const int j0 = tid * CPT;
      const bool row_first = (ti.tile_start == 0) && (tid == 0);
      const bool full = ti.valid == TILE;
      // last lane of this warp that holds a real cell (31 when the warp is full, or empty: all identity)
      const int wl_ = (ti.valid - 1) / CPT - 32 * warp, wlast = (wl_ < 0 || wl_ > 31) ? 31 : wl_;
      const double mu = va.mu[ti.dir];
      double2 xr[CPT];
#pragma unroll
      for (int jj = 0; jj < CPT; ++jj) xr[jj] = *reinterpret_cast<const double2 *>(sx + swz(j0 + jj));
      double up_x = 0.0;
      if (tid == 0) { if (ti.tile_start > 0) up_x = s_halo[buf].y; }
      else up_x = reinterpret_cast<const double2 *>(sx + swz(j0 - 1))->y;

      // the cell's material enters every stage through sigma and the source terms qa, qb only, and the source terms do not
      // depend on the stage kind (same tau = dt/2): read the staged tables ONCE per tile, keep 3 values per cell
      double sigc[CPT], qac[CPT], qbc[CPT];
      {
         const RowConst<KIND_BE, CORR> rq(mu, va.dx, va.tau, va.dt, va.beta);
#pragma unroll
         for (int pp = 0; pp < CPT / 2; ++pp)
         {
            double2 m[5];
#pragma unroll
            for (int t = 0; t < 5; ++t)
               m[t] = (t < NT) ? *reinterpret_cast<const double2 *>(sm + t * MAT_BYTES + swz((j0 >> 1) + pp)) : make_double2(0.0, 0.0);
            sigc[2 * pp] = m[0].x; sigc[2 * pp + 1] = m[0].y;
            rq.source_terms(m[1].x, m[2].x, m[3].x, m[4].x, qac[2 * pp], qbc[2 * pp]);
            rq.source_terms(m[1].y, m[2].y, m[3].y, m[4].y, qac[2 * pp + 1], qbc[2 * pp + 1]);
         }
      }

      // ---- pass 1 of one stage: per-cell constants, right-hand sides, thread-local pair (A, W).
      // x1 / x2 / xs as in stage_pass1 (sweep_device.cuh); cells past the end of the row are the identity element.
      double alpha[CPT], gamma[CPT], beta[CPT], U[CPT];
      auto pass1m = [&](auto kindc, auto maskc, const double2 (&x1)[CPT], const double2 (&x2)[CPT], const double (&xs)[CPT], double up1,
                        double up2, double &A, double &W) {
         constexpr int KIND = decltype(kindc)::value;
         constexpr bool MASKED = decltype(maskc)::value;
         const RowConst<KIND, CORR> rc(mu, va.dx, va.tau, va.dt, va.beta);
         A = 1.0; W = 0.0;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const RowCoef c = rc.cell_q(sigc[jj], qac[jj], qbc[jj]);
            double2 X1[1] = {x1[jj]}, X2[1] = {x2[jj]};
            double XS[1] = {xs[jj]}, wl[1], Uc[1];
            stage_pass1<KIND, CORR, 1>(c, X1, X2, XS, up1, up2, row_first && jj == 0, wl, Uc);
            const bool ok = !MASKED || (j0 + jj) < ti.valid;
            alpha[jj] = ok ? c.alpha : 1.0;
            gamma[jj] = ok ? c.gamma : 0.0;
            beta[jj] = ok ? wl[0] : 0.0;
            U[jj] = ok ? Uc[0] : 0.0;
            W = fma(alpha[jj], W, beta[jj]);
            A *= alpha[jj];
            if (KIND == KIND_BDF) up1 = x1[jj].y;
            if (KIND != KIND_BE) up2 = x2[jj].y;
         }
      };

      // full tiles (all but the last tile of a row) take the copy without the identity selects
      auto pass1 = [&](auto kindc, const double2 (&x1)[CPT], const double2 (&x2)[CPT], const double (&xs)[CPT], double up1, double up2,
                       double &A, double &W) {
         if (full) pass1m(kindc, std::false_type{}, x1, x2, xs, up1, up2, A, W);
         else pass1m(kindc, std::true_type{}, x1, x2, xs, up1, up2, A, W);
      };

      // ---- block scan of the pairs + look-back chain k: returns the true inflow of this thread's chunk
      auto scan = [&](auto kc, const double kbc, double A, double W) -> double {
         constexpr int k = decltype(kc)::value;
         double Ai = A, Wi = W;
#pragma unroll
         for (int d = 1; d < (1 << NSTEPS); d <<= 1)
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
         // The warp's aggregate.  A row's last tile is padded with identity elements (alpha = 1), whose products do not die
         // out: with the short scan the aggregate would not travel from the last real cell to lane 31 (a reflective left
         // boundary reads the tile's aggregate), so it is taken where it is complete -- at the warp's last real lane.
         if (lane == (NSTEPS < 5 ? wlast : 31)) { s_wA[warp] = Ai; s_wW[warp] = Wi; }
         __syncthreads(); // (C)
         if (MB == 1 && k == 0 && tid == PROD && nxt < total_tiles) issue_mat(s_desc[slot_next], buf ^ 1); // the tables were read at the tile's start
         double Ap = 1.0, Wp = 0.0;
         for (int w2 = 0; w2 < warp; ++w2)
         {
            Wp = fma(s_wA[w2], Wp, s_wW[w2]);
            Ap *= s_wA[w2];
         }
         const double Wx = fma(Ae, Wp, We), Ax = Ae * Ap;
         if (tid == 0)
         {
            double At = 1.0, Wt = 0.0;
            for (int w2 = 0; w2 < WARPS; ++w2)
            {
               Wt = fma(s_wA[w2], Wt, s_wW[w2]);
               At *= s_wA[w2];
            }
            double ctile;
            if (ti.p == 0) ctile = kbc * bdry0;
            else
            {
               double2 w = early[k];
               long long spins = 0;
               while ((int)(status_flag(w) - flag_incl) < 0)
               {
                  __nanosleep(32);
                  w = ld_status(status - 4 + k);
                  if (++spins > SPIN_LIMIT) __trap();
               }
               ctile = ti.junction ? kbc * w.x : w.x;
               if (k == 0) pred_v0 = w.x;
               if (k == 1) pred_v1 = w.x;
            }
            st_status(status + k, fma(At, ctile, Wt), flag_incl);
            s_ctile = ctile;
         }
         __syncthreads(); // (E)
         return fma(Ax, s_ctile, Wx);
      };

      double2 r[CPT];
      double xsum[CPT], A, W;

      // ================= stage 0: BE predictor, ends^n -> S0
      {
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = xr[jj].x + xr[jj].y;
         pass1(std::integral_constant<int, KIND_BE>{}, xr, xr, xsum, 0.0, 0.0, A, W);
         double v = scan(std::integral_constant<int, 0>{}, 1.0, A, W);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double vn = fma(alpha[jj], v, beta[jj]);
            r[jj] = make_double2(fma(gamma[jj], v, U[jj]), vn);
            *reinterpret_cast<double2 *>(t0 + swz(j0 + jj)) = r[jj];
            v = vn;
         }
      }
      // ================= stage 1: CN corrector, x1 = S0 (own cells, registers), x2 = X -> S1
      {
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
         pass1(std::integral_constant<int, KIND_CN>{}, r, xr, xsum, 0.0, up_x, A, W);
         double v = scan(std::integral_constant<int, 1>{}, 2.0, A, W);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double vn = fma(alpha[jj], v, beta[jj]);
            r[jj] = make_double2(fma(gamma[jj], v, U[jj]), vn);
            *reinterpret_cast<double2 *>(t1 + swz(j0 + jj)) = r[jj];
            v = vn;
         }
      }
      // ================= stage 2: BE predictor, x1 = S1 -> S2 (only u+v survives: the BDF stage's correction source; without
      // that source nothing reads the stage -- it is left out, its scan and look-back chain included, cf. step_warp_kernel.cu)
      if (CORR)
      {
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
         pass1(std::integral_constant<int, KIND_BE>{}, r, r, xsum, 0.0, 0.0, A, W);
         double v = scan(std::integral_constant<int, 2>{}, 1.0, A, W);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double vn = fma(alpha[jj], v, beta[jj]);
            xsum[jj] = fma(gamma[jj], v, U[jj]) + vn;
            v = vn;
         }
      }
      else __syncthreads(); // stage 3 reads the neighbouring thread's cell of S1, written in stage 1: the order stage 2's barriers gave
      // ================= stage 3: BDF corrector, x1 = half (S1 for mu<0 rows, S0 for mu>0 rows), x2 = X -> ends^(n+1)
      {
         const unsigned char *const sh = ti.sgn ? t0 : t1;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) r[jj] = *reinterpret_cast<const double2 *>(sh + swz(j0 + jj));
         double up_h = 0.0;
         if (tid == 0) { if (ti.tile_start > 0) up_h = ti.sgn ? pred_v0 : pred_v1; } // the predecessor tile's last cell (F7)
         else up_h = reinterpret_cast<const double2 *>(sh + swz(j0 - 1))->y;
         pass1(std::integral_constant<int, KIND_BDF>{}, r, xr, xsum, up_h, up_x, A, W);
         double v = scan(std::integral_constant<int, 3>{}, 6.0, A, W);
         unsigned char *const so = ti.sgn ? t1 : t0; // the tile that is NOT half_ends: nobody reads it any more
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double vn = fma(alpha[jj], v, beta[jj]);
            *reinterpret_cast<double2 *>(so + swz(j0 + jj)) = make_double2(fma(gamma[jj], v, U[jj]), vn);
            v = vn;
         }
         fence_proxy_async();
         __syncthreads(); // (F)
         if (tid == PROD)
         {
            tma_store_3d(&maps.m[a.iout], 0, ti.tile_start >> 3, ti.row, so); // rows that end inside the tile are clipped
            bulk_commit();
         }
      }
      cur = nxt;
      buf ^= 1;
      slot = slot_next;
   }
   if (tid == PROD)
   {
      bulk_wait_read<0>();
      tq.finish(a, total_tiles);
   }
}

// ------------------------------------------------------------------------------------------ host side
// The largest number of warp-scan steps any row of the solver needs.  Step s folds in the lane 2^s below with the product
// of the alphas of cpt 2^s cells; alpha falls with sigma, so over a group it is at most alpha(sigma_min); once that bound
// is below 2^-70 the step cannot change a result by more than 2^-70 of a neighbouring value.  sigma never changes during
// a solve (the temperature update moves T, not rho kappa), so this is evaluated once.
__global__ void var_scan_steps_kernel(const double *__restrict__ sigma_fwd, int pitch, int N, const double *__restrict__ mu, int M,
                                      double dx, double tau, double dt, double beta, int cpt, int *__restrict__ max_steps)
{
   __shared__ double s_min[32];
   const int gl = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   double m = 1e300;
   for (int c = tid; c < N; c += blockDim.x) m = fmin(m, sigma_fwd[(size_t)gl * pitch + c]);
   for (int d = 16; d > 0; d >>= 1) m = fmin(m, __shfl_xor_sync(0xffffffffu, m, d));
   if (lane == 0) s_min[warp] = m;
   __syncthreads();
   if (tid == 0)
   {
      for (int w = 1; w < (int)(blockDim.x >> 5); ++w) m = fmin(m, s_min[w]);
      s_min[0] = m;
   }
   __syncthreads();
   const double smin = s_min[0];
   for (int i = tid; i < M; i += blockDim.x)
   {
      const double a = fmax(fmax(RowConst<KIND_BE, false>(mu[i], dx, tau, dt, beta).cell_q(smin, 0.0, 0.0).alpha,
                                 RowConst<KIND_CN, false>(mu[i], dx, tau, dt, beta).cell_q(smin, 0.0, 0.0).alpha),
                            RowConst<KIND_BDF, false>(mu[i], dx, tau, dt, beta).cell_q(smin, 0.0, 0.0).alpha);
      double f = a; // alpha_max^(cpt 2^s), s = 0, 1, ...
      for (int j = 1; j < cpt; ++j) f *= a;
      int n = 0;
      for (int sidx = 0; sidx < 5; ++sidx)
      {
         if (!(f < 0x1p-70)) n = sidx + 1;
         f = f * f;
      }
      atomicMax(max_steps, n);
   }
}

void launch_var_scan_steps(const double *sigma_fwd, int pitch, int N, int groups, const double *mu, int M, double dx, double tau, double dt,
                           double beta, int cpt, int *max_steps, cudaStream_t st)
{
   cudaMemsetAsync(max_steps, 0, sizeof(int), st);
   var_scan_steps_kernel<<<groups, 256, 0, st>>>(sigma_fwd, pitch, N, mu, M, dx, tau, dt, beta, cpt, max_steps);
}

template <bool CORR, int THREADS, int CPT, int MB, int NSTEPS>
static cudaError_t var_step_launch_one(const SweepMaps &maps, const MatMaps &mats, const VarStepArgs &va, int num_sms, cudaStream_t st,
                                       int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   int ctas_per_sm = plan_slot->load(std::memory_order_acquire);
   auto kern = var_step_kernel<CORR, THREADS, CPT, MB, NSTEPS>;
   const size_t smem = (size_t)var_step_smem_bytes<CORR, THREADS, CPT, MB>();
   if (ctas_per_sm <= 0)
   {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      int occ = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem);
      if (e != cudaSuccess) return e;
      ctas_per_sm = occ > 0 ? occ : 1;
      plan_slot->store(ctas_per_sm, std::memory_order_release);
   }
   const long long total = (long long)va.s.n_chains * va.s.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, THREADS, smem, st>>>(maps, mats, va);
   return cudaGetLastError();
}

bool var_step_shape_supported(int threads, int cpt) { return (threads == 64 || threads == 128) && cpt == 4; }

cudaError_t launch_var_step(bool corr, int threads, int cpt, int mat_slots, int scan_steps, const SweepMaps &maps, const MatMaps &mats,
                            const VarStepArgs &va, int num_sms, cudaStream_t st, int *grid_out)
{
   if (cpt != 4) return cudaErrorInvalidValue;
   // the short-scan instantiation exists for the default shape only
   if (threads == 64 && mat_slots == 1 && scan_steps <= 2)
      return corr ? var_step_launch_one<true, 64, 4, 1, 2>(maps, mats, va, num_sms, st, grid_out)
                  : var_step_launch_one<false, 64, 4, 1, 2>(maps, mats, va, num_sms, st, grid_out);
   if (threads == 64 && mat_slots == 1) return corr ? var_step_launch_one<true, 64, 4, 1, 5>(maps, mats, va, num_sms, st, grid_out)
                                                    : var_step_launch_one<false, 64, 4, 1, 5>(maps, mats, va, num_sms, st, grid_out);
   if (threads == 64) return corr ? var_step_launch_one<true, 64, 4, 2, 5>(maps, mats, va, num_sms, st, grid_out)
                                  : var_step_launch_one<false, 64, 4, 2, 5>(maps, mats, va, num_sms, st, grid_out);
   if (threads == 128) return corr ? var_step_launch_one<true, 128, 4, 2, 5>(maps, mats, va, num_sms, st, grid_out)
                                   : var_step_launch_one<false, 128, 4, 2, 5>(maps, mats, va, num_sms, st, grid_out);
   return cudaErrorInvalidValue;
}
} // namespace rtk
