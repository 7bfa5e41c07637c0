// var_tile_kernel.cu -- the stage sweep with CELL-DEPENDENT material fields rho(c), T(c), kappa_g(c) on the machinery of the
// uniform sweep (sweep_kernel.cu): persistent CTAs, tickets, TMA tile loads / stores through SWIZZLE_128B tensor maps,
// mbarrier completion, decoupled look-back across the tiles of a chain.  EXTENSION: the
// reference's slab is uniform (solver.cpp:150-157 TODO), the oracle is the cell-dependent restatement in oracle/rt_oracle.cpp.
//
// What differs from the uniform kernel:
//   * the stepper constants depend on the cell.  The five per-(group, cell) material tables (sigma, sigma*B, cor1..3) are
//     kept TWICE, once per sweep direction ([sign][group][j], j = storage index along the sweep), so that every row reads
//     its tile of each table as one aligned TMA box, in its own order; the boxes land in shared memory next to the
//     time-level tiles, on the same mbarrier ("shared-memory staging of cross sections").  They cost 80 B per (group,
//     cell) against 64*M B of state and are re-read by the M/2 rows of a sign out of L2.
//   * the recurrence v_j = alpha_j v_(j-1) + beta_j has a cell-dependent factor, so the scan composes PAIRS
//     (A, W) = (product of alphas, local solution): (A2, W2) o (A1, W1) = (A2 A1, A2 W1 + W2) -- thread-local over CPT cells,
//     warp shuffles, cross-warp through shared memory.  Cells past the end of a row are the identity (1, 0).
//   * the look-back needs no change: a tile waits for its predecessor's INCLUSIVE value (the true outflow node upstream of
//     the tile), so the status word stays {value, flag}; the tile's own factor A never leaves the CTA.
// The per-cell constants come from the very functions of the uniform path (coef_from_material, stage_pass1).
#include <cstdlib>

#include "rt_common.cuh"
#include "sweep_device.cuh"
#include "sweep_kernel.cuh"
#include "var_kernel.cuh"
#include "var_device.cuh"

namespace rtk
{
// dynamic shared memory of one CTA: 3 X1/result slots, 2 each for X2 / XC, ONE slot of material tables (+ alignment slack)
template <int KIND, bool CORR, int THREADS, int CPT>
constexpr int var_smem_bytes()
{
   const int narr = 1 + (KIND != KIND_BE ? 1 : 0) + ((KIND == KIND_BDF && CORR) ? 1 : 0);
   return (3 + 2 * (narr - 1)) * THREADS * CPT * 16 + (CORR ? 5 : 2) * THREADS * CPT * 8 + 1024;
}
// resident CTAs per SM the register allocation aims for: what shared memory allows, at most 16 warps
template <int KIND, bool CORR, int THREADS, int CPT>
constexpr int var_min_blocks()
{
   const int by_smem = (227 * 1024) / (var_smem_bytes<KIND, CORR, THREADS, CPT>() + 1024), by_warps = 512 / THREADS;
   return by_smem < by_warps ? by_smem : by_warps;
}

template <int KIND, bool CORR, int THREADS, int CPT>
__global__ void __launch_bounds__(THREADS, var_min_blocks<KIND, CORR, THREADS, CPT>()) var_tile_kernel(const __grid_constant__ SweepMaps maps, const __grid_constant__ MatMaps mats,
                                                           const VarTileArgs va)
{
   using TS = TileShape<THREADS, CPT>;
   constexpr int TILE = TS::TILE, TILE_BYTES = TS::TILE_BYTES, WARPS = TS::WARPS;
   constexpr bool NEED_X2 = (KIND != KIND_BE);
   constexpr bool NEED_XC = (KIND == KIND_BDF) && CORR;
   constexpr int NARR = 1 + (NEED_X2 ? 1 : 0) + (NEED_XC ? 1 : 0);
   constexpr int NT = CORR ? 5 : 2;        // material tables staged per tile
   constexpr int MAT_BYTES = TILE * 8;     // one table tile: TILE doubles = TILE/2 "double2 cells" of the tensor map
   constexpr int RING = 3;
   static_assert(CPT == 4 || CPT == 8, "a thread reads its table entries as 16-byte pairs; the swizzled images are conflict-free for 4 and 8");
   static_assert(MAT_BYTES % 1024 == 0, "SWIZZLE_128B images must be 1024-byte aligned");

   const SweepArgs &a = va.s;
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   auto x1_slot = [&](int r) { return tiles + r * TILE_BYTES; };
   auto x2_slot = [&](int b) { return tiles + (RING + b) * TILE_BYTES; };
   auto xc_slot = [&](int b) { return tiles + (RING + (NEED_X2 ? 2 : 0) + b) * TILE_BYTES; };
   // ONE slot for the material tables: they are consumed by pass 1, so the next tile's boxes are issued right after the
   // barrier that ends pass 1 (half a tile of lead time) -- a second slot would cost resident CTAs
   unsigned char *const mat = tiles + (RING + 2 * (NARR - 1)) * TILE_BYTES;

   __shared__ __align__(16) double2 s_halo[2][2]; // [buffer][0: X1, 1: X2]: the cell upstream of the tile
   __shared__ __align__(8) unsigned long long s_full[2];
   __shared__ TileDesc s_desc[3];
   __shared__ long long s_ticket[3];
   __shared__ double s_wA[WARPS], s_wW[WARPS];
   __shared__ double s_ctile;

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   // two single-thread roles in DIFFERENT warps, so that neither warp carries all the serial work: thread 0 looks back and
   // publishes; the first lane of the last warp (PROD) draws and decodes tickets and issues every TMA load and store
   constexpr int PROD = THREADS - 32;
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;
   const double kbc = (KIND == KIND_BE) ? 1.0 : (KIND == KIND_CN ? 2.0 : 6.0);

   // ---- PROD only: TMA / bulk loads of one tile into pipeline buffer b (X1 into ring slot r)
   auto issue_loads = [&](const TileDesc &ti, int b, int r) {
      const bool halo = NEED_X2 && ti.tile_start > 0;
      const unsigned bytes = (unsigned)(NARR * TILE_BYTES + NT * MAT_BYTES) + (halo ? (KIND == KIND_BDF ? 32u : 16u) : 0u);
      mbar_expect_tx(&s_full[b], bytes);
      const int chunk0 = ti.tile_start >> 3;
      tma_load_3d(x1_slot(r), &maps.m[a.ix1[ti.sgn]], 0, chunk0, ti.row, &s_full[b]);
      if (NEED_X2) tma_load_3d(x2_slot(b), &maps.m[a.ix2], 0, chunk0, ti.row, &s_full[b]);
      if (NEED_XC) tma_load_3d(xc_slot(b), &maps.m[a.ixc], 0, chunk0, ti.row, &s_full[b]);
      if (halo)
      {
         bulk_load(&s_halo[b][1], a.x2 + ti.row_off - 1, 16u, &s_full[b]);
         if (KIND == KIND_BDF) bulk_load(&s_halo[b][0], a.x1[ti.sgn] + ti.row_off - 1, 16u, &s_full[b]);
      }
   };

   // PROD only: the tile's boxes of the material tables (counted in the expect_tx of issue_loads)
   auto issue_mat = [&](const TileDesc &ti, int b) {
      const int trow = ti.sgn * va.Gl + ti.gl; // the table copy in this row's sweep order
#pragma unroll
      for (int t = 0; t < NT; ++t) tma_load_3d(mat + t * MAT_BYTES, &mats.m[t], 0, ti.tile_start >> 4, trow, &s_full[b]);
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
      const long long t0 = tq.take(a, total_tiles), t1 = tq.take(a, total_tiles);
      s_ticket[0] = t0;
      s_ticket[1] = t1;
      if (t0 < total_tiles)
      {
         s_desc[0] = decode_ticket(a, cursor, t0, TILE);
         issue_loads(s_desc[0], 0, 0);
         issue_mat(s_desc[0], 0);
      }
      if (t1 < total_tiles) s_desc[1] = decode_ticket(a, cursor, t1, TILE);
      ahead = tq.take(a, total_tiles);
   }
   __syncthreads();
   long long cur = s_ticket[0];
   int buf = 0, ring = 0, slot = 0;
   unsigned phase = 0;

   while (cur < total_tiles)
   {
      const int ring_next = (ring == RING - 1) ? 0 : ring + 1;
      const int slot_next = (slot == 2) ? 0 : slot + 1, slot_ahead = (slot_next == 2) ? 0 : slot_next + 1;
      const long long nxt = s_ticket[slot_next];
      if (tid == PROD)
      {
         if (nxt < total_tiles)
         {
            bulk_wait_read<1>(); // the store that last used the target X1 slot was issued a full iteration ago
            const TileDesc dn = s_desc[slot_next];
            issue_loads(dn, buf ^ 1, ring_next);
         }
         s_ticket[slot_ahead] = ahead;
         if (ahead < total_tiles) s_desc[slot_ahead] = decode_ticket(a, cursor, ahead, TILE);
         ahead = tq.take(a, total_tiles);
      }

      const TileDesc ti = s_desc[slot];
      unsigned char *const s1 = x1_slot(ring);
      unsigned char *const s2 = x2_slot(buf);
      unsigned char *const sc = xc_slot(buf);
      const unsigned char *const sm = mat;
      double2 *status = a.tile_status + ti.status_idx;
      const int i_dir = ti.dir;
      const double mu = va.mu[i_dir];

      // early look-back probe / the chain head's boundary inflow
      double2 early = make_double2(0.0, 0.0);
      double inflow0 = 0.0;
      if (tid == 0)
      {
         if (ti.p > 0) early = ld_status(status - 1);
         else inflow0 = kbc * va.bdry[(size_t)i_dir + (size_t)a.M * (va.G0 + ti.gl)];
      }

      mbar_wait(&s_full[buf], (phase >> buf) & 1u);
      phase ^= 1u << buf;

      // ---- pass 1: per-cell stepper constants, right-hand sides and the thread-local composition (A, W)
      double alpha[CPT], gamma[CPT], beta[CPT], U[CPT];
      double A = 1.0, W = 0.0;
      {
         const int j0 = tid * CPT;
         double up1 = 0.0, up2 = 0.0;
         if (tid == 0)
         {
            if (NEED_X2 && ti.tile_start > 0) { up1 = s_halo[buf][0].y; up2 = s_halo[buf][1].y; }
         }
         else
         {
            if (KIND == KIND_BDF) up1 = reinterpret_cast<const double2 *>(s1 + swz(j0 - 1))->y;
            if (NEED_X2) up2 = reinterpret_cast<const double2 *>(s2 + swz(j0 - 1))->y;
         }
         // the thread's CPT entries of every table are CPT/2 16-byte pairs of the swizzled image (conflict-free for 4 and 8)
         auto mat_pair = [&](int t, int pp) { return *reinterpret_cast<const double2 *>(sm + t * MAT_BYTES + swz((j0 >> 1) + pp)); };
         const RowConst<KIND, CORR> rc(mu, va.dx, va.tau, va.dt, va.beta);
         const bool row_head = ti.tile_start == 0 && tid == 0;
         auto cell = [&](int jj, bool masked, double sig, double srcB, double cor1, double cor2, double cor3) {
            const int j = j0 + jj;
            const RowCoef c = rc.cell(sig, srcB, cor1, cor2, cor3);
            const int o = swz(j);
            double2 X1[1], X2[1];
            double xs[1], wl[1], Uc[1];
            X1[0] = *reinterpret_cast<const double2 *>(s1 + o);
            X2[0] = NEED_X2 ? *reinterpret_cast<const double2 *>(s2 + o) : X1[0];
            if (KIND == KIND_BDF)
            {
               xs[0] = 0.0;
               if (CORR) { const double2 q2 = *reinterpret_cast<const double2 *>(sc + o); xs[0] = q2.x + q2.y; }
            }
            else xs[0] = X1[0].x + X1[0].y;
            stage_pass1<KIND, CORR, 1>(c, X1, X2, xs, up1, up2, row_head && jj == 0, wl, Uc);
            const bool ok = !masked || j < ti.valid; // cells past the end of the row are the identity element
            alpha[jj] = ok ? c.alpha : 1.0;
            gamma[jj] = ok ? c.gamma : 0.0;
            beta[jj] = ok ? wl[0] : 0.0;
            U[jj] = ok ? Uc[0] : 0.0;
            W = fma(alpha[jj], W, beta[jj]);
            A *= alpha[jj];
            if (KIND == KIND_BDF) up1 = X1[0].y;
            if (NEED_X2) up2 = X2[0].y;
         };
         auto run = [&](bool masked) {
#pragma unroll
            for (int pp = 0; pp < CPT / 2; ++pp)
            {
               double2 m[5];
#pragma unroll
               for (int t = 0; t < 5; ++t) m[t] = (t < NT) ? mat_pair(t, pp) : make_double2(0.0, 0.0);
               cell(2 * pp, masked, m[0].x, m[1].x, m[2].x, m[3].x, m[4].x);
               cell(2 * pp + 1, masked, m[0].y, m[1].y, m[2].y, m[3].y, m[4].y);
            }
         };
         if (ti.valid == TILE) run(false);
         else run(true);
      }

      // ---- block scan of the pairs
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
      if (lane == 31) { s_wA[warp] = Ai; s_wW[warp] = Wi; }
      __syncthreads(); // (C) warp totals visible; nobody reads the material tables of this tile any more
      if (tid == PROD && nxt < total_tiles) issue_mat(s_desc[slot_next], buf ^ 1);
      double Ap = 1.0, Wp = 0.0; // composition of the warps before this one
      for (int w2 = 0; w2 < warp; ++w2)
      {
         Wp = fma(s_wA[w2], Wp, s_wW[w2]);
         Ap *= s_wA[w2];
      }
      const double Wx = fma(Ae, Wp, We), Ax = Ae * Ap; // everything upstream of this thread inside the tile

      // ---- decoupled look-back (one thread): always the predecessor's inclusive value (bitwise reproducible)
      if (tid == 0)
      {
         double At = 1.0, Wt = 0.0;
         for (int w2 = 0; w2 < WARPS; ++w2)
         {
            Wt = fma(s_wA[w2], Wt, s_wW[w2]);
            At *= s_wA[w2];
         }
         double ctile;
         if (ti.p == 0) ctile = inflow0;
         else
         {
            double2 w = early;
            long long spins = 0;
            while ((int)(status_flag(w) - flag_incl) < 0)
            {
               __nanosleep(32);
               w = ld_status(status - 1);
               if (++spins > SPIN_LIMIT) __trap();
            }
            ctile = ti.junction ? kbc * w.x : w.x;
         }
         st_status(status, fma(At, ctile, Wt), flag_incl);
         s_ctile = ctile;
      }
      __syncthreads(); // (E)

      // ---- pass 2: true nodes, written back into the X1 tile image
      {
         double v = fma(Ax, s_ctile, Wx);
         const int j0 = tid * CPT;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double vn = fma(alpha[jj], v, beta[jj]);
            *reinterpret_cast<double2 *>(s1 + swz(j0 + jj)) = make_double2(fma(gamma[jj], v, U[jj]), vn);
            v = vn;
         }
      }
      fence_proxy_async();
      __syncthreads(); // (F)

      if (tid == PROD)
      {
         tma_store_3d(&maps.m[a.iout], 0, ti.tile_start >> 3, ti.row, s1); // rows that end inside the tile are clipped
         bulk_commit();
      }
      cur = nxt;
      buf ^= 1;
      ring = ring_next;
      slot = slot_next;
   }
   if (tid == PROD)
   {
      bulk_wait_read<0>();
      tq.finish(a, total_tiles);
   }
}

// ------------------------------------------------------------------------------------------ host side
template <int KIND, bool CORR, int THREADS, int CPT>
static cudaError_t var_launch_one(const SweepMaps &maps, const MatMaps &mats, const VarTileArgs &va, int num_sms, cudaStream_t st,
                                  int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   int ctas_per_sm = plan_slot->load(std::memory_order_acquire);
   auto k = var_tile_kernel<KIND, CORR, THREADS, CPT>;
   const size_t sm = (size_t)var_smem_bytes<KIND, CORR, THREADS, CPT>();
   if (ctas_per_sm <= 0)
   {
      cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
      if (e != cudaSuccess) return e;
      int occ = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, THREADS, sm);
      if (e != cudaSuccess) return e;
      ctas_per_sm = occ < 1 ? 1 : occ;
      plan_slot->store(ctas_per_sm, std::memory_order_release);
   }
   const long long total = (long long)va.s.n_chains * va.s.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   k<<<(unsigned)grid, THREADS, sm, st>>>(maps, mats, va);
   return cudaGetLastError();
}

template <int T, int CPT>
static cudaError_t var_launch_shape(int kind, bool corr, const SweepMaps &maps, const MatMaps &mats, const VarTileArgs &va, int num_sms,
                                    cudaStream_t st, int *grid_out)
{
   switch (kind * 2 + (corr ? 1 : 0))
   {
   case 0: return var_launch_one<KIND_BE, false, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   case 1: return var_launch_one<KIND_BE, true, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   case 2: return var_launch_one<KIND_CN, false, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   case 3: return var_launch_one<KIND_CN, true, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   case 4: return var_launch_one<KIND_BDF, false, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   case 5: return var_launch_one<KIND_BDF, true, T, CPT>(maps, mats, va, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}

bool var_tile_shape_supported(int threads, int cpt) { return (threads == 64 && (cpt == 4 || cpt == 8)) || (threads == 128 && cpt == 4); }

cudaError_t launch_var_tile(int kind, bool corr, int threads, int cpt, const SweepMaps &maps, const MatMaps &mats, const VarTileArgs &va,
                            int num_sms, cudaStream_t st, int *grid_out)
{
   if (threads == 64 && cpt == 4) return var_launch_shape<64, 4>(kind, corr, maps, mats, va, num_sms, st, grid_out);
   if (threads == 64 && cpt == 8) return var_launch_shape<64, 8>(kind, corr, maps, mats, va, num_sms, st, grid_out);
   if (threads == 128 && cpt == 4) return var_launch_shape<128, 4>(kind, corr, maps, mats, va, num_sms, st, grid_out);
   return cudaErrorInvalidValue;
}
} // namespace rtk
