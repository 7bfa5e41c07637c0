// step_kernel.cu -- one whole BDF2 time step (BE predictor, CN corrector, BE predictor, BDF corrector;
// reference src/solver.cpp:606-820 with `_it % 4` = 0..3) in ONE pass over the state.
//
// Why: run stage by stage (sweep_kernel.cu) a step moves 32 + 48 + 32 + 64 = 176 bytes per unknown through HBM --
// every stage re-reads what the previous one has just written.  But all four stages sweep a row in the SAME order,
// and stage k+1 of a cell needs only stage-k values of that cell and old-level values of the upstream cell.  So a CTA
// that holds a tile of ends^n in shared memory can run all four stage scans on it back to back and write only the
// final ends^(n+1): 16 bytes in, 16 bytes out per unknown and step -- 5.5x less traffic, the kernel becomes fp64-bound.
//
//   stage 0  BE (tau=dt/2)   x1 = X (ends^n)                                   -> S0   (kept in shared memory)
//   stage 1  CN              x1 = S0, x2 = X, upstream X          -> S1   (kept in shared memory)
//   stage 2  BE              x1 = S1                                            -> S2   (only u+v kept, in registers)
//   stage 3  BDF             x1 = half = S1 for mu<0 rows, S0 for mu>0 rows, x2 = X, source from S2,
//                            upstream half and X                                -> ends^(n+1)  (TMA store)
//
// Each stage has its own affine recurrence along the row, hence its own look-back chain: a tile owns FOUR status
// words.  The old-level outflow nodes of the cell upstream of a tile that the per-stage kernel reads from the half
// buffer in HBM do not exist in memory here -- they are exactly the predecessor tile's inclusive look-back values of
// stage 0 / stage 1, which the look-back has fetched anyway.  Reflective junctions, boundary rule, tickets, TMA
// staging, L2 prefetch: as in sweep_kernel.cu.  prev_ends stays available (the step is out of place); half_ends is
// re-materialised on demand by two per-stage launches (rt_capi.cu).
#include <cstdlib>
#include <type_traits>

#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"

namespace rtk
{
template <bool CORR, int THREADS, int CPT>
__global__ void __launch_bounds__(THREADS, (THREADS * CPT == 512) ? 384 / THREADS : 1) step_kernel(const __grid_constant__ SweepMaps maps, const StepArgs a)
{
   using TS = TileShape<THREADS, CPT>;
   constexpr int TILE = TS::TILE, TILE_BYTES = TS::TILE_BYTES, WARPS = TS::WARPS;
   static_assert(CPT == 8 || CPT == 4, "the swizzled tile image is conflict-free for 4 or 8 cells per thread");
   static_assert((WARPS & (WARPS - 1)) == 0, "WARPS must be a power of two");

   // dynamic shared memory, 1024-byte aligned by hand: [2 x X][S0][S1].  The result tile of stage 3 is written over
   // whichever of S0 / S1 is NOT this row's half_ends (S1 for mu>0 rows, S0 for mu<0 rows): that tile has no reader
   // left in shared memory, so no separate output slot is needed (4 tiles instead of 6 -> 6 resident CTAs per SM).
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   auto x_slot = [&](int b) { return tiles + b * TILE_BYTES; };
   unsigned char *const t0 = tiles + 2 * TILE_BYTES;
   unsigned char *const t1 = tiles + 3 * TILE_BYTES;

   __shared__ __align__(16) RowCoef s_coef[2][3]; // [buffer][BE, CN, BDF]
   __shared__ __align__(16) double2 s_halo[2];    // [buffer]: X at the cell upstream of the tile
   __shared__ __align__(8) unsigned long long s_full[2];
   __shared__ TileDesc s_desc[3];
   __shared__ long long s_ticket[3];
   __shared__ double s_warp_tot[WARPS];
   __shared__ double s_agg, s_alast, s_ctile;

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;

   // ---- thread 0 only: TMA / bulk loads of one tile into pipeline buffer b
   auto issue_loads = [&](const TileDesc &ti, int b) {
      const bool halo = ti.tile_start > 0;
      mbar_expect_tx(&s_full[b], (unsigned)(TILE_BYTES + 3 * sizeof(RowCoef)) + (halo ? 16u : 0u));
      tma_load_3d(x_slot(b), &maps.m[a.ix], 0, ti.tile_start >> 3, ti.row, &s_full[b]);
#pragma unroll
      for (int k = 0; k < 3; ++k)
         bulk_load(&s_coef[b][k], a.coef[k] + (size_t)ti.row * (sizeof(RowCoef) / 16), (unsigned)sizeof(RowCoef), &s_full[b]);
      if (halo) bulk_load(&s_halo[b], a.x + ti.row_off - 1, 16u, &s_full[b]);
   };

   // ---- prologue: barriers armed, two tickets drawn and decoded, the first tile in flight, a third ticket drawn
   TicketQueue tq;
   TicketCursor cursor;
   cursor.ticket = -2;
   long long ahead = 0;
   if (tid == 0)
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
      double pred_v0 = 0.0, pred_v1 = 0.0; // thread 0: stage-0 / stage-1 outflow node of the cell upstream of the tile
      if (tid == 0)
      {
         if (nxt < total_tiles)
         {
            const TileDesc dn = s_desc[slot_next];
            issue_loads(dn, buf ^ 1); // X slot buf^1 was last read in the previous tile, three barriers ago
         }
         s_ticket[slot_ahead] = ahead;
         if (ahead < total_tiles)
         {
            const TileDesc d_ahead = decode_ticket(a, cursor, ahead, TILE);
            if (a.l2_prefetch) prefetch_l2_bulk(a.x + d_ahead.row_off, (unsigned)d_ahead.valid * 16u);
            s_desc[slot_ahead] = d_ahead;
         }
         ahead = tq.take(a, total_tiles);
         // early look-back probes: by the time a stage's local scan is done the predecessor's word has arrived
#pragma unroll
         for (int k = 0; k < 4; ++k) early[k] = (ti.p > 0) ? ld_status(status - 4 + k) : make_double2(0.0, 0.0);
      }

      mbar_wait(&s_full[buf], (phase >> buf) & 1u); // this tile's bytes have landed
      phase ^= 1u << buf;

      const unsigned char *const sx = x_slot(buf);
      const int j0 = tid * CPT;
      const bool row_first = (ti.tile_start == 0) && (tid == 0);
      // own cells of X = ends^n (prev_ends of this step) and the X outflow node of the upstream cell
      double2 xr[CPT];
#pragma unroll
      for (int jj = 0; jj < CPT; ++jj) xr[jj] = *reinterpret_cast<const double2 *>(sx + swz(j0 + jj));
      double up_x = 0.0;
      if (tid == 0) { if (ti.tile_start > 0) up_x = s_halo[buf].y; }
      else up_x = reinterpret_cast<const double2 *>(sx + swz(j0 - 1))->y;

      // ---- block scan of one stage + its look-back: returns the true inflow of this thread's chunk
      auto scan = [&](auto kc, const double alpha, const double kbc, const double inflow0, const double (&wl)[CPT]) -> double {
         constexpr int k = decltype(kc)::value; // stage index = look-back chain
         const double a2 = alpha * alpha, a4 = a2 * a2, achunk = (CPT == 8) ? a4 * a4 : a4; // alpha^CPT
         double pw[5];
         double incl = wl[CPT - 1], f = achunk;
#pragma unroll
         for (int s = 0; s < 5; ++s)
         {
            pw[s] = f;
            const double o = __shfl_up_sync(0xffffffffu, incl, 1 << s);
            if (lane >= (1 << s)) incl = fma(f, o, incl);
            f = f * f;
         }
         const double warp_factor = f; // alpha^(32*CPT)
         double excl = __shfl_up_sync(0xffffffffu, incl, 1);
         if (lane == 0) excl = 0.0;
         double lane_pow = 1.0; // alpha^(CPT*lane)
#pragma unroll
         for (int s = 0; s < 5; ++s)
            if ((lane >> s) & 1) lane_pow *= pw[s];
         double afull = warp_factor; // alpha^TILE
#pragma unroll
         for (int w2 = 1; w2 < WARPS; w2 <<= 1) afull *= afull;
         if (lane == 31) s_warp_tot[warp] = incl;
         if (ti.valid != TILE)
         {
            const int last = ti.valid - 1;
            if (tid == last / CPT)
            {
               const int jo = last % CPT;
               double wsel = wl[0], apow = alpha;
#pragma unroll
               for (int jj = 1; jj < CPT; ++jj)
                  if (jj <= jo) { wsel = wl[jj]; apow *= alpha; }
               s_agg = fma(apow, excl, wsel);
               s_alast = apow * lane_pow;
            }
         }
         __syncthreads(); // (C)
         double warp_carry = 0.0, warp_pow = 1.0;
         for (int w2 = 0; w2 < warp; ++w2)
         {
            warp_carry = fma(warp_factor, warp_carry, s_warp_tot[w2]);
            warp_pow *= warp_factor;
         }
         const double E_t = fma(lane_pow, warp_carry, excl);
         const double tpow = lane_pow * warp_pow;
         if (tid == 0)
         {
            double agg, alast;
            if (ti.valid == TILE)
            {
               agg = 0.0;
               for (int w2 = 0; w2 < WARPS; ++w2) agg = fma(warp_factor, agg, s_warp_tot[w2]);
               alast = afull;
            }
            else
            {
               const int wl_ = ((ti.valid - 1) / CPT) >> 5;
               double wc = 0.0, wp = 1.0;
               for (int w2 = 0; w2 < wl_; ++w2) { wc = fma(warp_factor, wc, s_warp_tot[w2]); wp *= warp_factor; }
               agg = fma(s_alast, wc, s_agg);
               alast = s_alast * wp;
            }
            double ctile;
            if (ti.p == 0) ctile = inflow0;
            else
            {
               // always the predecessor's inclusive value of THIS stage (bitwise run-to-run reproducible)
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
            st_status(status + k, fma(alast, ctile, agg), flag_incl);
            s_ctile = ctile;
            if (k == 0) bulk_wait_read<0>(); // the previous tile's store has read S0 / S1 before pass 2 overwrites them
         }
         __syncthreads(); // (E)
         return fma(tpow, s_ctile, E_t);
      };

      double wl[CPT], U[CPT];
      double2 r[CPT];     // result of the stage just finished (own cells)
      double xsum[CPT];   // (u+v) of the cell the next stage's correction source uses

      // ================= stage 0: BE predictor, ends^n -> S0
      {
         const RowCoef c = fold_coef<KIND_BE, CORR>(s_coef[buf][KIND_BE]);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = xr[jj].x + xr[jj].y;
         stage_pass1<KIND_BE, CORR, CPT>(c, xr, xr, xsum, 0.0, 0.0, row_first, wl, U);
         const double cin = scan(std::integral_constant<int, 0>{}, c.alpha, c.kbc, c.inflow0, wl);
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            r[jj] = make_double2(fma(c.gamma, vin, U[jj]), v);
            *reinterpret_cast<double2 *>(t0 + swz(j0 + jj)) = r[jj];
            vin = v;
            ap *= c.alpha;
         }
      }
      // ================= stage 1: CN corrector, x1 = S0 (own cells, registers), x2 = X -> S1
      {
         const RowCoef c = fold_coef<KIND_CN, CORR>(s_coef[buf][KIND_CN]);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
         stage_pass1<KIND_CN, CORR, CPT>(c, r, xr, xsum, 0.0, up_x, row_first, wl, U);
         const double cin = scan(std::integral_constant<int, 1>{}, c.alpha, c.kbc, c.inflow0, wl);
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            r[jj] = make_double2(fma(c.gamma, vin, U[jj]), v);
            *reinterpret_cast<double2 *>(t1 + swz(j0 + jj)) = r[jj];
            vin = v;
            ap *= c.alpha;
         }
      }
      // ================= stage 2: BE predictor, x1 = S1 -> S2 (only u+v survives: the BDF stage's correction source)
      {
         const RowCoef c = fold_coef<KIND_BE, CORR>(s_coef[buf][KIND_BE]);
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
         stage_pass1<KIND_BE, CORR, CPT>(c, r, r, xsum, 0.0, 0.0, row_first, wl, U);
         const double cin = scan(std::integral_constant<int, 2>{}, c.alpha, c.kbc, c.inflow0, wl);
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            xsum[jj] = fma(c.gamma, vin, U[jj]) + v;
            vin = v;
            ap *= c.alpha;
         }
      }
      // ================= stage 3: BDF corrector, x1 = half (S1 for mu<0 rows, S0 for mu>0 rows), x2 = X -> ends^(n+1)
      {
         const RowCoef c = fold_coef<KIND_BDF, CORR>(s_coef[buf][KIND_BDF]);
         const unsigned char *const sh = ti.sgn ? t0 : t1; // written two / four barriers ago
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj) r[jj] = *reinterpret_cast<const double2 *>(sh + swz(j0 + jj));
         double up_h = 0.0;
         if (tid == 0) { if (ti.tile_start > 0) up_h = ti.sgn ? pred_v0 : pred_v1; } // the predecessor tile's last cell (F7)
         else up_h = reinterpret_cast<const double2 *>(sh + swz(j0 - 1))->y;
         stage_pass1<KIND_BDF, CORR, CPT>(c, r, xr, xsum, up_h, up_x, row_first, wl, U);
         const double cin = scan(std::integral_constant<int, 3>{}, c.alpha, c.kbc, c.inflow0, wl);
         unsigned char *const so = ti.sgn ? t1 : t0; // the tile that is NOT half_ends: nobody reads it any more
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            *reinterpret_cast<double2 *>(so + swz(j0 + jj)) = make_double2(fma(c.gamma, vin, U[jj]), v);
            vin = v;
            ap *= c.alpha;
         }
         fence_proxy_async(); // generic-proxy writes -> visible to the TMA store
         __syncthreads();     // (F)
         if (tid == 0)
         {
            tma_store_3d(&maps.m[a.iout], 0, ti.tile_start >> 3, ti.row, so); // rows that end inside the tile are clipped
            bulk_commit();
         }
      }
      cur = nxt;
      buf ^= 1;
      slot = slot_next;
   }
   if (tid == 0)
   {
      bulk_wait_read<0>();
      tq.finish(a, total_tiles);
   }
}

// ------------------------------------------------------------------------------------------ host side
template <bool CORR, int THREADS, int CPT>
static cudaError_t launch_one(const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   int ctas_per_sm = plan_slot->load(std::memory_order_acquire);
   auto kern = step_kernel<CORR, THREADS, CPT>;
   const size_t smem = (size_t)4 * TileShape<THREADS, CPT>::TILE_BYTES + 1024; // 2 X + S0 + S1 + alignment slack
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
   const long long total = (long long)a.n_chains * a.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, THREADS, smem, st>>>(maps, a);
   return cudaGetLastError();
}

bool step_shape_supported(int threads, int cpt) { return (threads == 32 || threads == 64 || threads == 128) && (cpt == 4 || cpt == 8); }

cudaError_t launch_step(bool corr, int threads, int cpt, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms,
                        cudaStream_t st, int *grid_out)
{
   if (threads == 32) return launch_step_warp(corr, cpt, chain_warps, maps, a, num_sms, st, grid_out);
   switch ((threads * 16 + cpt) * 2 + (corr ? 1 : 0))
   {
   case (64 * 16 + 8) * 2 + 0: return launch_one<false, 64, 8>(maps, a, num_sms, st, grid_out);
   case (64 * 16 + 8) * 2 + 1: return launch_one<true, 64, 8>(maps, a, num_sms, st, grid_out);
   case (64 * 16 + 4) * 2 + 0: return launch_one<false, 64, 4>(maps, a, num_sms, st, grid_out);
   case (64 * 16 + 4) * 2 + 1: return launch_one<true, 64, 4>(maps, a, num_sms, st, grid_out);
   case (128 * 16 + 8) * 2 + 0: return launch_one<false, 128, 8>(maps, a, num_sms, st, grid_out);
   case (128 * 16 + 8) * 2 + 1: return launch_one<true, 128, 8>(maps, a, num_sms, st, grid_out);
   case (128 * 16 + 4) * 2 + 0: return launch_one<false, 128, 4>(maps, a, num_sms, st, grid_out);
   case (128 * 16 + 4) * 2 + 1: return launch_one<true, 128, 4>(maps, a, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}
} // namespace rtk
