// sweep_kernel.cu -- the stage sweep (BE / CN / BDF) as a single-pass affine-recurrence scan, fed by TMA.
//
// What the reference does (src/solver.cpp:628-811 with the steppers at 319-587): for every
// direction i and group g it walks the N cells in upwind order and solves a 2x2 system per cell;
// the outflow node of cell j is the inflow of cell j+1 -- a serial first-order recurrence.
//
// What this kernel does: in all three steppers the matrix is [[d, o], [-o, d]] with o = k/2 and k
// the stepper's streaming constant, so with (u,v) = (inflow node, outflow node)
//
//        v_j = alpha * v_{j-1} + beta_j ,   alpha = (o/det) * k ,  det = d^2 + o^2
//        u_j = gamma * v_{j-1} + U_j    ,   gamma = (d/det) * k
//
// where beta_j, U_j depend only on old time levels of cell j (and, for CN/BDF, on old-time-level
// outflow nodes of cell j-1).  alpha is constant along a row (uniform rho, kappa_g, dx).  Rows are
// cut into tiles of TILE cells; a tile is scanned by one CTA (thread-serial chunks of CPT cells,
// warp-shuffle scan of the chunk aggregates, cross-warp combine) and tiles of a row are chained
// with a decoupled look-back through global memory, so every cell is read once and written once.
//
// Data movement (the kernel is HBM-bound and, sustained, POWER-bound: on a 1 kW B200 every instruction
// that is not needed costs bandwidth, so the copies are done by the TMA engine, not by the threads):
//   * every time-level buffer is described by a 3-D tensor map (16 doubles = 8 cells = 128 bytes, Np/8 chunks,
//     rows) with SWIZZLE_128B.  ONE thread issues one cp.async.bulk.tensor load per time level and tile; the
//     bytes land in dense shared memory, XOR-swizzled by the hardware so that the blocked access of the scan
//     (thread t owns cells CPT*t .. CPT*t+CPT-1) is bank-conflict free without padding; completion is signalled
//     on an mbarrier that all threads wait on.  Rows that end inside a tile are clipped / zero-filled by TMA.
//   * the row's stepper constants (144 bytes) and the old-time-level outflow nodes of the cell upstream of the
//     tile travel with the tile as plain bulk copies on the same mbarrier.
//   * results are written back into the staged tile and leave with ONE cp.async.bulk.tensor store per tile.
//   * the tile two tickets ahead is pulled from HBM into L2 with cp.async.bulk.prefetch.L2.
//   Shared memory: a 3-slot ring for the X1/result tile (a slot is reloaded one full iteration after its store was
//   issued, so the store never has to be waited for), 2 slots for the X2 / XC tiles.
//
// Work distribution: persistent CTAs draw tickets from a global counter (several tickets per atomic); ticket t maps
// to (tile position p = t / n_chains, chain = t % n_chains), i.e. all chains advance together and the predecessor
// tile of a chain has normally been finished long before its successor starts -- the look-back then costs one
// 16-byte status read, issued before the local scan so its latency is hidden.  Tickets are handed out and consumed
// in dependency order, so a CTA only ever waits on a CTA that is already running: no deadlock, no cooperative
// launch.  One thread decodes tickets (incrementally -- consecutive tickets need no division).
//
// Reflective left boundary (solver.cpp:677-685): the mu>0 row takes its inflow from the SAME
// stage's result of the mirrored mu<0 row at cell 0, i.e. from the end of that row's sweep.  The two
// rows are simply linked into one chain of 2*tiles_per_row tiles.
//
// Boundary rule common to all BCs (solver.cpp:695-697): at the first cell of a row the three inflow
// carries (n+1, n+1/2, n) all equal bdry_cond.  In scan form: effective inflow = kbc * bdry with
// kbc = 1 (BE), 2 (CN), 6 (BDF), and no old-time-level upstream terms for that cell.
#include <cstdlib>

#include "rt_common.cuh"
#include "sweep_device.cuh"
#include "sweep_kernel.cuh"

namespace rtk
{
__device__ __forceinline__ RowCoef make_coef(int kind, bool corr, const CoefArgs &a, int g, int i)
{
   RowCoef c = coef_from_material(kind, corr, a.mu[i], a.sigma[g], a.srcB[g], corr ? a.cor1[g] : 0.0, corr ? a.cor2[g] : 0.0,
                                  corr ? a.cor3[g] : 0.0, a.dx, a.tau, a.dt, a.beta);
   c.inflow0 = c.kbc * a.bdry[(size_t)i + (size_t)a.M * g];
   return c;
}

// one thread per owned (group, direction): table[(gl*M + i)] for gl in [0, groups)
__global__ void coef_table_kernel(int kind, int corr, CoefArgs a, RowCoef *__restrict__ table)
{
   const int idx = blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= a.groups * a.M) return;
   const int gl = idx / a.M, i = idx % a.M;
   table[idx] = make_coef(kind, corr != 0, a, a.G0 + gl, i);
}

void launch_coef_table(int kind, bool corr, const CoefArgs &a, double *table, cudaStream_t st)
{
   const int n = a.groups * a.M;
   coef_table_kernel<<<(n + 127) / 128, 128, 0, st>>>(kind, corr ? 1 : 0, a, reinterpret_cast<RowCoef *>(table));
}

// RING = number of shared-memory slots for the X1 / result tile:
//   3: a slot is reloaded one full iteration after its store was issued -> the store is never waited for;
//   2: (when 3 would cost a resident CTA) the reload is issued after this tile's look-back, by which time the
//      store of the previous tile has long finished reading its slot.
template <int KIND, bool CORR, int THREADS, int CPT, int RING>
__global__ void __launch_bounds__(THREADS) sweep_kernel(const __grid_constant__ SweepMaps maps, const SweepArgs a)
{
   using TS = TileShape<THREADS, CPT>;
   constexpr int TILE = TS::TILE, TILE_BYTES = TS::TILE_BYTES, WARPS = TS::WARPS;
   constexpr bool NEED_X2 = (KIND != KIND_BE);
   constexpr bool NEED_XC = (KIND == KIND_BDF) && CORR;
   constexpr int NARR = 1 + (NEED_X2 ? 1 : 0) + (NEED_XC ? 1 : 0);
   static_assert(CPT == 8 || CPT == 4, "the swizzled tile image is conflict-free for 4 or 8 cells per thread");
   static_assert((WARPS & (WARPS - 1)) == 0, "WARPS must be a power of two");
   static_assert(TILE_BYTES % 1024 == 0, "SWIZZLE_128B images must be 1024-byte aligned");

   static_assert(RING == 2 || RING == 3, "X1 ring depth");
   // dynamic shared memory, 1024-byte aligned by hand: [RING x X1/result][2 x X2][2 x XC]
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   auto x1_slot = [&](int r) { return tiles + r * TILE_BYTES; };
   auto x2_slot = [&](int b) { return tiles + (RING + b) * TILE_BYTES; };
   auto xc_slot = [&](int b) { return tiles + (RING + (NEED_X2 ? 2 : 0) + b) * TILE_BYTES; };

   __shared__ __align__(16) RowCoef s_coef[2];
   __shared__ __align__(16) double2 s_halo[2][2]; // [buffer][0: X1, 1: X2]: the cell upstream of the tile
   __shared__ __align__(8) unsigned long long s_full[2];
   // tickets and their decoded descriptors rotate through 3 slots: iteration k scans slot k%3, loads slot (k+1)%3 and
   // thread 0 publishes the ticket two ahead into slot (k+2)%3 -- last read two barriers ago, first read after the next
   __shared__ TileDesc s_desc[3];
   __shared__ long long s_ticket[3];
   __shared__ double s_warp_tot[WARPS];
   __shared__ double s_agg, s_alast, s_ctile;

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const unsigned flag_agg = a.epoch * 2u + 1u, flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;

   // ---- thread 0 only: TMA / bulk loads of one tile into pipeline buffer b (X1 into ring slot r)
   auto issue_loads = [&](const TileDesc &ti, int b, int r) {
      const bool halo = NEED_X2 && ti.tile_start > 0;
      const unsigned bytes = (unsigned)(NARR * TILE_BYTES + sizeof(RowCoef)) + (halo ? (KIND == KIND_BDF ? 32u : 16u) : 0u);
      mbar_expect_tx(&s_full[b], bytes);
      const int chunk0 = ti.tile_start >> 3;
      tma_load_3d(x1_slot(r), &maps.m[a.ix1[ti.sgn]], 0, chunk0, ti.row, &s_full[b]);
      if (NEED_X2) tma_load_3d(x2_slot(b), &maps.m[a.ix2], 0, chunk0, ti.row, &s_full[b]);
      if (NEED_XC) tma_load_3d(xc_slot(b), &maps.m[a.ixc], 0, chunk0, ti.row, &s_full[b]);
      bulk_load(&s_coef[b], a.coef + (size_t)ti.row * (sizeof(RowCoef) / 16), (unsigned)sizeof(RowCoef), &s_full[b]);
      if (halo)
      {
         // old-time-level outflow nodes of the cell upstream of the tile (CN: prev ; BDF: half and prev)
         bulk_load(&s_halo[b][1], a.x2 + ti.row_off - 1, 16u, &s_full[b]);
         if (KIND == KIND_BDF) bulk_load(&s_halo[b][0], a.x1[ti.sgn] + ti.row_off - 1, 16u, &s_full[b]);
      }
   };

   // thread 0: pull the time levels of a tile two tickets ahead from HBM into L2 (bulk prefetch, 16-byte granules)
   auto prefetch_tile = [&](const TileDesc &ti) {
      if (!a.l2_prefetch) return;
      const unsigned bytes = (unsigned)ti.valid * (unsigned)sizeof(double2);
      prefetch_l2_bulk(a.x1[ti.sgn] + ti.row_off, bytes);
      if (NEED_X2) prefetch_l2_bulk(a.x2 + ti.row_off, bytes);
      if (NEED_XC) prefetch_l2_bulk(a.xc + ti.row_off, bytes);
   };

   // Tickets are drawn in batches of a.ticket_batch consecutive tickets per atomic (one atomic per tile on a single
   // address would be ~4e8 atomics/s at 6 TB/s with 8 KB tiles).  A batch is consumed in increasing order, so the
   // dependency-order argument above still holds.  Host accounting: the counter advances by one per draw; every CTA
   // makes exactly ONE draw that lies past the end (at the latest when it exits).
   // (no branch depends on a freshly drawn value, so the atomic's latency is only paid where the ticket is used)
   long long last_taken = 0, q_next = 0, q_end = 0, q_start = -1;
   auto take_ticket = [&]() -> long long {
      if (last_taken >= total_tiles) return last_taken;
      if (q_next == q_end)
      {
         q_start = (long long)(atomicAdd(a.ticket, 1ull) - a.ticket_base) * a.ticket_batch;
         q_next = q_start;
         q_end = q_start + a.ticket_batch;
      }
      last_taken = q_next++;
      return last_taken;
   };

   // ---- prologue: barriers armed, two tickets drawn and decoded, the first tile in flight, a third ticket drawn
   TicketCursor cursor;
   cursor.ticket = -2;
   long long ahead = 0; // thread 0: the ticket two tiles ahead (already drawn, value known)
   if (tid == 0)
   {
      mbar_init(&s_full[0], 1);
      mbar_init(&s_full[1], 1);
      fence_proxy_async();
      const long long t0 = take_ticket(), t1 = take_ticket();
      s_ticket[0] = t0;
      s_ticket[1] = t1;
      if (t0 < total_tiles)
      {
         s_desc[0] = decode_ticket(a, cursor, t0, TILE);
         issue_loads(s_desc[0], 0, 0);
      }
      if (t1 < total_tiles) s_desc[1] = decode_ticket(a, cursor, t1, TILE);
      ahead = take_ticket();
   }
   __syncthreads();
   long long cur = s_ticket[0];
   int buf = 0, ring = 0, slot = 0;
   unsigned phase = 0; // bit b: parity the next wait on s_full[b] uses

   while (cur < total_tiles)
   {
      const int ring_next = (ring == RING - 1) ? 0 : ring + 1;
      const int slot_next = (slot == 2) ? 0 : slot + 1, slot_ahead = (slot_next == 2) ? 0 : slot_next + 1;
      const long long nxt = s_ticket[slot_next];
      // thread 0: next tile's loads.  The target X1 slot must no longer be read by the store that last used it.
      auto issue_next = [&]() {
         if (nxt >= total_tiles) return;
         if (RING == 3) bulk_wait_read<1>(); // that store was issued a full iteration ago; the newest may still fly
         else bulk_wait_read<0>();
         const TileDesc dn = s_desc[slot_next];
         issue_loads(dn, buf ^ 1, ring_next);
      };
      if (tid == 0)
      {
         if (RING == 3) issue_next();
         s_ticket[slot_ahead] = ahead;
         if (ahead < total_tiles)
         {
            const TileDesc d_ahead = decode_ticket(a, cursor, ahead, TILE);
            prefetch_tile(d_ahead); // HBM -> L2 for the tile after next
            s_desc[slot_ahead] = d_ahead;
         }
         ahead = take_ticket(); // three ahead; the atomic's latency hides behind this tile's work
      }

      const TileDesc ti = s_desc[slot];
      unsigned char *const s1 = x1_slot(ring); // X1 tile, later the result tile
      unsigned char *const s2 = x2_slot(buf);  // X2 tile
      unsigned char *const sc = xc_slot(buf);  // ends-before-stage tile for the BDF correction source
      double2 *status = a.tile_status + ti.status_idx;

      // early look-back probe: by the time the local scan is done the predecessor's word has arrived
      double2 early = make_double2(0.0, 0.0);
      if (tid == 0 && ti.p > 0) early = ld_status(status - 1);

      mbar_wait(&s_full[buf], (phase >> buf) & 1u); // this tile's bytes have landed
      phase ^= 1u << buf;

      // the row's constants from shared memory; entries that are structurally fixed for this stage kind are folded
      // at compile time (BE: no cross terms, no X2 ; CN: no X2 cell terms ; symmetric / antisymmetric pairs)
      RowCoef c;
      {
         const RowCoef &cs = s_coef[buf];
         c.alpha = cs.alpha; c.gamma = cs.gamma; c.id = cs.id; c.io = cs.io;
         c.qa = cs.qa;
         c.qb = CORR ? cs.qb : 0.0;
         c.c_uu = cs.c_uu; c.c_vv = c.c_uu;
         c.c_uv = (KIND == KIND_BE) ? 0.0 : cs.c_uv; c.c_vu = -c.c_uv;
         c.p_uu = (KIND == KIND_BDF) ? cs.p_uu : 0.0; c.p_vv = c.p_uu;
         c.p_uv = (KIND == KIND_BDF) ? cs.p_uv : 0.0; c.p_vu = -c.p_uv;
         c.up1 = (KIND == KIND_BDF) ? cs.up1 : 0.0;
         c.up2 = (KIND != KIND_BE) ? cs.up2 : 0.0;
         c.kbc = (KIND == KIND_BE) ? 1.0 : (KIND == KIND_CN ? 2.0 : 6.0);
         c.inflow0 = cs.inflow0;
      }

      // ---- pass 1: per-cell beta_j / U_j and the thread-local scan with zero inflow
      double wl[CPT], U[CPT];
      {
         const int j0 = tid * CPT;
         double up1 = 0.0, up2 = 0.0; // X1(j-1).v and X2(j-1).v
         if (tid == 0)
         {
            if (NEED_X2 && ti.tile_start > 0) { up1 = s_halo[buf][0].y; up2 = s_halo[buf][1].y; }
         }
         else
         {
            if (KIND == KIND_BDF) up1 = reinterpret_cast<const double2 *>(s1 + swz(j0 - 1))->y;
            if (NEED_X2) up2 = reinterpret_cast<const double2 *>(s2 + swz(j0 - 1))->y;
         }
         const bool row_first = (ti.tile_start == 0) && (tid == 0);
         double w = 0.0;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const int o = swz(j0 + jj);
            const double2 x1 = *reinterpret_cast<const double2 *>(s1 + o);
            double2 x2 = make_double2(0.0, 0.0);
            if (NEED_X2) x2 = *reinterpret_cast<const double2 *>(s2 + o);
            double q = c.qa;
            if (CORR)
            {
               const double2 xs = NEED_XC ? *reinterpret_cast<const double2 *>(sc + o) : x1;
               q = fma(c.qb, xs.x + xs.y, c.qa);
            }
            double ru, rv;
            if (KIND == KIND_BE)
            {
               ru = fma(c.c_uu, x1.x, q);
               rv = fma(c.c_uu, x1.y, q);
            }
            else
            {
               ru = q + c.c_uu * x1.x + c.c_uv * x1.y;
               rv = q + c.c_vu * x1.x + c.c_vv * x1.y;
            }
            if (KIND == KIND_BDF)
            {
               ru += c.p_uu * x2.x + c.p_uv * x2.y;
               rv += c.p_vu * x2.x + c.p_vv * x2.y;
            }
            if (NEED_X2 && !(row_first && jj == 0))
            {
               if (KIND == KIND_BDF) ru += c.up1 * up1 + c.up2 * up2;
               else ru += c.up2 * up2;
            }
            const double beta = c.io * ru + c.id * rv;
            U[jj] = c.id * ru - c.io * rv;
            w = fma(c.alpha, w, beta);
            wl[jj] = w;
            if (KIND == KIND_BDF) up1 = x1.y;
            if (NEED_X2) up2 = x2.y;
         }
      }

      // ---- block scan of the chunk aggregates (factor alpha^CPT per chunk)
      const double a2 = c.alpha * c.alpha, a4 = a2 * a2, achunk = (CPT == 8) ? a4 * a4 : a4; // alpha^CPT
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
         // row ends inside the tile: zero-inflow value and inflow factor at the last valid cell need the warp carry,
         // so the owning thread publishes its pieces and thread 0 finishes the job after the barrier
         const int last = ti.valid - 1;
         if (tid == last / CPT)
         {
            const int jo = last % CPT;
            double wsel = wl[0], apow = c.alpha;
#pragma unroll
            for (int jj = 1; jj < CPT; ++jj)
               if (jj <= jo) { wsel = wl[jj]; apow *= c.alpha; }
            // value(last) = apow * (lane_pow * warp_carry + excl) + wsel ; factor(last) = apow * lane_pow * warp_pow
            s_agg = fma(apow, excl, wsel);
            s_alast = apow * lane_pow;
         }
      }
      __syncthreads(); // (C) warp totals (and the partial-tile pieces) visible
      double warp_carry = 0.0, warp_pow = 1.0; // inflow of this warp for zero tile inflow ; alpha^(32*CPT*warp)
      for (int w2 = 0; w2 < warp; ++w2)
      {
         warp_carry = fma(warp_factor, warp_carry, s_warp_tot[w2]);
         warp_pow *= warp_factor;
      }
      const double E_t = fma(lane_pow, warp_carry, excl); // inflow of this thread's chunk for zero tile inflow
      const double tpow = lane_pow * warp_pow;            // alpha^(CPT*tid)

      // ---- decoupled look-back (one thread)
      if (tid == 0)
      {
         // tile aggregate: zero-inflow value and inflow factor at the tile's last valid cell
         double agg, alast;
         if (ti.valid == TILE)
         {
            agg = 0.0;
            for (int w2 = 0; w2 < WARPS; ++w2) agg = fma(warp_factor, agg, s_warp_tot[w2]);
            alast = afull;
         }
         else
         {
            const int wl_ = ((ti.valid - 1) / CPT) >> 5; // warp that owns the last valid cell
            double wc = 0.0, wp = 1.0;
            for (int w2 = 0; w2 < wl_; ++w2) { wc = fma(warp_factor, wc, s_warp_tot[w2]); wp *= warp_factor; }
            agg = fma(s_alast, wc, s_agg);
            alast = s_alast * wp;
         }
         double ctile;
         if (ti.p == 0) ctile = c.inflow0;
         else
         {
            double2 w = early;
            if (ti.junction || !a.lookback_walk)
            {
               // deterministic: always the predecessor's inclusive value (bitwise run-to-run reproducible)
               long long spins = 0;
               while ((int)(status_flag(w) - flag_incl) < 0)
               {
                  __nanosleep(32);
                  w = ld_status(status - 1);
                  if (++spins > SPIN_LIMIT) __trap();
               }
               ctile = ti.junction ? c.kbc * w.x : w.x;
            }
            else if ((int)(status_flag(w) - flag_incl) >= 0) ctile = w.x;
            else
            {
               // classic walk-back over aggregates (not bitwise reproducible; for very few chains)
               st_status(status, agg, flag_agg);
               double carry = 0.0, fac = 1.0;
               const double2 *qp = status - 1;
               while (true)
               {
                  w = ld_status(qp);
                  long long spins = 0;
                  while ((int)(status_flag(w) - flag_agg) < 0)
                  {
                     __nanosleep(32);
                     w = ld_status(qp);
                     if (++spins > SPIN_LIMIT) __trap();
                  }
                  if ((int)(status_flag(w) - flag_incl) >= 0) { carry = fma(fac, w.x, carry); break; }
                  carry = fma(fac, w.x, carry);
                  fac *= afull;
                  if (fac == 0.0) break; // everything further upstream has underflowed out of fp64 range
                  --qp; // never runs past the junction / chain head: those tiles always publish inclusive values
               }
               ctile = carry;
            }
         }
         st_status(status, fma(alast, ctile, agg), flag_incl);
         s_ctile = ctile;
         if (RING == 2) issue_next();
      }
      __syncthreads(); // (E)

      // ---- pass 2: true nodes, written back into the X1 tile image
      {
         const double cin = fma(tpow, s_ctile, E_t); // true inflow of this thread's chunk
         const int j0 = tid * CPT;
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            const double u = fma(c.gamma, vin, U[jj]);
            *reinterpret_cast<double2 *>(s1 + swz(j0 + jj)) = make_double2(u, v);
            vin = v;
            ap *= c.alpha;
         }
      }
      fence_proxy_async(); // generic-proxy writes -> visible to the TMA store
      __syncthreads();     // (F)

      if (tid == 0)
      {
         tma_store_3d(&maps.m[a.iout], 0, ti.tile_start >> 3, ti.row, s1); // rows that end inside the tile are clipped
         bulk_commit();
      }
      cur = nxt;
      buf ^= 1;
      ring = ring_next;
      slot = slot_next;
   }
   if (tid == 0)
   {
      bulk_wait_read<0>(); // the last stores must have read their tiles before the CTA's shared memory goes away
      if (q_start < total_tiles) atomicAdd(a.ticket, 1ull); // no draw of this CTA lay past the end yet (q_start = -1: none at all)
   }
}

// ------------------------------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 3-D view of one time-level buffer [rows][Np] of double2: (16 doubles = 8 cells, Np/8 chunks, rows), box = one tile
cudaError_t make_tile_map(CUtensorMap *map, void *base, int rows, int Np, int tile_cells)
{
   static EncodeTiledFn encode = nullptr;
   if (!encode)
   {
      void *fn = nullptr;
      cudaDriverEntryPointQueryResult q;
      cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
      if (e != cudaSuccess) return e;
      if (q != cudaDriverEntryPointSuccess || !fn) return cudaErrorNotSupported;
      encode = (EncodeTiledFn)fn;
   }
   const cuuint64_t dims[3] = {16, (cuuint64_t)(Np / 8), (cuuint64_t)rows};
   const cuuint64_t strides[2] = {128, (cuuint64_t)Np * 16}; // bytes, dims 1 and 2
   const cuuint32_t box[3] = {16, (cuuint32_t)(tile_cells / 8), 1};
   const cuuint32_t estr[3] = {1, 1, 1};
   const CUresult r = encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
   return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

template <int KIND, bool CORR, int THREADS, int CPT, int RING>
static size_t smem_bytes()
{
   const int narr = 1 + (KIND != KIND_BE ? 1 : 0) + ((KIND == KIND_BDF && CORR) ? 1 : 0);
   return (size_t)(RING + 2 * (narr - 1)) * TileShape<THREADS, CPT>::TILE_BYTES + 1024; // + alignment slack
}

template <int KIND, bool CORR, int THREADS, int CPT>
static cudaError_t launch_one(int ring_pref, const SweepMaps &maps, const SweepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   // plan once per instantiation: X1 ring depth and resident CTAs per SM
   static DevicePlanCache plans; // per device: ctas_per_sm * 8 + ring
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   const int packed = plan_slot->load(std::memory_order_acquire);
   int ctas_per_sm = packed >> 3, ring = packed & 7;
   auto k3 = sweep_kernel<KIND, CORR, THREADS, CPT, 3>;
   auto k2 = sweep_kernel<KIND, CORR, THREADS, CPT, 2>;
   const size_t s3 = smem_bytes<KIND, CORR, THREADS, CPT, 3>(), s2 = smem_bytes<KIND, CORR, THREADS, CPT, 2>();
   if (packed <= 0)
   {
      cudaError_t e = cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s3);
      if (e != cudaSuccess) return e;
      e = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s2);
      if (e != cudaSuccess) return e;
      int o3 = 0, o2 = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o3, k3, THREADS, s3);
      if (e != cudaSuccess) return e;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o2, k2, THREADS, s2);
      if (e != cudaSuccess) return e;
      // measured (exp9): with >= 4 resident CTAs the never-wait 3-slot ring wins (CN: 5 CTAs x ring 3 beats 6 x ring 2);
      // below that the extra CTA is worth more (BDF: 4 CTAs x ring 2 beats 3 x ring 3)
      ring = (o3 < 4 && o2 > o3) ? 2 : 3;
      if (ring_pref == 2 || ring_pref == 3) ring = ring_pref;
      ctas_per_sm = (ring == 2) ? o2 : o3;
      if (ctas_per_sm < 1) ctas_per_sm = 1;
      plan_slot->store(ctas_per_sm * 8 + ring, std::memory_order_release);
   }
   const long long total = (long long)a.n_chains * a.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   if (ring == 3) k3<<<(unsigned)grid, THREADS, s3, st>>>(maps, a);
   else k2<<<(unsigned)grid, THREADS, s2, st>>>(maps, a);
   return cudaGetLastError();
}

template <int T, int CPT>
static cudaError_t launch_shape(int kind, bool corr, int ring, const SweepMaps &maps, const SweepArgs &a, int num_sms,
                                cudaStream_t st, int *grid_out)
{
   switch (kind * 2 + (corr ? 1 : 0))
   {
   case 0: return launch_one<KIND_BE, false, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   case 1: return launch_one<KIND_BE, true, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   case 2: return launch_one<KIND_CN, false, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   case 3: return launch_one<KIND_CN, true, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   case 4: return launch_one<KIND_BDF, false, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   case 5: return launch_one<KIND_BDF, true, T, CPT>(ring, maps, a, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}

bool sweep_shape_supported(int threads, int cpt)
{
   return (threads == 64 || threads == 128 || threads == 256) && (cpt == 4 || cpt == 8);
}

cudaError_t launch_sweep(int kind, bool corr, int threads, int cpt, int ring, const SweepMaps &maps, const SweepArgs &a,
                         int num_sms, cudaStream_t st, int *grid_out)
{
   switch (threads * 16 + cpt)
   {
   case 64 * 16 + 8: return launch_shape<64, 8>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   case 64 * 16 + 4: return launch_shape<64, 4>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   case 128 * 16 + 8: return launch_shape<128, 8>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   case 128 * 16 + 4: return launch_shape<128, 4>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   case 256 * 16 + 8: return launch_shape<256, 8>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   case 256 * 16 + 4: return launch_shape<256, 4>(kind, corr, ring, maps, a, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}
} // namespace rtk
