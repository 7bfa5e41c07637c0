// sweep_kernel.cu -- the stage sweep (BE / CN / BDF) as a single-pass affine-recurrence scan.
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
// cut into tiles of TILE cells; a tile is scanned by one CTA (thread-serial chunks of 8 cells,
// warp-shuffle scan of the chunk aggregates, cross-warp combine) and tiles of a row are chained
// with a decoupled look-back through global memory, so every cell is read once and written once.
//
// Work distribution: persistent CTAs draw tickets from a global counter; ticket t maps to
// (tile position p = t / n_chains, chain = t % n_chains), i.e. all chains advance together and the
// predecessor tile of a chain has normally been finished long before its successor starts -- the
// look-back then costs one 16-byte status read, issued before the local scan so its latency is hidden.  With few chains (small problems) the same code degrades
// gracefully to the classic aggregate walk-back.  Tickets are handed out in dependency order, so a
// CTA only ever waits on a CTA that is already running: no deadlock, no cooperative launch.
//
// Pipeline: each CTA double-buffers its tiles in shared memory -- the cp.async copies of the NEXT ticket's tile are
// in flight while the current tile is scanned, and the ticket after that is being drawn -- so HBM reads never wait
// for the scan, the look-back or the barriers of the tile in progress.
//
// Reflective left boundary (solver.cpp:677-685): the mu>0 row takes its inflow from the SAME
// stage's result of the mirrored mu<0 row at cell 0, i.e. from the end of that row's sweep.  The two
// rows are simply linked into one chain of 2*tiles_per_row tiles.
//
// Boundary rule common to all BCs (solver.cpp:695-697): at the first cell of a row the three inflow
// carries (n+1, n+1/2, n) all equal bdry_cond.  In scan form: effective inflow = kbc * bdry with
// kbc = 1 (BE), 2 (CN), 6 (BDF), and no old-time-level upstream terms for that cell.
#include "rt_common.cuh"
#include "sweep_kernel.cuh"

namespace rtk
{
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem)
{
   const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
   asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}

// padded shared-memory slot of cell j of the tile: one 16-byte pad slot per 8 cells makes both the
// striped (coalesced) and the blocked (8 consecutive cells per thread) access patterns conflict-free
__device__ __forceinline__ int slot(int j) { return j + (j >> 3); }

template <int THREADS, int CPT_>
struct TileShape
{
   static constexpr int CPT = CPT_; // cells per thread: 8 or 4 (4 doubles the warps per shared-memory byte)
   static constexpr int TILE = THREADS * CPT;
   static constexpr int SLOTS = TILE + TILE / 8;
   static constexpr int WARPS = THREADS / 32;
};

// Per-row stepper constants (identical formulas for both signs of mu, see rt_common.cuh).  They depend on
// (stage kind, group, direction) only, so they are evaluated ONCE per solver by coef_table_kernel into a table of
// 144-byte entries; the sweep kernel stages the entry of a tile's row in shared memory with the tile itself
// (one cp.async group) -- no divisions, no dependent global loads of cross sections on the per-tile critical path.
struct RowCoef
{
   double alpha, gamma, id, io; // recurrence factors and d/det, o/det
   double qa, qb;               // q = qa + qb*(u+v) of the cell the source is evaluated at
   double c_uu, c_uv, c_vu, c_vv;     // weights of X1 (ends or half_ends) in (ru, rv)
   double p_uu, p_uv, p_vu, p_vv;     // weights of X2 (prev_ends) in (ru, rv)   [BDF]
   double up1, up2;                   // weights of the upstream old-level outflow nodes X1(j-1).v, X2(j-1).v in ru
   double kbc;                        // boundary multiplier
   double inflow0;                    // kbc * bdry_cond of this (direction, group): inflow of the row's first cell
};
static_assert(sizeof(RowCoef) == COEF_ENTRY_DOUBLES * sizeof(double), "coefficient table entry size");
static_assert(sizeof(RowCoef) % 16 == 0, "entries are staged with 16-byte cp.async copies");

__device__ __forceinline__ RowCoef make_coef(int kind, bool corr, const CoefArgs &a, int g, int i)
{
   RowCoef c;
   const double mu = a.mu[i], amu = fabs(mu), sig = a.sigma[g];
   const double ctau = C_LIGHT * a.tau;
   const double hq = 0.5 * ctau * a.dx; // 0.5*c*timestep*dx (solver.cpp:337)
   c.qa = hq * a.srcB[g];
   c.qb = 0.0;
   if (corr)
   {
      // total_correction = (cor1*psi + cor2)*mu*beta - cor3*mu^2*beta^2 (correction.cpp:392-394), psi = (u+v)/2
      const double mb = mu * a.beta;
      c.qa += hq * (a.cor2[g] * mb - a.cor3[g] * mb * mb);
      c.qb = 0.5 * hq * a.cor1[g] * mb;
   }
   double k, d;
   c.p_uu = c.p_uv = c.p_vu = c.p_vv = 0.0;
   c.up1 = c.up2 = 0.0;
   if (kind == KIND_BE)
   {
      // solver.cpp:324-334,344-345,381-382
      k = ctau * amu;
      d = 0.5 * ((1.0 + ctau * sig) * a.dx + k);
      c.c_uu = 0.5 * a.dx; c.c_uv = 0.0; c.c_vu = 0.0; c.c_vv = 0.5 * a.dx;
      c.kbc = 1.0;
   }
   else if (kind == KIND_CN)
   {
      // solver.cpp:411-425,435-436,469-470
      const double t = 0.5 * ctau * sig;
      k = 0.5 * C_LIGHT * amu * a.tau;
      d = 0.5 * ((1.0 + t) * a.dx + k);
      const double e1 = 0.5 * ((1.0 - t) * a.dx - k);
      c.c_uu = e1; c.c_uv = -0.5 * k; c.c_vu = 0.5 * k; c.c_vv = e1;
      c.up2 = k; // a * in_prev, in_prev = prev_ends(j-1).v
      c.kbc = 2.0;
   }
   else
   {
      // solver.cpp:498-512,522-527,561-566 ; streaming constant uses the FULL dt
      const double t = C_LIGHT * sig * a.tau / 6.0;
      k = C_LIGHT * amu * a.dt / 6.0;
      d = 0.5 * ((1.0 + t) * a.dx + k);
      const double f1 = 0.5 * ((1.0 - 4.0 * t) * a.dx - 4.0 * k), g1 = 0.5 * (k + t * a.dx);
      c.c_uu = f1; c.c_uv = -2.0 * k; c.c_vu = 2.0 * k; c.c_vv = f1;
      c.p_uu = -g1; c.p_uv = -0.5 * k; c.p_vu = 0.5 * k; c.p_vv = -g1;
      c.up1 = 4.0 * k; // b * 4 * in_half, in_half = half_ends(j-1).v
      c.up2 = k;       // b * in_prev
      c.kbc = 6.0;
   }
   const double o = 0.5 * k, det = d * d + o * o;
   c.id = d / det;
   c.io = o / det;
   c.alpha = c.io * k;
   c.gamma = c.id * k;
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

// ---- tile status word: {value, flag} in ONE aligned 16-byte word, written and read with single 128-bit
// accesses (the same single-copy-atomicity CUB's decoupled look-back relies on for 16-byte descriptors), so no
// fence is needed between value and flag.
__device__ __forceinline__ double2 ld_status(const double2 *p)
{
   double2 v;
   asm volatile("ld.volatile.global.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
   return v;
}
__device__ __forceinline__ void st_status(double2 *p, double value, unsigned flag)
{
   const double f = __longlong_as_double((long long)flag);
   asm volatile("st.volatile.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(value), "d"(f) : "memory");
}
__device__ __forceinline__ unsigned status_flag(const double2 &v) { return (unsigned)__double_as_longlong(v.y); }

__device__ __forceinline__ void cp_async8(void *smem, const void *gmem)
{
   const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
   asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
}
// HBM -> L2 only (no shared-memory landing zone needed): used two tiles ahead of the tile being scanned
__device__ __forceinline__ void prefetch_l2_bulk(const void *gmem, unsigned bytes)
{
   asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Everything the CTA needs to know about one ticket; decoded by ONE thread (the 64-bit divisions are not cheap) and
// handed to the others through shared memory.
struct TileDesc
{
   long long row_off;    // element offset of the tile's first cell in a time-level buffer
   long long status_idx; // index of the tile's look-back word
   int row, p, t, sgn, tile_start, valid;
};

// ticket -> (tile position p, chain) -> (row, tile in row); p-major so that all chains advance together
__device__ __forceinline__ TileDesc decode_ticket(const SweepArgs &a, long long ticket, int tile_cells)
{
   TileDesc ti;
   ti.p = (int)(ticket / a.n_chains);
   const int chain = (int)(ticket % a.n_chains);
   if (a.reflective_left)
   {
      const int half_m = a.M >> 1, gl = chain / half_m, pp = chain % half_m;
      if (ti.p < a.tiles_per_row) { ti.row = gl * a.M + (half_m - 1 - pp); ti.t = ti.p; }
      else { ti.row = gl * a.M + (half_m + pp); ti.t = ti.p - a.tiles_per_row; }
   }
   else { ti.row = chain; ti.t = ti.p; }
   ti.sgn = ((ti.row % a.M) >= (a.M >> 1)) ? 1 : 0;
   ti.tile_start = ti.t * tile_cells;
   ti.valid = min(tile_cells, a.N - ti.tile_start);
   ti.row_off = (long long)ti.row * a.Np + ti.tile_start;
   ti.status_idx = (long long)chain * a.tiles_per_chain + ti.p;
   return ti;
}

// (A single-buffered variant that relied on the L2 prefetch alone and doubled the resident CTAs was measured
// 7 % slower on C4 -- gpurun_out exp2 -- and removed.)
template <int KIND, bool CORR, int THREADS, int CPT>
__global__ void __launch_bounds__(THREADS) sweep_kernel(const SweepArgs a)
{
   using TS = TileShape<THREADS, CPT>;
   constexpr int TILE = TS::TILE, SLOTS = TS::SLOTS, WARPS = TS::WARPS;
   constexpr bool NEED_X2 = (KIND != KIND_BE);
   constexpr bool NEED_XC = (KIND == KIND_BDF) && CORR;
   constexpr int NARR = 1 + (NEED_X2 ? 1 : 0) + (NEED_XC ? 1 : 0);
   static_assert(CPT == 8 || CPT == 4, "the padded shared-memory layout is conflict-free for 4 or 8 cells per thread");
   static_assert((WARPS & (WARPS - 1)) == 0, "WARPS must be a power of two");

   extern __shared__ double2 smem[]; // [2 pipeline buffers][NARR arrays][SLOTS]
   __shared__ double s_warp_tot[WARPS];
   __shared__ double s_halo[2][2];
   __shared__ double s_agg, s_alast, s_ctile;
   __shared__ long long s_ticket[2];
   __shared__ TileDesc s_desc[3]; // [0],[1]: the tiles in the two pipeline buffers ; [2]: the tile two tickets ahead
   __shared__ __align__(16) RowCoef s_coef[2];

   const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
   const unsigned flag_agg = a.epoch * 2u + 1u, flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;

   // coalesced 16-byte async copies of one tile's time levels into pipeline buffer b
   auto issue_loads = [&](const TileDesc &ti, int b) {
      double2 *d1 = smem + (size_t)b * NARR * SLOTS;
      double2 *d2 = d1 + SLOTS;
      double2 *dc = d1 + (NEED_X2 ? 2 : 1) * SLOTS;
      const double2 *g1 = a.x1[ti.sgn] + ti.row_off;
      const double2 *g2 = NEED_X2 ? a.x2[ti.sgn] + ti.row_off : nullptr;
      const double2 *gc = NEED_XC ? a.xc[ti.sgn] + ti.row_off : nullptr;
#pragma unroll
      for (int k = 0; k < CPT; ++k)
      {
         const int j = k * THREADS + tid, s = slot(j);
         if (j < ti.valid)
         {
            cp_async16(&d1[s], g1 + j);
            if (NEED_X2) cp_async16(&d2[s], g2 + j);
            if (NEED_XC) cp_async16(&dc[s], gc + j);
         }
         else
         {
            d1[s] = make_double2(0.0, 0.0);
            if (NEED_X2) d2[s] = make_double2(0.0, 0.0);
            if (NEED_XC) dc[s] = make_double2(0.0, 0.0);
         }
      }
      // the row's stepper constants travel with the tile (9 x 16 bytes)
      if (tid < (int)(sizeof(RowCoef) / 16))
         cp_async16(reinterpret_cast<double2 *>(&s_coef[b]) + tid, a.coef + (size_t)ti.row * (sizeof(RowCoef) / 16) + tid);
      if (NEED_X2 && tid == 0)
      {
         // old-time-level outflow nodes of the cell upstream of the tile (CN: prev ; BDF: half and prev)
         if (ti.tile_start > 0)
         {
            cp_async8(&s_halo[b][1], &(g2 - 1)->y);
            if (KIND == KIND_BDF) cp_async8(&s_halo[b][0], &(g1 - 1)->y);
            else s_halo[b][0] = 0.0;
         }
         else { s_halo[b][0] = 0.0; s_halo[b][1] = 0.0; }
      }
   };

   // one thread: pull the time levels of a tile two tickets ahead from HBM into L2 (bulk prefetch, 16-byte granules)
   auto prefetch_tile = [&](const TileDesc &ti) {
      if (!a.l2_prefetch) return;
      const unsigned bytes = (unsigned)ti.valid * (unsigned)sizeof(double2);
      prefetch_l2_bulk(a.x1[ti.sgn] + ti.row_off, bytes);
      if (NEED_X2) prefetch_l2_bulk(a.x2[ti.sgn] + ti.row_off, bytes);
      if (NEED_XC) prefetch_l2_bulk(a.xc[ti.sgn] + ti.row_off, bytes);
   };

   // every CTA draws tickets until it has seen ONE that is past the end (the host's ticket accounting relies on it)
   // Tickets are drawn in batches of a.ticket_batch consecutive tickets per atomic (one atomic per tile on a single
   // address would be ~4e8 atomics/s at 6 TB/s with 8 KB tiles).  A batch is consumed in increasing order, so the
   // dependency-order argument above still holds.  Host accounting: the counter advances by one per draw; every CTA
   // makes exactly ONE draw that lies past the end (at the latest when it exits).
   long long last_taken = 0, q_next = 0, q_end = 0;
   bool drew_past_end = false;
   auto take_ticket = [&]() -> long long {
      if (last_taken >= total_tiles) return last_taken;
      if (q_next == q_end)
      {
         q_next = (long long)(atomicAdd(a.ticket, 1ull) - a.ticket_base) * a.ticket_batch;
         q_end = q_next + a.ticket_batch;
         if (q_next >= total_tiles) drew_past_end = true;
      }
      last_taken = q_next++;
      return last_taken;
   };

   // ---- prologue: two tickets drawn and decoded, the first tile in flight, a third ticket drawn
   long long ahead = 0; // thread 0: the ticket two tiles ahead (already drawn, value known)
   if (tid == 0)
   {
      const long long t0 = take_ticket(), t1 = take_ticket();
      s_ticket[0] = t0;
      s_ticket[1] = t1;
      if (t0 < total_tiles) s_desc[0] = decode_ticket(a, t0, TILE);
      if (t1 < total_tiles) s_desc[1] = decode_ticket(a, t1, TILE);
      ahead = take_ticket();
   }
   __syncthreads();
   long long cur = s_ticket[0];
   int buf = 0;
   if (cur < total_tiles) issue_loads(s_desc[0], 0);
   cp_async_commit();

   while (cur < total_tiles)
   {
      __syncthreads(); // (A) next ticket visible ; everybody is done with the other pipeline buffer
      const long long nxt = s_ticket[buf ^ 1];
      if (nxt < total_tiles) issue_loads(s_desc[buf ^ 1], buf ^ 1);
      cp_async_commit();
      long long newer = 0;
      if (tid == 0)
      {
         if (ahead < total_tiles)
         {
            const TileDesc d_ahead = decode_ticket(a, ahead, TILE);
            prefetch_tile(d_ahead); // HBM -> L2 for the tile after next
            s_desc[2] = d_ahead;
         }
         newer = take_ticket(); // three ahead; latency hidden behind this tile's work, consumed at the loop end
      }

      const TileDesc ti = s_desc[buf];
      double2 *s1 = smem + (size_t)buf * NARR * SLOTS; // X1 tile, later the result tile
      double2 *s2 = s1 + SLOTS;                          // X2 tile
      double2 *sc = s1 + (NEED_X2 ? 2 : 1) * SLOTS;      // ends-before-stage tile for the BDF correction source
      double2 *status = a.tile_status + ti.status_idx;

      // early look-back probe: by the time the local scan is done the predecessor's word has arrived
      double2 early = make_double2(0.0, 0.0);
      if (tid == 0 && ti.p > 0) early = ld_status(status - 1);

      cp_async_wait<1>(); // this tile's copies (the group before the prefetch) have landed
      __syncthreads();    // (B)
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
         double up1, up2; // X1(j-1).v and X2(j-1).v
         if (tid == 0) { up1 = s_halo[buf][0]; up2 = s_halo[buf][1]; }
         else
         {
            up1 = (KIND == KIND_BDF) ? s1[slot(j0 - 1)].y : 0.0;
            up2 = NEED_X2 ? s2[slot(j0 - 1)].y : 0.0;
         }
         const bool row_first = (ti.tile_start == 0) && (tid == 0);
         double w = 0.0;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const int s = slot(j0 + jj);
            const double2 x1 = s1[s];
            double2 x2 = make_double2(0.0, 0.0);
            if (NEED_X2) x2 = s2[s];
            double q = c.qa;
            if (CORR)
            {
               const double2 xs = NEED_XC ? sc[s] : x1;
               q = fma(c.qb, xs.x + xs.y, c.qa);
            }
            double ru = q + c.c_uu * x1.x + c.c_uv * x1.y;
            double rv = q + c.c_vu * x1.x + c.c_vv * x1.y;
            if (KIND == KIND_BDF)
            {
               ru += c.p_uu * x2.x + c.p_uv * x2.y;
               rv += c.p_vu * x2.x + c.p_vv * x2.y;
            }
            if (NEED_X2 && !(row_first && jj == 0)) ru += c.up1 * up1 + c.up2 * up2;
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
      double warp_carry = 0.0, warp_pow = 1.0; // inflow of this warp for zero tile inflow ; alpha^(32*CPT*warp)
      if (WARPS > 1)
      {
         if (lane == 31) s_warp_tot[warp] = incl;
         __syncthreads(); // (C)
         for (int w2 = 0; w2 < warp; ++w2)
         {
            warp_carry = fma(warp_factor, warp_carry, s_warp_tot[w2]);
            warp_pow *= warp_factor;
         }
      }
      const double E_t = fma(lane_pow, warp_carry, excl); // inflow of this thread's chunk for zero tile inflow
      const double tpow = lane_pow * warp_pow;            // alpha^(CPT*tid)
      double afull = warp_factor;                         // alpha^TILE
#pragma unroll
      for (int w2 = 1; w2 < WARPS; w2 <<= 1) afull *= afull;

      // ---- tile aggregate (zero-inflow value and inflow factor at the tile's last valid cell)
      if (ti.valid == TILE)
      {
         if (tid == THREADS - 1) { s_agg = fma(achunk, E_t, wl[CPT - 1]); s_alast = afull; }
      }
      else
      {
         const int last = ti.valid - 1;
         if (tid == last / CPT)
         {
            const int jo = last % CPT;
            double wsel = wl[0], apow = c.alpha;
#pragma unroll
            for (int jj = 1; jj < CPT; ++jj)
               if (jj <= jo) { wsel = wl[jj]; apow *= c.alpha; }
            s_agg = fma(apow, E_t, wsel);
            s_alast = apow * tpow;
         }
      }
      __syncthreads(); // (D)

      // ---- decoupled look-back (one thread)
      if (tid == 0)
      {
         double ctile;
         if (ti.p == 0) ctile = c.inflow0;
         else
         {
            const bool junction = (ti.t == 0); // reflective: first tile of the mu>0 row, fed by the mirrored mu<0 row
            double2 w = early;
            if (junction || !a.lookback_walk)
            {
               // deterministic: always the predecessor's inclusive value (bitwise run-to-run reproducible)
               while ((int)(status_flag(w) - flag_incl) < 0) { __nanosleep(32); w = ld_status(status - 1); }
               ctile = junction ? c.kbc * w.x : w.x;
            }
            else if ((int)(status_flag(w) - flag_incl) >= 0) ctile = w.x;
            else
            {
               // classic walk-back over aggregates (not bitwise reproducible; for very few chains)
               st_status(status, s_agg, flag_agg);
               double carry = 0.0, fac = 1.0;
               int q = ti.t - 1;
               const double2 *qp = status - 1;
               while (true)
               {
                  w = ld_status(qp);
                  while ((int)(status_flag(w) - flag_agg) < 0) { __nanosleep(32); w = ld_status(qp); }
                  if ((int)(status_flag(w) - flag_incl) >= 0) { carry = fma(fac, w.x, carry); break; }
                  if (q == 0)
                  {
                     while ((int)(status_flag(w) - flag_incl) < 0) { __nanosleep(32); w = ld_status(qp); }
                     carry = fma(fac, w.x, carry);
                     break;
                  }
                  carry = fma(fac, w.x, carry);
                  fac *= afull;
                  if (fac == 0.0) break; // everything further upstream has underflowed out of fp64 range
                  --q;
                  --qp;
               }
               ctile = carry;
            }
         }
         st_status(status, fma(s_alast, ctile, s_agg), flag_incl);
         s_ctile = ctile;
      }
      __syncthreads(); // (E)

      // ---- pass 2: true nodes, written back into the X1 tile (blocked, conflict-free)
      {
         const double cin = fma(tpow, s_ctile, E_t); // true inflow of this thread's chunk
         const int j0 = tid * CPT;
         double vin = cin, ap = c.alpha;
#pragma unroll
         for (int jj = 0; jj < CPT; ++jj)
         {
            const double v = fma(ap, cin, wl[jj]);
            const double u = fma(c.gamma, vin, U[jj]);
            s1[slot(j0 + jj)] = make_double2(u, v);
            vin = v;
            ap *= c.alpha;
         }
      }
      __syncthreads(); // (F)

      // ---- coalesced store
      {
         double2 *go = a.out[ti.sgn] + ti.row_off;
#pragma unroll
         for (int k = 0; k < CPT; ++k)
         {
            const int j = k * THREADS + tid;
            if (j < ti.valid) __stcs(go + j, s1[slot(j)]);
         }
      }
      if (tid == 0)
      {
         s_ticket[buf] = ahead;
         if (ahead < total_tiles) s_desc[buf] = s_desc[2];
         ahead = newer;
      }
      cur = nxt;
      buf ^= 1;
   }
   cp_async_wait<0>();
   if (tid == 0 && !drew_past_end) atomicAdd(a.ticket, 1ull);
}

// ------------------------------------------------------------------------------------------ host side
template <int KIND, bool CORR, int THREADS, int CPT>
static size_t smem_bytes()
{
   const int narr = 1 + (KIND != KIND_BE ? 1 : 0) + ((KIND == KIND_BDF && CORR) ? 1 : 0);
   return (size_t)2 * narr * TileShape<THREADS, CPT>::SLOTS * sizeof(double2); // two pipeline buffers
}

template <int KIND, bool CORR, int THREADS, int CPT>
static cudaError_t launch_one(const SweepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   static int ctas_per_sm = -1;
   const size_t smem = smem_bytes<KIND, CORR, THREADS, CPT>();
   auto kern = sweep_kernel<KIND, CORR, THREADS, CPT>;
   if (ctas_per_sm < 0)
   {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      int occ = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem);
      if (e != cudaSuccess) return e;
      ctas_per_sm = occ > 0 ? occ : 1;
   }
   const long long total = (long long)a.n_chains * a.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, THREADS, smem, st>>>(a);
   return cudaGetLastError();
}

template <int T, int CPT>
static cudaError_t launch_shape(int kind, bool corr, const SweepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   switch (kind * 2 + (corr ? 1 : 0))
   {
   case 0: return launch_one<KIND_BE, false, T, CPT>(a, num_sms, st, grid_out);
   case 1: return launch_one<KIND_BE, true, T, CPT>(a, num_sms, st, grid_out);
   case 2: return launch_one<KIND_CN, false, T, CPT>(a, num_sms, st, grid_out);
   case 3: return launch_one<KIND_CN, true, T, CPT>(a, num_sms, st, grid_out);
   case 4: return launch_one<KIND_BDF, false, T, CPT>(a, num_sms, st, grid_out);
   case 5: return launch_one<KIND_BDF, true, T, CPT>(a, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}

bool sweep_shape_supported(int threads, int cpt)
{
   return (threads == 64 || threads == 128 || threads == 256) && (cpt == 4 || cpt == 8);
}

cudaError_t launch_sweep(int kind, bool corr, int threads, int cpt, const SweepArgs &a, int num_sms, cudaStream_t st,
                         int *grid_out)
{
   switch (threads * 16 + cpt)
   {
   case 64 * 16 + 8: return launch_shape<64, 8>(kind, corr, a, num_sms, st, grid_out);
   case 64 * 16 + 4: return launch_shape<64, 4>(kind, corr, a, num_sms, st, grid_out);
   case 128 * 16 + 8: return launch_shape<128, 8>(kind, corr, a, num_sms, st, grid_out);
   case 128 * 16 + 4: return launch_shape<128, 4>(kind, corr, a, num_sms, st, grid_out);
   case 256 * 16 + 8: return launch_shape<256, 8>(kind, corr, a, num_sms, st, grid_out);
   case 256 * 16 + 4: return launch_shape<256, 4>(kind, corr, a, num_sms, st, grid_out);
   }
   return cudaErrorInvalidValue;
}
} // namespace rtk
