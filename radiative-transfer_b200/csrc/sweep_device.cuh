// sweep_device.cuh -- device-side building blocks shared by the per-stage sweep kernel (sweep_kernel.cu) and the fused
// BDF2-step kernel (step_kernel.cu): PTX wrappers (mbarrier, TMA, bulk copies), the swizzled tile image, the per-row
// stepper constants, look-back status words and the ticket decoder.
#pragma once
#include "rt_common.cuh"

namespace rtk
{
// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
   asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
   asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long *bar, unsigned parity)
{
   unsigned ok;
   asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
   return ok != 0;
}
// Every spin in this kernel is bounded: a wait that does not end within seconds of polling is a bug (or a dead
// predecessor) and traps -- the launch fails loudly instead of hanging the device.
constexpr long long SPIN_LIMIT = 1ll << 22;
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
   long long spins = 0;
   while (!mbar_try_wait(bar, parity))
      if (++spins > SPIN_LIMIT) __trap();
}
// 3-D tiled tensor load: coordinates (element in chunk, chunk in row, row)
__device__ __forceinline__ void tma_load_3d(void *dst, const CUtensorMap *map, int c0, int c1, int c2, unsigned long long *bar)
{
   asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar))
                : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap *map, int c0, int c1, int c2, const void *src)
{
   asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3}], [%4];"
                ::"l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(src))
                : "memory");
}
// plain bulk copy global -> shared (16-byte granules), completing on an mbarrier
__device__ __forceinline__ void bulk_load(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
   asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar))
                : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// HBM -> L2 only (no shared-memory landing zone needed): used two tiles ahead of the tile being scanned
__device__ __forceinline__ void prefetch_l2_bulk(const void *gmem, unsigned bytes)
{
   asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem), "r"(bytes) : "memory");
}

template <int THREADS, int CPT_>
struct TileShape
{
   static constexpr int CPT = CPT_; // cells per thread: 8 or 4
   static constexpr int TILE = THREADS * CPT;
   static constexpr int TILE_BYTES = TILE * 16;
   static constexpr int WARPS = THREADS / 32;
};

// Byte offset of cell j of a tile in its SWIZZLE_128B shared-memory image (tile base 1024-byte aligned):
// 128-byte line r = j/8 holds 8 cells; the 16-byte chunk index is XORed with (r mod 8).
__device__ __forceinline__ int swz(int j) { return ((j >> 3) << 7) | ((((j & 7) ^ ((j >> 3) & 7))) << 4); }

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

// The stepper constants of one (stage kind, direction, material state): sigma = rho*kappa_g, srcB = sigma*B_g, cor1..3 the
// correction terms of the group.  Used once per row by coef_table_kernel (the reference's uniform slab) and once per
// CELL by the cell-dependent kernels (var_kernel.cu).  inflow0 is left to the caller.
__device__ __forceinline__ RowCoef coef_from_material(int kind, bool corr, double mu, double sig, double srcB, double cor1,
                                                      double cor2, double cor3, double dx, double tau, double dt, double beta)
{
   RowCoef c;
   const double amu = fabs(mu);
   const double ctau = C_LIGHT * tau;
   const double hq = 0.5 * ctau * dx; // 0.5*c*timestep*dx (solver.cpp:337)
   c.qa = hq * srcB;
   c.qb = 0.0;
   if (corr)
   {
      // total_correction = (cor1*psi + cor2)*mu*beta - cor3*mu^2*beta^2 (correction.cpp:392-394), psi = (u+v)/2
      const double mb = mu * beta;
      c.qa += hq * (cor2 * mb - cor3 * mb * mb);
      c.qb = 0.5 * hq * cor1 * mb;
   }
   double k, d;
   c.p_uu = c.p_uv = c.p_vu = c.p_vv = 0.0;
   c.up1 = c.up2 = 0.0;
   if (kind == KIND_BE)
   {
      // solver.cpp:324-334,344-345,381-382
      k = ctau * amu;
      d = 0.5 * ((1.0 + ctau * sig) * dx + k);
      c.c_uu = 0.5 * dx; c.c_uv = 0.0; c.c_vu = 0.0; c.c_vv = 0.5 * dx;
      c.kbc = 1.0;
   }
   else if (kind == KIND_CN)
   {
      // solver.cpp:411-425,435-436,469-470
      const double t = 0.5 * ctau * sig;
      k = 0.5 * C_LIGHT * amu * tau;
      d = 0.5 * ((1.0 + t) * dx + k);
      const double e1 = 0.5 * ((1.0 - t) * dx - k);
      c.c_uu = e1; c.c_uv = -0.5 * k; c.c_vu = 0.5 * k; c.c_vv = e1;
      c.up2 = k; // a * in_prev, in_prev = prev_ends(j-1).v
      c.kbc = 2.0;
   }
   else
   {
      // solver.cpp:498-512,522-527,561-566 ; streaming constant uses the FULL dt
      const double t = C_LIGHT * sig * tau / 6.0;
      k = C_LIGHT * amu * dt / 6.0;
      d = 0.5 * ((1.0 + t) * dx + k);
      const double f1 = 0.5 * ((1.0 - 4.0 * t) * dx - 4.0 * k), g1 = 0.5 * (k + t * dx);
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
   c.inflow0 = 0.0;
   return c;
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

// Everything the CTA needs to know about one ticket; decoded by ONE thread and handed to the others through
// shared memory.
struct TileDesc
{
   long long row_off;    // element offset of the tile's first cell in a time-level buffer
   long long status_idx; // index of the tile's look-back word
   int row, p, sgn, tile_start, valid, junction;
   int gl, dir;          // local group and direction of the row (row = gl*M + dir)
};

// ticket -> (tile position p, chain) -> (row, tile in row); p-major so that all chains advance together.
// Consecutive tickets are decoded by increments; the divisions only run after a jump (new batch of tickets).
struct TicketCursor
{
   long long ticket;
   int chain, p, gl, pp; // gl = chain / dm, pp = chain % dm with dm = M/2 (reflective left) or M
};

template <typename Args>
__device__ __forceinline__ TileDesc decode_ticket(const Args &a, TicketCursor &cu, long long ticket, int tile_cells)
{
   const int half_m = a.M >> 1, dm = a.reflective_left ? half_m : a.M;
   if (ticket == cu.ticket + 1)
   {
      ++cu.chain;
      if (++cu.pp == dm) { cu.pp = 0; ++cu.gl; }
      if (cu.chain == a.n_chains) { cu.chain = 0; cu.gl = 0; cu.pp = 0; ++cu.p; }
   }
   else
   {
      cu.p = (int)(ticket / a.n_chains);
      cu.chain = (int)(ticket - (long long)cu.p * a.n_chains);
      cu.gl = cu.chain / dm;
      cu.pp = cu.chain - cu.gl * dm;
   }
   cu.ticket = ticket;
   TileDesc ti;
   int t;
   ti.p = cu.p;
   if (a.reflective_left)
   {
      ti.sgn = (cu.p >= a.tiles_per_row) ? 1 : 0;
      ti.row = cu.gl * a.M + (ti.sgn ? half_m + cu.pp : half_m - 1 - cu.pp);
      t = ti.sgn ? cu.p - a.tiles_per_row : cu.p;
   }
   else
   {
      ti.sgn = (cu.pp >= half_m) ? 1 : 0;
      ti.row = cu.chain;
      t = cu.p;
   }
   ti.junction = (t == 0 && cu.p > 0) ? 1 : 0; // reflective: first tile of the mu>0 row, fed by the mirrored mu<0 row
   ti.tile_start = t * tile_cells;
   ti.valid = min(tile_cells, a.N - ti.tile_start);
   ti.row_off = (long long)ti.row * a.Np + ti.tile_start;
   ti.status_idx = (long long)cu.chain * a.tiles_per_chain + cu.p;
   ti.gl = cu.gl;
   ti.dir = ti.row - cu.gl * a.M;
   return ti;
}

// Tickets are drawn in batches of `ticket_batch` consecutive tickets per atomic (one atomic per tile on a single
// address would be ~4e8 atomics/s at 6 TB/s with 8 KB tiles).  A batch is consumed in increasing order, so a CTA only
// ever waits on tickets smaller than its own, all held by running CTAs.  Host accounting: the counter advances by one
// per draw; every CTA makes exactly ONE draw that lies past the end (at the latest when it exits).
// No branch depends on a freshly drawn value, so the atomic's latency is only paid where the ticket is used.
struct TicketQueue
{
   long long last_taken = 0, q_next = 0, q_end = 0, q_start = -1;

   template <typename Args>
   __device__ __forceinline__ long long take(const Args &a, long long total_tiles)
   {
      if (last_taken >= total_tiles) return last_taken;
      if (q_next == q_end)
      {
         q_start = (long long)(atomicAdd(a.ticket, 1ull) - a.ticket_base) * a.ticket_batch;
         q_next = q_start;
         q_end = q_start + a.ticket_batch;
      }
      last_taken = q_next++;
      return last_taken;
   }
   // at CTA exit: true when none of this CTA's draws lay past the end yet (q_start = -1: no draw at all)
   template <typename Args>
   __device__ __forceinline__ void finish(const Args &a, long long total_tiles)
   {
      if (q_start < total_tiles) atomicAdd(a.ticket, 1ull);
   }
};

// Pass 1 of one stage for the CPT cells a thread owns: per-cell right-hand sides -> (beta_j, U_j) and the thread-local
// scan with zero inflow (wl).  x1 / x2 are the stage's time levels at the cell (BE, CN: x1 = ends before the stage ;
// BDF: x1 = half_ends, x2 = prev_ends), xsum = (u+v) of the cell the correction source is evaluated at, up1 / up2 the
// outflow nodes of x1 / x2 at the cell upstream of the thread's first cell, skip_up = true for the first cell of a row
// (solver.cpp:695-697: all inflow carries equal bdry_cond there).  Formulas: solver.cpp:344-345,381-382 (BE),
// 435-436,469-470 (CN), 522-527,561-566 (BDF) in the (u,v) frame of rt_common.cuh.
template <int KIND, bool CORR, int CPT>
__device__ __forceinline__ void stage_pass1(const RowCoef &c, const double2 (&x1)[CPT], const double2 (&x2)[CPT],
                                            const double (&xsum)[CPT], double up1, double up2, bool skip_up,
                                            double (&wl)[CPT], double (&U)[CPT])
{
   double w = 0.0;
#pragma unroll
   for (int jj = 0; jj < CPT; ++jj)
   {
      double q = c.qa;
      if (CORR) q = fma(c.qb, xsum[jj], c.qa);
      double ru, rv;
      if (KIND == KIND_BE)
      {
         ru = fma(c.c_uu, x1[jj].x, q);
         rv = fma(c.c_uu, x1[jj].y, q);
      }
      else
      {
         ru = q + c.c_uu * x1[jj].x + c.c_uv * x1[jj].y;
         rv = q + c.c_vu * x1[jj].x + c.c_vv * x1[jj].y;
      }
      if (KIND == KIND_BDF)
      {
         ru += c.p_uu * x2[jj].x + c.p_uv * x2[jj].y;
         rv += c.p_vu * x2[jj].x + c.p_vv * x2[jj].y;
      }
      if (KIND != KIND_BE && !(skip_up && jj == 0))
      {
         if (KIND == KIND_BDF) ru += c.up1 * up1 + c.up2 * up2;
         else ru += c.up2 * up2;
      }
      const double beta = c.io * ru + c.id * rv;
      U[jj] = c.id * ru - c.io * rv;
      w = fma(c.alpha, w, beta);
      wl[jj] = w;
      if (KIND == KIND_BDF) up1 = x1[jj].y;
      if (KIND != KIND_BE) up2 = x2[jj].y;
   }
}

// the row's constants from shared memory; entries that are structurally fixed for the stage kind are folded at compile
// time (BE: no cross terms, no X2 ; CN: no X2 cell terms ; symmetric / antisymmetric pairs)
template <int KIND, bool CORR>
__device__ __forceinline__ RowCoef fold_coef(const RowCoef &cs)
{
   RowCoef c;
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
   return c;
}

} // namespace rtk
