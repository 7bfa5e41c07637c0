// step4_kernel.cu -- the fused BDF2-step kernel with ONE scan per sub-tile instead of four (default since round 2).
//
// step_warp_kernel.cu runs the four stage sweeps of a step (BE, CN, BE, BDF; solver.cpp:706-809) back to back on a tile
// held on chip, each with its own warp scan and its own carry hand-off: 4 x (5 dependent shuffle steps + a hop from warp
// to warp).  ncu on that kernel (profiles/r01_v6_step_ncu_full_wide.txt): no wasted bytes, issue slots 55 % busy, fp64
// pipe 35 % -- the bound is the latency of that dependent chain at 4 warps per scheduler.  This kernel removes the chain
// instead of hiding it.
//
// Stage k+1 of a cell needs stage-k values of the SAME cell (plus old-level / half-level values of the upstream cell), and
// every stage is linear in the state.  A whole step is therefore ONE affine recurrence in the 4-vector
// c = (v0, v1, v2, v3) of the four stages' outflow nodes:
//
//     c_out = T c_in + e ,          (u3, v3)_j = local_j + sum_l R[l][j] c_in[l]
//
// with a lower-triangular 4x4 transfer matrix T per chunk of 8 cells and a response R of the stage-3 result to the
// chunk's inflow that are ROW CONSTANTS (uniform slab: rho, kappa_g, dx do not depend on the cell) -- tabulated once per
// solver by pack_step4_coef_kernel together with T^2, T^4, T^8, T^16.  Per sub-tile of 32 lanes x 8 cells:
//
//   pass A   every lane runs all four stages over its 8 cells with ZERO inflow: ~300 fused multiply-adds of straight-line
//            code, no shuffle, no wait (the row's first lane starts from the boundary inflow instead);
//   scan     ONE seeded Kogge-Stone scan of 4-vectors with the tabulated powers of T (5 steps, 10 multiply-adds each,
//            independent of one another) -- and ONE carry hand-off from warp to warp / ONE look-back per tile;
//   pass B   (u3, v3)_j += R[.][j] c_in : 8 multiply-adds per cell, the scalar-intensity partial sum of the step's
//            per-step output rides along (PHI).
//
// Everything around it -- warp tiles chained through shared memory, tickets, TMA tile loads / stores, decoupled look-back
// with four {value, flag} words per tile, bounded spins -- is step_warp_kernel.cu's.  The row constants (1280 bytes) are
// shared by the warps of a CTA (they work on consecutive sub-tiles of one row): a 3-entry ring filled by warp 0.
// Ragged sub-tiles (a row's last, N not a multiple of 256) finish with a short cell-by-cell loop from the true inflows.
// Formulation cross-check in numpy: tools/step4_form_check.py.
#include <atomic>
#include <cstdlib>
#include <type_traits>

#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"
#include "warp_tile.cuh"

#ifndef RT_SPIN_SHIFT
#define RT_SPIN_SHIFT 22
#endif

namespace rtk
{
namespace s4
{
constexpr int CPT = 8, WT = 32 * CPT, WT_BYTES = WT * 16;
constexpr long long SPIN_GLOBAL = 1ll << RT_SPIN_SHIFT;
constexpr long long SPIN_SHARED = 1ll << (RT_SPIN_SHIFT + 3); // a shared-memory poll is ~8x quicker than a global one

// packed row entry (doubles)
enum Index
{
   BE = 0,    // alpha gamma | B0 Bx By | U0 Ux Uy
   CN = 8,    // alpha gamma | B0 Bx By Bup2 | U0 Ux Uy Uup2
   BDF = 18,  // alpha gamma | B0 Bx By Bq BX BY Bup1 Bup2 | U0 Ux Uy Uq UX UY Uup1 Uup2
   IN = 36,   // inflow of the row's first cell, stages 0..3 (kbc * bdry_cond, solver.cpp:695-697)
   WQ = 40,   // 0.5 * quadrature weight of the row's direction (psi = (u+v)/2), one pad
   TM = 42,   // T^(2^s), s = 0..4: t00 | t10 t11 | t20 t21 t22 | t30 t31 t32 t33
   RS = 92,   // R: per cell j (Ru0 Rv0 Ru1 Rv1 Ru2 Rv2 Ru3 Rv3)
   DOUBLES = STEP4_COEF_DOUBLES
};
static_assert(RS + 8 * CPT <= DOUBLES && (TM % 2) == 0 && (RS % 2) == 0, "packed entry layout");

// all four stages over one chunk of 8 cells, homogeneous problem (old level zero, no sources): the response to a unit
// inflow of stage l.  k = the packed BE/CN/BDF constants of the row; sgn: mu > 0 row (half_ends = stage 0).
__device__ void chunk_response(const double *k, bool sgn, bool corr, int l, double *ru, double *rv, double *e)
{
   const double *be = k + BE, *cn = k + CN, *bd = k + BDF;
   double v0 = (l == 0), v1 = (l == 1), v2 = (l == 2), v3 = (l == 3);
   double up1 = (l == (sgn ? 0 : 1)) ? 1.0 : 0.0; // the half level's outflow node upstream of the chunk IS that stage's inflow
   for (int j = 0; j < CPT; ++j)
   {
      // stage 0 (old level zero)
      const double u0 = be[1] * v0;
      v0 = be[0] * v0;
      // stage 1 from (u0, v0); the upstream old-level node is zero
      double b = cn[4] * v0 + cn[3] * u0, U = cn[8] * v0 + cn[7] * u0;
      const double u1 = cn[1] * v1 + U;
      v1 = cn[0] * v1 + b;
      // stage 2 from (u1, v1)
      double xs = 0.0;
      if (corr)
      {
         b = be[4] * v1 + be[3] * u1;
         U = be[7] * v1 + be[6] * u1;
         const double u2 = be[1] * v2 + U;
         v2 = be[0] * v2 + b;
         xs = u2 + v2;
      }
      // stage 3: x1 = half level, source from stage 2
      const double hx = sgn ? u0 : u1, hy = sgn ? v0 : v1;
      b = bd[4] * hy + bd[3] * hx + bd[5] * xs + bd[8] * up1;
      U = bd[12] * hy + bd[11] * hx + bd[13] * xs + bd[16] * up1;
      ru[j] = bd[1] * v3 + U;
      v3 = bd[0] * v3 + b;
      rv[j] = v3;
      up1 = hy;
   }
   e[0] = v0; e[1] = v1; e[2] = corr ? v2 : 0.0; e[3] = v3;
}
} // namespace s4

__global__ void pack_step4_coef_kernel(const RowCoef *__restrict__ be, const RowCoef *__restrict__ cn, const RowCoef *__restrict__ bdf,
                                       const double *__restrict__ wt, double *__restrict__ out, int rows, int M, int corr)
{
   using namespace s4;
   const int r = blockIdx.x * blockDim.x + threadIdx.x;
   if (r >= rows) return;
   double *o = out + (size_t)r * DOUBLES;
   for (int i = 0; i < DOUBLES; ++i) o[i] = 0.0;
   for (int kind = 0; kind < 3; ++kind)
   {
      // beta and U as linear forms of the cell's inputs (the folding of step_warp_kernel.cu, pack_step_coef_kernel)
      const RowCoef c = (kind == 0) ? be[r] : (kind == 1 ? cn[r] : bdf[r]);
      const double io = c.io, id = c.id, qb = corr ? c.qb : 0.0;
      const double cuu = c.c_uu, cuv = (kind == 0) ? 0.0 : c.c_uv;
      const double puu = (kind == 2) ? c.p_uu : 0.0, puv = (kind == 2) ? c.p_uv : 0.0;
      const double u1 = (kind == 2) ? c.up1 : 0.0, u2 = (kind != 0) ? c.up2 : 0.0;
      const double sb = io + id, su = id - io;
      const double B0 = sb * c.qa, Bq = sb * qb, Bx = io * cuu - id * cuv, By = io * cuv + id * cuu;
      const double U0 = su * c.qa, Uq = su * qb, Ux = id * cuu + io * cuv, Uy = id * cuv - io * cuu;
      if (kind == 0)
      {
         double *k = o + BE;
         k[0] = c.alpha; k[1] = c.gamma;
         k[2] = B0; k[3] = Bx + Bq; k[4] = By + Bq;
         k[5] = U0; k[6] = Ux + Uq; k[7] = Uy + Uq;
         o[IN + 0] = c.inflow0; o[IN + 2] = c.inflow0;
      }
      else if (kind == 1)
      {
         double *k = o + CN;
         k[0] = c.alpha; k[1] = c.gamma;
         k[2] = B0; k[3] = Bx + Bq; k[4] = By + Bq; k[5] = io * u2;
         k[6] = U0; k[7] = Ux + Uq; k[8] = Uy + Uq; k[9] = id * u2;
         o[IN + 1] = c.inflow0;
      }
      else
      {
         double *k = o + BDF;
         k[0] = c.alpha; k[1] = c.gamma;
         k[2] = B0; k[3] = Bx; k[4] = By; k[5] = Bq; k[6] = io * puu - id * puv; k[7] = io * puv + id * puu;
         k[8] = io * u1; k[9] = io * u2;
         k[10] = U0; k[11] = Ux; k[12] = Uy; k[13] = Uq; k[14] = id * puu + io * puv; k[15] = id * puv - io * puu;
         k[16] = id * u1; k[17] = id * u2;
         o[IN + 3] = c.inflow0;
      }
   }
   const int dir = r % M;
   const bool sgn = dir >= (M >> 1);
   o[WQ] = 0.5 * wt[dir];
   // chunk transfer matrix, column by column, and the response of the stage-3 result
   double T[4][4];
   for (int l = 0; l < 4; ++l)
   {
      double ru[CPT], rv[CPT], e[4];
      chunk_response(o, sgn, corr != 0, l, ru, rv, e);
      for (int k = 0; k < 4; ++k) T[k][l] = (k >= l) ? e[k] : 0.0;
      for (int j = 0; j < CPT; ++j)
      {
         o[RS + j * 8 + l * 2 + 0] = ru[j];
         o[RS + j * 8 + l * 2 + 1] = rv[j];
      }
   }
   for (int s = 0; s < 5; ++s)
   {
      double *t = o + TM + 10 * s;
      int n = 0;
      for (int k = 0; k < 4; ++k)
         for (int l = 0; l <= k; ++l) t[n++] = T[k][l];
      double S[4][4];
      for (int k = 0; k < 4; ++k)
         for (int l = 0; l < 4; ++l)
         {
            double acc = 0.0;
            for (int m = l; m <= k; ++m) acc += T[k][m] * T[m][l];
            S[k][l] = (k >= l) ? acc : 0.0;
         }
      for (int k = 0; k < 4; ++k)
         for (int l = 0; l < 4; ++l) T[k][l] = S[k][l];
   }
}

void launch_pack_step4_coef(const double *be, const double *cn, const double *bdf, const double *wt, double *out, int rows, int M, bool corr,
                            cudaStream_t st)
{
   pack_step4_coef_kernel<<<(rows + 63) / 64, 64, 0, st>>>(reinterpret_cast<const RowCoef *>(be), reinterpret_cast<const RowCoef *>(cn),
                                                          reinterpret_cast<const RowCoef *>(bdf), wt, out, rows, M, corr ? 1 : 0);
}

namespace s4
{
// one tile descriptor of the ring: 32 bytes
struct Tile
{
   int row;      // row of the batch, -1 = ticket past the end
   int chain;    // look-back chain (row, or pair of mirrored rows with a reflective left boundary)
   int p;        // tile position in the chain
   int tf;       // tile in row * 16 + run_last * 8 + run_first * 4 + sgn * 2 + junction
   int part_row; // row of the scalar-intensity partial sums this tile's run adds to
   int pad[3];
};

__device__ __forceinline__ double2 lds2(const double *p) { return *reinterpret_cast<const double2 *>(p); }
// every polling loop leaves through a warp vote: all lanes leave in the same iteration (see step_warp_kernel.cu on why a
// 16-byte poll is not warp-uniform on the hardware although ptxas takes it for one)
__device__ __forceinline__ bool again(bool c) { return __any_sync(0xffffffffu, c); }
__device__ __forceinline__ void mbar_wait_v(unsigned long long *bar, unsigned parity)
{
   long long spins = 0;
   while (again(!mbar_try_wait(bar, parity)))
      if (++spins > SPIN_SHARED) __trap();
}
__device__ __forceinline__ void stg2(double *p, double a, double b)
{
   asm volatile("st.global.v2.f64 [%0], {%1, %2};" ::"l"(p), "d"(a), "d"(b) : "memory");
}
} // namespace s4

template <bool CORR, bool PHI, int NW>
__global__ void __launch_bounds__(32 * NW, 16 / NW) step4_kernel(const __grid_constant__ SweepMaps maps, const StepArgs a)
{
   using namespace s4;
   constexpr int TILE = WT * NW; // the CTA's tile: NW consecutive sub-tiles of one row

   // dynamic shared memory, 1024-byte aligned by hand: per warp [X slot 0][X slot 1][H / OUT]
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   __shared__ __align__(16) double s_coef[3][DOUBLES]; // row constants of tiles seq, seq+1 (and seq-1 for a lagging warp)
   __shared__ __align__(16) double2 s_halo[NW][2];     // X at the cell upstream of the sub-tile
   __shared__ __align__(16) double2 s_carry[4][4][NW]; // [tile sequence & 3][stage][warp]: inclusive value of a sub-tile
   __shared__ __align__(16) Tile s_desc[16];
   __shared__ __align__(8) unsigned long long s_full[NW][2];
   __shared__ __align__(8) unsigned long long s_cfull[3];
   __shared__ int s_decoded;  // ring entries decoded by warp 0 so far
   __shared__ int s_cissued;  // tiles whose row constants warp 0 has requested so far (the other warps never run ahead of it)
   __shared__ int s_done[NW]; // tiles each warp has finished (or sat out)

   // warp index through a shuffle: ptxas then KNOWS it is warp-uniform (step_warp_kernel.cu, "convergence is stated")
   const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;
   const int half_m = a.M >> 1, dm = a.reflective_left ? half_m : a.M;
   const int batch = a.ticket_batch; // <= 8
   unsigned char *const wt = tiles + warp * 3 * WT_BYTES;
   unsigned char *const th = wt + 2 * WT_BYTES;

   // ---- tickets (warp 0): one atomic per batch, the batch decoded by `batch` lanes in parallel into the ring
   long long q_start = -1;
   bool ended = false;
   int dec = 0;
   auto draw = [&]() {
      unsigned long long t0 = 0;
      if (lane == 0) t0 = atomicAdd(a.ticket, 1ull);
      t0 = __shfl_sync(0xffffffffu, t0, 0);
      q_start = (long long)(t0 - a.ticket_base) * batch;
      if (lane < batch)
      {
         const long long t = q_start + lane;
         Tile d;
         d.row = -1; d.chain = 0; d.p = 0; d.tf = 0; d.part_row = 0; d.pad[0] = d.pad[1] = d.pad[2] = 0;
         if (t < total_tiles)
         {
            int p, chain;
            if (total_tiles < 0x7fffffffll)
            {
               p = (int)((unsigned)t / (unsigned)a.n_chains);
               chain = (int)((unsigned)t - (unsigned)p * (unsigned)a.n_chains);
            }
            else
            {
               p = (int)(t / a.n_chains);
               chain = (int)(t - (long long)p * a.n_chains);
            }
            const int gl = chain / dm, pp = chain - gl * dm;
            int sgn, row, tr;
            if (a.reflective_left)
            {
               sgn = (p >= a.tiles_per_row) ? 1 : 0;
               row = gl * a.M + (sgn ? half_m + pp : half_m - 1 - pp);
               tr = sgn ? p - a.tiles_per_row : p;
            }
            else
            {
               sgn = (pp >= half_m) ? 1 : 0;
               row = chain;
               tr = p;
            }
            d.row = row; d.chain = chain; d.p = p;
            // a batch is one run of rows of the same (group, sign) at the same tile position (the host makes the batch
            // size divide M/2): its partial sum of w psi goes to one row of the partial array
            const bool run_first = (lane == 0), run_last = (lane == batch - 1) || (t + 1 >= total_tiles);
            d.tf = tr * 16 + (run_last ? 8 : 0) + (run_first ? 4 : 0) + sgn * 2 + ((tr == 0 && p > 0) ? 1 : 0);
            if (PHI)
            {
               const int dir = row - gl * a.M, isg = sgn ? dir - half_m : dir;
               d.part_row = (gl * 2 + sgn) * a.phi_slots + isg / batch;
            }
         }
         s_desc[(dec + lane) & 15] = d;
      }
      if (q_start + batch > total_tiles) ended = true;
      dec += batch;
      __syncwarp();
      if (NW > 1 && lane == 0)
      {
         __threadfence_block();
         st_flag(&s_decoded, dec);
      }
      __syncwarp();
   };

   // ---- lane 0 of each warp: TMA / bulk loads of the warp's sub-tile of tile d into its pipeline buffer b
   auto issue_loads = [&](const Tile &d, int sub_start, int b) {
      const bool halo = sub_start > 0;
      mbar_expect_tx(&s_full[warp][b], (unsigned)WT_BYTES + (halo ? 16u : 0u));
      tma_load_3d(wt + b * WT_BYTES, &maps.m[a.ix], 0, sub_start >> 3, d.row, &s_full[warp][b]);
      if (halo) bulk_load(&s_halo[warp][b], a.x + (size_t)d.row * a.Np + sub_start - 1, 16u, &s_full[warp][b]);
   };

   if (lane == 0)
   {
      mbar_init(&s_full[warp][0], 1);
      mbar_init(&s_full[warp][1], 1);
      if (warp == 0)
         for (int k = 0; k < 3; ++k) mbar_init(&s_cfull[k], 1);
      fence_proxy_async();
   }
   for (int i = threadIdx.x; i < 4 * 4 * NW; i += 32 * NW) (&s_carry[0][0][0])[i] = make_double2(0.0, 0.0);
   if (threadIdx.x < NW) s_done[threadIdx.x] = 0;
   if (threadIdx.x == 0) { s_decoded = 0; s_cissued = 0; }
   __syncthreads(); // the only CTA-wide barrier

   double acc[CPT]; // PHI: sum over the run's rows of w_m psi at this lane's cells
#pragma unroll
   for (int jj = 0; jj < CPT; ++jj) acc[jj] = 0.0;

   unsigned phase = 0;
   int issued = 0; // tiles whose loads this warp has issued (or skipped: sub-tile past the end of the row)
   for (int seq = 0;; ++seq)
   {
      int avail;
      if (warp == 0)
      {
         while (!ended && dec < seq + 3) draw();
         avail = dec;
         if (NW > 1)
         {
            // back-pressure: every warp has finished tile seq-2 (bounds the constants', carry and descriptor rings: the
            // loads issued below refill the constants of tile seq-2 with those of tile seq+1)
            long long spins = 0;
            while (again(ld_flag(&s_done[lane % NW]) < seq - 1))
               if (++spins > SPIN_SHARED) __trap();
         }
      }
      else
      {
         // wait until warp 0 has requested this tile's row constants (hence decoded it).  Waiting for the decode alone is not
         // enough: a warp that sits tiles out (row ends before its sub-tile) would run ahead of warp 0 and find the
         // constants' mbarrier in an OLDER phase of the same parity -- and take the stale entry for the new one.
         long long spins = 0;
         while (again(ld_flag(&s_cissued) <= seq))
            if (++spins > SPIN_SHARED) __trap();
         avail = ld_flag(&s_decoded); // the warp's own tile loads may run ahead as far as the descriptors are decoded
         __threadfence_block();
      }
      const Tile ti = s_desc[seq & 15];
      if (NW > 1 && warp == 0 && ti.row < 0)
      {
         // no more tiles: release the warps waiting for this tile's constants (a CTA that drew no tile at all never gets
         // to the publishing store below)
         if (lane == 0) st_flag(&s_cissued, 0x7fffffff);
         __syncwarp();
      }
      if (ti.row < 0) break;
      // warp 0: probe the predecessor tile's four look-back words first of all (lane k holds the word of stage k), so
      // that the L2 round trip overlaps the issue of the next tile's loads and pass A
      double2 *const status = a.tile_status + 4 * ((size_t)ti.chain * a.tiles_per_chain + ti.p);
      double2 wpre = make_double2(0.0, 0.0);
      if (warp == 0 && ti.p > 0) wpre = ld_status(status - 4 + (lane & 3)); // (all lanes load: no divergent branch)
      __syncwarp(); // every lane is done with the X slot and halo that are refilled now
      while (issued <= seq + 1 && issued < avail)
      {
         if (lane == 0)
         {
            const Tile d = s_desc[issued & 15];
            const int sub = (d.tf >> 4) * TILE + warp * WT;
            if (d.row >= 0 && sub < a.N) issue_loads(d, sub, issued & 1);
            if (warp == 0 && d.row >= 0)
            {
               // the row's constants, once per CTA tile
               const int cs = issued % 3;
               mbar_expect_tx(&s_cfull[cs], (unsigned)(DOUBLES * 8));
               bulk_load(&s_coef[cs][0], a.coef4 + (size_t)d.row * DOUBLES, (unsigned)(DOUBLES * 8), &s_cfull[cs]);
            }
         }
         ++issued;
      }
      __syncwarp();
      if (NW > 1 && warp == 0)
      {
         if (lane == 0)
         {
            __threadfence_block();
            st_flag(&s_cissued, issued);
         }
         __syncwarp();
      }
      const int tile_start = (ti.tf >> 4) * TILE, sub_start = tile_start + warp * WT;
      const int valid_total = min(TILE, a.N - tile_start);
      const int valid = min(WT, valid_total - warp * WT); // cells of this warp's sub-tile inside the row (<= 0: none)
      if (valid > 0)
      {
         const int buf = seq & 1;
         const bool sgn = (ti.tf & 2) != 0, junction = (ti.tf & 1) != 0 && warp == 0;
         const bool is_last = valid_total <= (warp + 1) * WT; // the row's tile ends inside (or at the end of) this sub-tile
         const bool row_first_sub = (sub_start == 0);         // starts from the boundary (or the mirrored row: junction)
         double2 *const carry_slot = &s_carry[seq & 3][0][0];

         mbar_wait_v(&s_full[warp][buf], (phase >> buf) & 1u); // sub-tile and halo cell have landed
         phase ^= 1u << buf;
         mbar_wait_v(&s_cfull[seq % 3], (unsigned)(seq / 3) & 1u); // the row's constants have landed

         const unsigned char *const sx = wt + buf * WT_BYTES;
         const double *const cf = s_coef[seq % 3];
         const int j0 = lane * CPT;
         auto ldx = [&](int j) { return *reinterpret_cast<const double2 *>(sx + swz(j)); };
         // X outflow node of the cell upstream of this lane's first cell (a row's first cell has no upstream terms,
         // solver.cpp:695-697)
         double up2_0 = 0.0;
         if (lane == 0) { if (sub_start > 0) up2_0 = s_halo[warp][buf].y; }
         else up2_0 = ldx(j0 - 1).y;

         // ---- carry into the sub-tile: the previous CTA tile's four inclusive values (global look-back words, warp 0) or
         // the previous warp's (shared memory); lane k polls the word of stage k, all lanes leave together
         double c0 = 0.0, c1 = 0.0, c2 = 0.0, c3 = 0.0;
         auto fetch_carry = [&]() {
            double2 w = wpre;
            long long spins = 0;
            if (warp == 0)
            {
               while (again((int)(status_flag(w) - flag_incl) < 0))
               {
                  w = ld_status(status - 4 + (lane & 3));
                  if (++spins > SPIN_GLOBAL) __trap();
               }
            }
            else
            {
               // The writer publishes the words of stages 0..3 in that order.  Polling all four with one 16-byte load per lane
               // costs 8 shared-memory wavefronts per iteration (2-way conflicted) and ran ~12 times per sub-tile: a quarter of
               // the kernel's shared-memory traffic (profiles/r02_step4_cellmajor_ncu_digest.txt).  So: every lane polls the
               // 8-byte sequence number of the LAST word (one broadcast wavefront), then the four words are read once -- each
               // still checked against its own sequence number, so nothing rests on the order the stores become visible in.
               const double2 *src = carry_slot + (lane & 3) * NW + (warp - 1);
               const int *last_seq = reinterpret_cast<const int *>(carry_slot + 3 * NW + (warp - 1)) + 2; // low word of .y
               while (again(ld_flag(last_seq) != seq + 1))
                  if (++spins > SPIN_SHARED) __trap();
               w = lds_word(src);
               while (again((int)__double_as_longlong(w.y) != seq + 1))
               {
                  w = lds_word(src);
                  if (++spins > SPIN_SHARED) __trap();
               }
            }
            c0 = __shfl_sync(0xffffffffu, w.x, 0);
            c1 = __shfl_sync(0xffffffffu, w.x, 1);
            c2 = __shfl_sync(0xffffffffu, w.x, 2);
            c3 = __shfl_sync(0xffffffffu, w.x, 3);
         };

         // inflow of the row's first lane: the boundary values of the four stages, or (junction) the mirrored row's
         // outflow times the boundary multiplier kbc = 1, 2, 1, 6 (BE, CN, BE, BDF)
         double sd0 = 0.0, sd1 = 0.0, sd2 = 0.0, sd3 = 0.0;
         if (junction)
         {
            // (kept OUTSIDE the row_first_sub block: nested there, ptxas stops treating the tile loop as convergent and puts a
            // divergence check with an out-of-line WARPSYNC path around every shuffle and vote of the kernel)
            fetch_carry();
            c1 *= 2.0; c3 *= 6.0;
         }
         if (row_first_sub)
         {
            if (!junction)
            {
               const double2 i01 = lds2(cf + IN), i23 = lds2(cf + IN + 2);
               c0 = i01.x; c1 = i01.y; c2 = i23.x; c3 = i23.y;
            }
            if (lane == 0) { sd0 = c0; sd1 = c1; sd2 = c2; sd3 = c3; }
            c0 = c1 = c2 = c3 = 0.0;
         }

         // ================= pass A: the four stages of this lane's 8 cells from the lane's own inflow (zero, or the boundary)
         double2 r[CPT], o[CPT];
         double xs[CPT];
         double e0, e1, e2 = 0.0, e3;
         {
            // stage 0: BE predictor, x1 = X (ends^n)
            const double2 ag = lds2(cf + BE), k01 = lds2(cf + BE + 2), k23 = lds2(cf + BE + 4), k45 = lds2(cf + BE + 6);
            double v = sd0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double2 x = ldx(j0 + jj);
               const double b = fma(k23.x, x.y, fma(k01.y, x.x, k01.x));
               const double U = fma(k45.y, x.y, fma(k45.x, x.x, k23.y));
               r[jj].x = fma(ag.y, v, U);
               v = fma(ag.x, v, b);
               r[jj].y = v;
            }
            e0 = v;
            if (lane == 0) bulk_wait_read<0>(); // the previous tile's TMA store has read H / OUT
            __syncwarp();
            if (sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = r[jj];
            }
         }
         {
            // stage 1: CN corrector, x1 = stage-0 result, upstream X
            const double2 ag = lds2(cf + CN), k01 = lds2(cf + CN + 2), k23 = lds2(cf + CN + 4), k45 = lds2(cf + CN + 6), k67 = lds2(cf + CN + 8);
            double v = sd1, up2 = up2_0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double b = fma(k23.y, up2, fma(k23.x, r[jj].y, fma(k01.y, r[jj].x, k01.x)));
               const double U = fma(k67.y, up2, fma(k67.x, r[jj].y, fma(k45.y, r[jj].x, k45.x)));
               r[jj].x = fma(ag.y, v, U);
               v = fma(ag.x, v, b);
               r[jj].y = v;
               if (jj + 1 < CPT) up2 = ldx(j0 + jj).y;
            }
            e1 = v;
            if (!sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = r[jj];
            }
         }
         if (CORR)
         {
            // stage 2: BE predictor, x1 = stage-1 result; only u+v survives (the BDF stage's correction source)
            const double2 ag = lds2(cf + BE), k01 = lds2(cf + BE + 2), k23 = lds2(cf + BE + 4), k45 = lds2(cf + BE + 6);
            double v = sd2;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double b = fma(k23.x, r[jj].y, fma(k01.y, r[jj].x, k01.x));
               const double U = fma(k45.y, r[jj].y, fma(k45.x, r[jj].x, k23.y));
               const double u = fma(ag.y, v, U);
               v = fma(ag.x, v, b);
               xs[jj] = u + v;
            }
            e2 = v;
         }
         {
            // stage 3: BDF corrector, x1 = half_ends (parked in H), x2 = X, zero inflow -> the local part of ends^(n+1)
            const double *c = cf + BDF;
            const double2 ag = lds2(c), k01 = lds2(c + 2), k23 = lds2(c + 4), k45 = lds2(c + 6), k67 = lds2(c + 8), k89 = lds2(c + 10),
                          kab = lds2(c + 12), kcd = lds2(c + 14), kef = lds2(c + 16);
            double v = sd3, up1 = 0.0, up2 = up2_0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double2 h = *reinterpret_cast<const double2 *>(th + swz(j0 + jj));
               const double2 x = ldx(j0 + jj);
               double b = fma(k23.x, h.y, fma(k01.y, h.x, k01.x));
               double u = fma(kab.x, h.y, fma(k89.y, h.x, k89.x));
               if (CORR)
               {
                  b = fma(k23.y, xs[jj], b);
                  u = fma(kab.y, xs[jj], u);
               }
               b = fma(k67.y, up2, fma(k67.x, up1, fma(k45.y, x.y, fma(k45.x, x.x, b))));
               u = fma(kef.y, up2, fma(kef.x, up1, fma(kcd.y, x.y, fma(kcd.x, x.x, u))));
               o[jj].x = fma(ag.y, v, u);
               v = fma(ag.x, v, b);
               o[jj].y = v;
               up1 = h.y;
               up2 = x.y;
            }
            e3 = v;
         }

         // ================= one seeded scan of the 4-vector of chunk outflows with the tabulated powers of T
         if (!row_first_sub) fetch_carry();
#pragma unroll
         for (int s = -1; s < 5; ++s)
         {
            // s = -1: fold the carry into lane 0's chunk (same matrix as step 0); s >= 0: Kogge-Stone step over 2^s lanes
            const double *t = cf + TM + 10 * (s < 0 ? 0 : s);
            const double2 t0 = lds2(t), t1 = lds2(t + 2), t2 = lds2(t + 4), t3 = lds2(t + 6), t4 = lds2(t + 8);
            double q0, q1, q2, q3;
            bool take;
            if (s < 0)
            {
               q0 = c0; q1 = c1; q2 = c2; q3 = c3;
               take = (lane == 0);
            }
            else
            {
               q0 = __shfl_up_sync(0xffffffffu, e0, 1 << s);
               q1 = __shfl_up_sync(0xffffffffu, e1, 1 << s);
               q2 = CORR ? __shfl_up_sync(0xffffffffu, e2, 1 << s) : 0.0;
               q3 = __shfl_up_sync(0xffffffffu, e3, 1 << s);
               take = (lane >= (1 << s));
            }
            // t00 | t10 t11 | t20 t21 t22 | t30 t31 t32 t33  =  t0.x | t0.y t1.x | t1.y t2.x t2.y | t3.x t3.y t4.x t4.y
            if (take)
            {
               if (CORR)
               {
                  e3 = fma(t4.y, q3, fma(t3.x, q0, e3)) + fma(t4.x, q2, t3.y * q1);
                  e2 = fma(t2.y, q2, fma(t2.x, q1, fma(t1.y, q0, e2)));
               }
               else e3 = fma(t4.y, q3, fma(t3.y, q1, fma(t3.x, q0, e3)));
               e1 = fma(t1.x, q1, fma(t0.y, q0, e1));
               e0 = fma(t0.x, q0, e0);
            }
         }
         // true inflow of this lane's chunk
         double ci0 = __shfl_up_sync(0xffffffffu, e0, 1), ci1 = __shfl_up_sync(0xffffffffu, e1, 1);
         double ci2 = CORR ? __shfl_up_sync(0xffffffffu, e2, 1) : 0.0, ci3 = __shfl_up_sync(0xffffffffu, e3, 1);
         if (lane == 0) { ci0 = c0; ci1 = c1; ci2 = c2; ci3 = c3; } // (zero in a row's first sub-tile: its lane 0 is seeded)
         if (valid == WT)
         {
            if (lane == 31)
            {
               if (!is_last)
               {
                  sts_word(carry_slot + 0 * NW + warp, e0, seq + 1);
                  sts_word(carry_slot + 1 * NW + warp, e1, seq + 1);
                  sts_word(carry_slot + 2 * NW + warp, e2, seq + 1);
                  sts_word(carry_slot + 3 * NW + warp, e3, seq + 1);
               }
               else
               {
                  st_status(status + 0, e0, flag_incl);
                  st_status(status + 1, e1, flag_incl);
                  st_status(status + 2, e2, flag_incl);
                  st_status(status + 3, e3, flag_incl);
               }
            }
            __syncwarp();
            // ================= pass B: add the response to the chunk's inflow; the step's result leaves by TMA store
            const double wq = PHI ? cf[WQ] : 0.0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double *rr = cf + RS + jj * 8;
               const double2 r0 = lds2(rr), r1 = lds2(rr + 2), r3 = lds2(rr + 6);
               double u = fma(r1.x, ci1, fma(r0.x, ci0, o[jj].x));
               double v = fma(r1.y, ci1, fma(r0.y, ci0, o[jj].y));
               if (CORR)
               {
                  const double2 r2 = lds2(rr + 4);
                  u = fma(r2.x, ci2, u);
                  v = fma(r2.y, ci2, v);
               }
               u = fma(r3.x, ci3, u);
               v = fma(r3.y, ci3, v);
               *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = make_double2(u, v);
               if (PHI) acc[jj] = fma(wq, u + v, acc[jj]);
            }
         }
         else
         {
            // ---- ragged sub-tile (the row's last): cell by cell from the true inflows; the lane that holds the row's last
            // cell publishes the chain's inclusive values
            const bool rfl = row_first_sub && lane == 0;
            double v0 = rfl ? sd0 : ci0, v1 = rfl ? sd1 : ci1, v2 = rfl ? sd2 : ci2, v3 = rfl ? sd3 : ci3;
            double up1 = rfl ? 0.0 : (sgn ? ci0 : ci1), up2 = up2_0;
            const int last = valid - 1;
            const bool mine = (lane == last / CPT);
            const int jo = last % CPT;
            const double *be = cf + BE, *cn = cf + CN, *bd = cf + BDF;
            double l0 = 0.0, l1 = 0.0, l2 = 0.0, l3 = 0.0; // the four stages' outflow at the row's last cell
#pragma unroll 1
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double2 x = ldx(j0 + jj);
               double b = fma(be[4], x.y, fma(be[3], x.x, be[2])), U = fma(be[7], x.y, fma(be[6], x.x, be[5]));
               const double u0 = fma(be[1], v0, U);
               v0 = fma(be[0], v0, b);
               b = fma(cn[5], up2, fma(cn[4], v0, fma(cn[3], u0, cn[2])));
               U = fma(cn[9], up2, fma(cn[8], v0, fma(cn[7], u0, cn[6])));
               const double u1 = fma(cn[1], v1, U);
               v1 = fma(cn[0], v1, b);
               double xsum = 0.0;
               if (CORR)
               {
                  b = fma(be[4], v1, fma(be[3], u1, be[2]));
                  U = fma(be[7], v1, fma(be[6], u1, be[5]));
                  const double u2 = fma(be[1], v2, U);
                  v2 = fma(be[0], v2, b);
                  xsum = u2 + v2;
               }
               const double hx = sgn ? u0 : u1, hy = sgn ? v0 : v1;
               b = fma(bd[4], hy, fma(bd[3], hx, bd[2]));
               U = fma(bd[12], hy, fma(bd[11], hx, bd[10]));
               if (CORR)
               {
                  b = fma(bd[5], xsum, b);
                  U = fma(bd[13], xsum, U);
               }
               b = fma(bd[9], up2, fma(bd[8], up1, fma(bd[7], x.y, fma(bd[6], x.x, b))));
               U = fma(bd[17], up2, fma(bd[16], up1, fma(bd[15], x.y, fma(bd[14], x.x, U))));
               const double u3 = fma(bd[1], v3, U);
               v3 = fma(bd[0], v3, b);
               *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = make_double2(u3, v3);
               up1 = hy;
               up2 = x.y;
               if (jj == jo) { l0 = v0; l1 = v1; l2 = v2; l3 = v3; } // (uniform branch)
            }
            __syncwarp();
            if (mine)
            {
               st_status(status + 0, l0, flag_incl);
               st_status(status + 1, l1, flag_incl);
               st_status(status + 2, l2, flag_incl);
               st_status(status + 3, l3, flag_incl);
            }
            __syncwarp();
            if (PHI)
            {
               const double wq = cf[WQ];
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj)
               {
                  const double2 t = *reinterpret_cast<const double2 *>(th + swz(j0 + jj));
                  acc[jj] = fma(wq, t.x + t.y, acc[jj]);
               }
            }
         }
         fence_proxy_async(); // generic-proxy writes -> visible to the TMA store
         __syncwarp();
         if (lane == 0)
         {
            tma_store_3d(&maps.m[a.iout], 0, sub_start >> 3, ti.row, th); // cells past the row's end are clipped
            bulk_commit();
         }
         __syncwarp();
         if (PHI && (ti.tf & 8))
         {
            // end of the run: this lane's 8 partial sums (64 contiguous bytes; the row pitch covers whole tiles)
            double *dst = a.phi_part + (size_t)ti.part_row * (size_t)a.phi_pitch + sub_start + j0;
            stg2(dst + 0, acc[0], acc[1]);
            stg2(dst + 2, acc[2], acc[3]);
            stg2(dst + 4, acc[4], acc[5]);
            stg2(dst + 6, acc[6], acc[7]);
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj) acc[jj] = 0.0;
         }
      }
      if (NW > 1)
      {
         __syncwarp();
         if (lane == 0) st_flag(&s_done[warp], seq + 1);
         __syncwarp(); // reconverge BEFORE the back edge (a single-lane block that ends the loop body makes the loop divergent for ptxas)
      }
   }
   if (lane == 0) bulk_wait_read<0>();
   if (warp == 0 && lane == 0 && q_start < total_tiles) atomicAdd(a.ticket, 1ull); // every CTA makes exactly one draw past the end
}

// scalar intensity of the step just completed from the partial sums the step kernel left behind:
// phi[gl][c] = sum over slots of part[(gl*2+0)*slots + s][N-1-c] (mu<0 rows, stored in sweep order) + part[(gl*2+1)*slots + s][c]
__global__ void __launch_bounds__(256) phi_combine_kernel(const double *__restrict__ part, long long pitch, int slots, int N, int groups,
                                                          double *__restrict__ phi)
{
   const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
   if (idx >= (long long)groups * N) return;
   const int c = (int)(idx % N), gl = (int)(idx / N);
   const double *pn = part + (size_t)(gl * 2 + 0) * slots * pitch + (N - 1 - c);
   const double *pp = part + (size_t)(gl * 2 + 1) * slots * pitch + c;
   double sum = 0.0;
   for (int s = 0; s < slots; ++s) sum += __ldcs(pn + (size_t)s * pitch);
   for (int s = 0; s < slots; ++s) sum += __ldcs(pp + (size_t)s * pitch);
   phi[idx] = sum;
}

void launch_phi_combine(const double *part, long long pitch, int slots, int N, int groups, double *phi, cudaStream_t st)
{
   const long long total = (long long)groups * N;
   phi_combine_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(part, pitch, slots, N, groups, phi);
}

// ------------------------------------------------------------------------------------------ host side
template <bool CORR, bool PHI, int NW>
static cudaError_t launch_step4_one(const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   auto kern = step4_kernel<CORR, PHI, NW>;
   const size_t smem = (size_t)NW * 3 * s4::WT_BYTES + 1024; // per warp 2 X + H/OUT, + alignment slack
   int occ = plan_slot->load(std::memory_order_acquire);
   if (occ <= 0)
   {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * NW, smem);
      if (e != cudaSuccess) return e;
      if (occ < 1) occ = 1;
      plan_slot->store(occ, std::memory_order_release);
   }
   const char *cap = getenv("RT_CUDA_STEP_CTAS"); // resident CTAs per SM (experiments)
   if (cap && atoi(cap) >= 1 && atoi(cap) < occ) occ = atoi(cap);
   const long long total = (long long)a.n_chains * a.tiles_per_chain;
   long long grid = (long long)num_sms * occ;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, 32 * NW, smem, st>>>(maps, a);
   return cudaGetLastError();
}

bool step4_chain_supported(int nw) { return nw == 2 || nw == 4 || nw == 8; }

cudaError_t launch_step4(bool corr, bool phi, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st,
                         int *grid_out)
{
   if (a.ticket_batch < 1 || a.ticket_batch > 8 || !a.coef4) return cudaErrorInvalidValue;
   if (phi && (!a.phi_part || a.phi_slots < 1 || (a.M / 2) % a.ticket_batch != 0)) return cudaErrorInvalidValue;
#define RT_STEP4_CASE(C, P, W) \
   if (corr == C && phi == P && chain_warps == W) return launch_step4_one<C, P, W>(maps, a, num_sms, st, grid_out);
   RT_STEP4_CASE(true, false, 4) RT_STEP4_CASE(false, false, 4) RT_STEP4_CASE(true, true, 4) RT_STEP4_CASE(false, true, 4)
   RT_STEP4_CASE(true, false, 2) RT_STEP4_CASE(false, false, 2) RT_STEP4_CASE(true, true, 2) RT_STEP4_CASE(false, true, 2)
   RT_STEP4_CASE(true, false, 8) RT_STEP4_CASE(false, false, 8) RT_STEP4_CASE(true, true, 8) RT_STEP4_CASE(false, true, 8)
#undef RT_STEP4_CASE
   return cudaErrorInvalidValue;
}
} // namespace rtk
