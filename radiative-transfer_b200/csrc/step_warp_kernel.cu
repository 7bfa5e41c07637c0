// step_warp_kernel.cu -- the fused BDF2-step kernel (see step_kernel.cu for the stage algebra) built from WARP tiles.
//
// Profile of the CTA-wide version (profiles/r01_v5_step_cta_ncu.txt): 24 % of the warp time waits at __syncthreads,
// another ~40 % hangs on work only thread 0 does (tickets, TMA issue, look-back), 12 warps per SM.  A row's recurrence
// needs no CTA-wide barrier at all:
//
//   * a warp owns a sub-tile of 32 lanes x CPT cells (256 cells, 4 KB) with its own TMA pipeline (2 x X, one H/OUT
//     tile, one mbarrier per X slot).  The NW warps of a CTA take NW CONSECUTIVE sub-tiles of one row (the CTA's tile,
//     one ticket) and hand the recurrence carry from warp to warp through shared memory ({value, sequence} words, a
//     few hundred cycles per hop); only warp 0 looks back to the previous CTA tile through global memory (a few
//     thousand cycles per hop under load -- with single-warp tiles that hop bounds the whole launch once there are
//     more tiles in flight than chains).  No __syncthreads after the prologue: the warps of a CTA run skewed by one
//     scan each and never wait for each other in steady state.
//   * seeded scan: the carry into the sub-tile is folded into lane 0's chunk BEFORE the warp scan, so the scan
//     delivers true values directly: no alpha^(CPT*lane) powers, no fix-up pass.  Warp 0 probes the predecessor
//     tile's four status words (one per stage) when the tile starts, lanes 0..3 holding one word each.
//   * pass 2 re-runs the cell recurrence v_j = alpha v_(j-1) + beta_j from the true inflow (no alpha^j powers).
//   * tickets are drawn in batches (one atomic per batch) by warp 0 and the whole batch is decoded by as many lanes
//     in parallel into a 16-entry descriptor ring shared by the CTA's warps.
//   * only the stage result that is this row's half_ends (stage 0 for mu>0 rows, stage 1 for mu<0 rows) is
//     parked in shared memory, thread-private; the stage-3 result overwrites it in place and leaves by TMA store.
//   * the row constants of the three stage kinds come as one packed 256-byte entry (pack_step_coef_kernel).
#include <cstdlib>
#include <type_traits>

#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"
#include "warp_tile.cuh"

#ifndef STEP_SPIN_SLEEP
// nanoseconds a warp sleeps between two polls of the previous warp's carry word.  Measured on C4 at 20 steps, where the
// board runs into its power cap: 0 -> 100 ns leaves the rate where it is (+0.1 ... +0.4 %) and takes 5-10 % off the peak
// board power (870 -> 800-830 W; 250 ns: 725 W) -- the polls were instructions and watts, not progress.
#define STEP_SPIN_SLEEP 100
#endif

// Polling loops and warp convergence.  This kernel has no CTA barrier in its tile loop, and ptxas is told that the warp
// index is warp-uniform (see the kernel's first lines), so it treats the whole tile loop as convergent: no BRA.DIV check
// and no BSSY/BSYNC pair around the shuffles -- and none after the polling loops either, because a load of ONE address
// by all lanes counts as a uniform value.  For the 16-byte status words that is not true on the hardware: a 128-bit
// load of a warp is served in several passes, a store of the producer can land between two of them, and then some
// lanes leave the loop one iteration before the others and the warp never reconverges (found the hard way: the launch
// ran into the spin limit).  Polls of 16-byte words therefore take their exit condition through a warp vote -- all lanes
// leave in the same iteration, every lane has seen the complete word.  RT_VOTE_POLLS selects the loops (bit 1: mbarrier
// wait, 2: look-back word in global memory, 4: carry word in shared memory, 8: the 32-bit ring counters); measured on C4:
// 4 alone +1.6 %, all four +0.4 % against the build with divergence checks.  Default: all four (the build the GPU suite
// was run on); 1|2|4 leaves out the 32-bit counters, whose load is a single broadcast access.
#ifndef RT_VOTE_POLLS
#define RT_VOTE_POLLS 15
#endif
#ifndef RT_SPIN_SHIFT
#define RT_SPIN_SHIFT 22 // log2 of the spin limit of the loops below (a diagnostic build with 40 hangs instead of trapping)
#endif

namespace rtk
{
template <int BIT>
__device__ __forceinline__ bool poll_again(bool c)
{
   if (RT_VOTE_POLLS & BIT) return __any_sync(0xffffffffu, c);
   return c;
}
constexpr long long SPIN_LIMIT_W = 1ll << RT_SPIN_SHIFT;
constexpr long long SPIN_LIMIT_S = 1ll << (RT_SPIN_SHIFT + 3); // shared-memory polls turn around ~8x quicker than global ones
__device__ __forceinline__ void mbar_wait_w(unsigned long long *bar, unsigned parity)
{
   long long spins = 0;
   while (poll_again<1>(!mbar_try_wait(bar, parity)))
      if (++spins > SPIN_LIMIT_W) __trap();
}
// packed per-row constants of a BDF2 step: 64 doubles.  The per-cell right-hand sides are FOLDED: with
//   ru = q + c_uu x + c_uv y + p_uu X + p_uv Y + up1c up1 + up2c up2 ,  rv = q - c_uv x + c_uu y - p_uv X + p_uu Y ,
//   beta = io ru + id rv ,  U = id ru - io rv      (x, y = X1 nodes ; X, Y = X2 nodes ; q = qa + qb (u+v))
// (the formulas of stage_pass1, sweep_device.cuh) beta and U are linear in the cell's inputs, so their coefficients are
// formed once per row:
//   B0 = (io+id) qa, Bq = (io+id) qb, Bx = io c_uu - id c_uv, By = io c_uv + id c_uu, BX = io p_uu - id p_uv,
//   BY = io p_uv + id p_uu, Bup1 = io up1c, Bup2 = io up2c ;  U likewise with (id, -io) in place of (io, id).
// In stages 0..2 the source cell IS the X1 cell (u+v = x+y), so Bq / Uq are folded into Bx, By / Ux, Uy as well:
// a BE cell costs 2 + 2 fused multiply-adds instead of 9 operations.  *_POW: alpha^1 .. alpha^8 of the kind (pass 2,
// the scan factor alpha^CPT and alpha itself come from here: no power chains on the critical path).
enum StepCoefIndex
{
   SC_BE = 0,       // gamma inflow0 | B0 Bx By | U0 Ux Uy
   SC_BE_POW = 8,
   SC_CN = 16,      // gamma inflow0 | B0 Bx By Bup2 | U0 Ux Uy Uup2
   SC_CN_POW = 26,
   SC_BDF = 34,     // gamma inflow0 | B0 Bx By Bq BX BY Bup1 Bup2 | U0 Ux Uy Uq UX UY Uup1 Uup2
   SC_BDF_POW = 52,
   SC_LOCAL = 60,   // 1.0 when alpha^16 < 2^-70 for all three kinds: lanes two or more apart do not feel each other (CPT = 8)
   SC_DOUBLES = STEP_COEF_DOUBLES
};
static_assert(SC_BDF + 18 <= SC_BDF_POW && SC_BDF_POW + 8 <= SC_LOCAL && SC_LOCAL < SC_DOUBLES, "packed entry layout");

__global__ void pack_step_coef_kernel(const RowCoef *__restrict__ be, const RowCoef *__restrict__ cn,
                                      const RowCoef *__restrict__ bdf, double *__restrict__ out, int rows, int corr)
{
   const int r = blockIdx.x * blockDim.x + threadIdx.x;
   if (r >= rows) return;
   double *o = out + (size_t)r * SC_DOUBLES;
   for (int i = 0; i < SC_DOUBLES; ++i) o[i] = 0.0;
   // kind: 0 BE, 1 CN, 2 BDF ; entries that are structurally absent for a kind are forced to zero (cf. fold_coef)
   for (int kind = 0; kind < 3; ++kind)
   {
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
         double *k = o + SC_BE;
         k[0] = c.gamma; k[1] = c.inflow0;
         k[2] = B0; k[3] = Bx + Bq; k[4] = By + Bq;
         k[5] = U0; k[6] = Ux + Uq; k[7] = Uy + Uq;
      }
      else if (kind == 1)
      {
         double *k = o + SC_CN;
         k[0] = c.gamma; k[1] = c.inflow0;
         k[2] = B0; k[3] = Bx + Bq; k[4] = By + Bq; k[5] = io * u2;
         k[6] = U0; k[7] = Ux + Uq; k[8] = Uy + Uq; k[9] = id * u2;
      }
      else
      {
         double *k = o + SC_BDF;
         k[0] = c.gamma; k[1] = c.inflow0;
         k[2] = B0; k[3] = Bx; k[4] = By; k[5] = Bq; k[6] = io * puu - id * puv; k[7] = io * puv + id * puu;
         k[8] = io * u1; k[9] = io * u2;
         k[10] = U0; k[11] = Ux; k[12] = Uy; k[13] = Uq; k[14] = id * puu + io * puv; k[15] = id * puv - io * puu;
         k[16] = id * u1; k[17] = id * u2;
      }
      double *pw = o + (kind == 0 ? SC_BE_POW : (kind == 1 ? SC_CN_POW : SC_BDF_POW));
      double pa = 1.0;
      for (int j = 0; j < 8; ++j)
      {
         pa *= c.alpha;
         pw[j] = pa;
      }
      // Local regime (small time steps: alpha ~ 0.01 at the reference's own dt = 1e-5).  The warp scan's step s folds in
      // the lane 2^s below with the factor alpha^(8 2^s); once that is below 2^-70 the step cannot change a result by more
      // than 2^-70 of a neighbouring value, 20 orders of magnitude inside the tolerance: such rows run step 0 only.
      if (kind == 0) o[SC_LOCAL] = 1.0;
      if (!(pa * pa < 0x1p-70)) o[SC_LOCAL] = 0.0; // pa = alpha^8 here
   }
}

void launch_pack_step_coef(const double *be, const double *cn, const double *bdf, double *out, int rows, bool corr, cudaStream_t st)
{
   pack_step_coef_kernel<<<(rows + 127) / 128, 128, 0, st>>>(reinterpret_cast<const RowCoef *>(be), reinterpret_cast<const RowCoef *>(cn),
                                                            reinterpret_cast<const RowCoef *>(bdf), out, rows, corr ? 1 : 0);
}


template <bool CORR, int CPT, int NW>
__global__ void __launch_bounds__(32 * NW, 16 / NW) step_warp_kernel(const __grid_constant__ SweepMaps maps, const StepArgs a)
{
   constexpr int WT = 32 * CPT, WT_BYTES = WT * 16; // one warp's sub-tile
   constexpr int TILE = WT * NW;                    // the CTA's tile: NW consecutive sub-tiles of one row
   static_assert(CPT == 8 || CPT == 4, "the swizzled tile image is conflict-free for 4 or 8 cells per thread");

   // dynamic shared memory, 1024-byte aligned by hand: per warp [X slot 0][X slot 1][H / OUT]
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   __shared__ __align__(16) double s_coef[NW][2][SC_DOUBLES];
   __shared__ __align__(16) double2 s_halo[NW][2];     // X at the cell upstream of the sub-tile
   __shared__ __align__(16) double2 s_carry[4][4][NW]; // [tile sequence & 3][stage][warp]: inclusive value of a sub-tile
   __shared__ __align__(16) WarpTile s_desc[16];
   __shared__ __align__(8) unsigned long long s_full[NW][2];
   __shared__ int s_decoded; // ring entries decoded by warp 0 so far
   __shared__ int s_done[NW]; // tiles each warp has finished or sat out (a row may end before a warp's sub-tile)

   // The warp index goes through a shuffle so that ptxas KNOWS it is warp-uniform: with `threadIdx.x >> 5` every branch on
   // it counts as divergent, each shuffle below gets a BRA.DIV check with an out-of-line WARPSYNC.COLLECTIVE path, and the
   // register constraints of those paths cost two moves per 64-bit shuffle in the main path (-30 % instructions, SASS).
   const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;
   const int half_m = a.M >> 1, dm = a.reflective_left ? half_m : a.M;
   const int batch = a.ticket_batch; // <= 8 (launch_step_warp)
   unsigned char *const wt = tiles + warp * 3 * WT_BYTES;
   unsigned char *const th = wt + 2 * WT_BYTES;

   // ---- tickets (warp 0): one atomic per batch, the batch decoded by `batch` lanes in parallel into the ring
   long long q_start = -1; // start ticket of the last draw
   bool ended = false;     // a ticket past the end has been seen: no more draws
   int dec = 0;            // ring entries decoded so far (sequence numbers 0 .. dec-1)
   auto draw = [&]() {
      unsigned long long t0 = 0;
      if (lane == 0) t0 = atomicAdd(a.ticket, 1ull);
      t0 = __shfl_sync(0xffffffffu, t0, 0);
      q_start = (long long)(t0 - a.ticket_base) * batch;
      if (lane < batch)
      {
         const WarpTile d = decode_warp_tile(a, q_start + lane, total_tiles, half_m, dm);
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
   };

   // ---- lane 0 of each warp: TMA / bulk loads of the warp's sub-tile of tile d into its pipeline buffer b
   auto issue_loads = [&](const WarpTile &d, int sub_start, int b) {
      const bool halo = sub_start > 0;
      mbar_expect_tx(&s_full[warp][b], (unsigned)(WT_BYTES + SC_DOUBLES * 8) + (halo ? 16u : 0u));
      tma_load_3d(wt + b * WT_BYTES, &maps.m[a.ix], 0, sub_start >> 3, d.row, &s_full[warp][b]);
      bulk_load(&s_coef[warp][b][0], a.coef_step + (size_t)d.row * SC_DOUBLES, (unsigned)(SC_DOUBLES * 8), &s_full[warp][b]);
      if (halo) bulk_load(&s_halo[warp][b], a.x + (size_t)d.row * a.Np + sub_start - 1, 16u, &s_full[warp][b]);
   };

   if (lane == 0)
   {
      mbar_init(&s_full[warp][0], 1);
      mbar_init(&s_full[warp][1], 1);
      fence_proxy_async();
   }
   if (NW > 1)
   {
      for (int i = threadIdx.x; i < 4 * 4 * NW; i += 32 * NW) (&s_carry[0][0][0])[i] = make_double2(0.0, 0.0);
      if (threadIdx.x == 0) s_decoded = 0;
      if (threadIdx.x < NW) s_done[threadIdx.x] = 0;
      __syncthreads(); // the only CTA-wide barrier
   }
   else __syncwarp();

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
            // back-pressure: at most two tiles ahead of the SLOWEST warp (bounds the carry and descriptor rings).  Every warp
            // counts its own tiles: with one counter written by the last warp, a tile that warp sits out (row ends earlier)
            // let warp 0 run ahead of the warps that were still working and overwrite their carry slots.
            long long spins = 0;
            while (__any_sync(0xffffffffu, ld_flag(&s_done[lane % NW]) < seq - 2))
               if (++spins > SPIN_LIMIT_S) __trap();
         }
      }
      else
      {
         long long spins = 0;
         while (poll_again<8>((avail = ld_flag(&s_decoded)) <= seq))
            if (++spins > SPIN_LIMIT_S) __trap();
         __threadfence_block();
      }
      const WarpTile ti = s_desc[seq & 15];
      if (ti.row < 0) break;
      // warp 0: probe the predecessor tile's four look-back words first of all (lane k holds the word of stage k), so
      // that the L2 round trip overlaps the issue of the next tile's loads below and pass 1 of stage 0
      double2 *const status = a.tile_status + 4 * ((size_t)ti.chain * a.tiles_per_chain + ti.p);
      double2 wpre = make_double2(0.0, 0.0);
      if (warp == 0 && ti.p > 0 && lane < 4) wpre = ld_status(status - 4 + lane);
      __syncwarp(); // every lane is done with the X slot, constants and halo that are refilled now
      while (issued <= seq + 1 && issued < avail)
      {
         if (lane == 0)
         {
            const WarpTile d = s_desc[issued & 15];
            const int sub = (d.tf >> 2) * TILE + warp * WT;
            if (d.row >= 0 && sub < a.N) issue_loads(d, sub, issued & 1);
            if (a.l2_prefetch && issued + 1 < avail)
            {
               const WarpTile da = s_desc[(issued + 1) & 15];
               const int sa = (da.tf >> 2) * TILE + warp * WT;
               if (da.row >= 0 && sa < a.N) prefetch_l2_bulk(a.x + (size_t)da.row * a.Np + sa, (unsigned)min(WT, a.N - sa) * 16u);
            }
         }
         ++issued;
      }
      const int tile_start = (ti.tf >> 2) * TILE, sub_start = tile_start + warp * WT;
      const int valid_total = min(TILE, a.N - tile_start);
      const int valid = min(WT, valid_total - warp * WT); // cells of this warp's sub-tile inside the row (<= 0: none)
      if (valid > 0)
      {
         const int buf = seq & 1;
         const bool sgn = (ti.tf & 2) != 0, junction = (ti.tf & 1) != 0 && warp == 0;
         const bool is_last = valid_total <= (warp + 1) * WT; // the row's tile ends inside (or at the end of) this sub-tile
         const bool first_sub = (warp == 0);                  // carry from the previous CTA tile / the boundary
         double2 *const carry_slot = &s_carry[seq & 3][0][0];

         mbar_wait_w(&s_full[warp][buf], (phase >> buf) & 1u); // sub-tile, constants and halo cell have landed
         phase ^= 1u << buf;

         const unsigned char *const sx = wt + buf * WT_BYTES;
         const double *const cf = s_coef[warp][buf];
         // row-uniform by value; through a shuffle so that ptxas knows it (no divergence handling around the scans' shuffles)
         const bool local_row = CPT == 8 && __shfl_sync(0xffffffffu, (int)(cf[SC_LOCAL] != 0.0), 0) != 0;
         const int j0 = lane * CPT;
         const bool row_first = (sub_start == 0) && (lane == 0);
         auto ldx = [&](int j) { return *reinterpret_cast<const double2 *>(sx + swz(j)); };
         // X outflow node of the cell upstream of this lane's first cell
         double up_x = 0.0;
         if (lane == 0) { if (sub_start > 0) up_x = s_halo[warp][buf].y; }
         else up_x = ldx(j0 - 1).y;

         // ---- carry into the sub-tile for stage k.  Every lane runs the same loads (one broadcast access), so the warp
         // stays converged: no divergent spin, no reconvergence code around the shuffles that follow.
         auto carry_in = [&](auto kc, const double kbc, const double inflow0) -> double {
            constexpr int k = decltype(kc)::value;
            if (first_sub)
            {
               // boundary inflow, or the previous CTA tile's inclusive value (global look-back word)
               double2 w;
               w.x = __shfl_sync(0xffffffffu, wpre.x, k);
               w.y = __shfl_sync(0xffffffffu, wpre.y, k);
               if (ti.p == 0) return inflow0;
               long long spins = 0;
               while (poll_again<2>((int)(status_flag(w) - flag_incl) < 0))
               {
                  w = ld_status(status - 4 + k);
                  if (++spins > SPIN_LIMIT_W) __trap();
               }
               return junction ? kbc * w.x : w.x;
            }
            // the previous warp's inclusive value of this tile and stage, through shared memory
            const double2 *src = carry_slot + k * NW + (warp - 1);
            long long spins = 0;
            // wait on the 8-byte sequence number (one broadcast wavefront per poll instead of two for the 16-byte word:
            // the polls are a measurable share of the shared-memory traffic, profiles/r02_step4_cellmajor_ncu_digest.txt),
            // then read the word -- still checked against its own sequence number below
            const int *seq_word = reinterpret_cast<const int *>(src) + 2;
            while (poll_again<4>(ld_flag(seq_word) != seq + 1))
            {
#if STEP_SPIN_SLEEP > 0
               __nanosleep(STEP_SPIN_SLEEP); // a polling warp burns issue slots and power the working warps could use
#endif
               if (++spins > SPIN_LIMIT_S) __trap();
            }
            double2 w = lds_word(src);
            while (poll_again<4>((int)__double_as_longlong(w.y) != seq + 1))
            {
#if STEP_SPIN_SLEEP > 0
               __nanosleep(STEP_SPIN_SLEEP); // a polling warp competes for issue slots with the warps that do the work
#endif
               w = lds_word(src);
               if (++spins > SPIN_LIMIT_S) __trap();
            }
            return w.x;
         };
         // ---- seeded warp scan of one stage: incl = this lane's chunk total with zero inflow; returns the true inflow
         // of the lane's chunk and publishes the sub-tile's inclusive value (ragged sub-tiles publish after pass 2)
         auto scan = [&](auto kc, const double achunk, const double kbc, const double inflow0, double incl) -> double {
            constexpr int k = decltype(kc)::value;
            double f = achunk; // alpha^CPT
            const double ctile = carry_in(kc, kbc, inflow0);
            const double seeded = fma(f, ctile, incl);
            if (lane == 0) incl = seeded;
            if (local_row)
            {
               // alpha^(2 CPT) < 2^-70: only the neighbouring lane is felt (SC_LOCAL)
               const double o = __shfl_up_sync(0xffffffffu, incl, 1);
               const double t = fma(f, o, incl);
               if (lane >= 1) incl = t;
            }
            else
            {
#pragma unroll
               for (int s = 0; s < 5; ++s)
               {
                  const double o = __shfl_up_sync(0xffffffffu, incl, 1 << s);
                  const double t = fma(f, o, incl);
                  if (lane >= (1 << s)) incl = t;
                  f = f * f;
               }
            }
            double cin = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) cin = ctile;
            if (lane == 31)
            {
               if (!is_last) sts_word(carry_slot + k * NW + warp, incl, seq + 1);
               else if (valid == WT) st_status(status + k, incl, flag_incl);
            }
            return cin;
         };
         // ragged sub-tile: the outflow node of the row's last cell is the chain's inclusive value
         auto publish_ragged = [&](auto kc, const double (&v)[CPT]) {
            constexpr int k = decltype(kc)::value;
            const int last = valid - 1;
            if (lane == last / CPT)
            {
               const int jo = last % CPT;
               double vs = v[0];
#pragma unroll
               for (int jj = 1; jj < CPT; ++jj)
                  if (jj <= jo) vs = v[jj];
               st_status(status + k, vs, flag_incl);
            }
         };

         double wl[CPT], U[CPT], vv[CPT]; // wl: the lane's local scan with zero inflow
         double2 r[CPT];
         double up_h = 0.0;

         // ================= stage 0: BE predictor, x1 = X (ends^n)
         {
            const double *c = cf + SC_BE, *pw = cf + SC_BE_POW;
            const double alpha = pw[0], gamma = c[0];
            double w = 0.0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double2 x = ldx(j0 + jj);
               const double b = fma(c[4], x.y, fma(c[3], x.x, c[2]));
               U[jj] = fma(c[7], x.y, fma(c[6], x.x, c[5]));
               w = fma(alpha, w, b);
               wl[jj] = w;
            }
            const double cin = scan(std::integral_constant<int, 0>{}, pw[CPT - 1], 1.0, c[1], w);
            if (sgn) up_h = cin;
            if (lane == 0) bulk_wait_read<0>(); // the previous tile's TMA store has read H / OUT
            __syncwarp();
            double v = cin;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double vn = fma(pw[jj], cin, wl[jj]);
               r[jj] = make_double2(fma(gamma, v, U[jj]), vn);
               vv[jj] = vn;
               v = vn;
            }
            if (sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = r[jj];
            }
            if (valid != WT) publish_ragged(std::integral_constant<int, 0>{}, vv);
         }
         // ================= stage 1: CN corrector, x1 = stage-0 result, upstream X
         {
            const double *c = cf + SC_CN, *pw = cf + SC_CN_POW;
            const double alpha = pw[0], gamma = c[0];
            double w = 0.0, up2 = row_first ? 0.0 : up_x; // a row's first cell has no upstream term (solver.cpp:695-697)
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double b = fma(c[5], up2, fma(c[4], r[jj].y, fma(c[3], r[jj].x, c[2])));
               U[jj] = fma(c[9], up2, fma(c[8], r[jj].y, fma(c[7], r[jj].x, c[6])));
               w = fma(alpha, w, b);
               wl[jj] = w;
               if (jj + 1 < CPT) up2 = ldx(j0 + jj).y;
            }
            const double cin = scan(std::integral_constant<int, 1>{}, pw[CPT - 1], 2.0, c[1], w);
            if (!sgn) up_h = cin;
            double v = cin;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double vn = fma(pw[jj], cin, wl[jj]);
               r[jj] = make_double2(fma(gamma, v, U[jj]), vn);
               vv[jj] = vn;
               v = vn;
            }
            if (!sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = r[jj];
            }
            if (valid != WT) publish_ragged(std::integral_constant<int, 1>{}, vv);
         }
         // ================= stage 2: BE predictor, x1 = stage-1 result; only u+v survives (the BDF stage's source).
         // Without the correction source (V = 0 or use_correction off: the reference's defaults) nothing reads it: the BDF
         // stage takes half_ends and prev_ends, and `ends` between the stages of a step is not observable from a fused
         // step -- the whole stage, its scan and its look-back chain are left out (3 scans per sub-tile instead of 4).
         double xs[CPT];
         if (CORR)
         {
            const double *c = cf + SC_BE, *pw = cf + SC_BE_POW;
            const double alpha = pw[0], gamma = c[0];
            double w = 0.0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double b = fma(c[4], r[jj].y, fma(c[3], r[jj].x, c[2]));
               U[jj] = fma(c[7], r[jj].y, fma(c[6], r[jj].x, c[5]));
               w = fma(alpha, w, b);
               wl[jj] = w;
            }
            const double cin = scan(std::integral_constant<int, 2>{}, pw[CPT - 1], 1.0, c[1], w);
            double v = cin;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double vn = fma(pw[jj], cin, wl[jj]);
               xs[jj] = fma(gamma, v, U[jj]) + vn;
               vv[jj] = vn;
               v = vn;
            }
            if (valid != WT) publish_ragged(std::integral_constant<int, 2>{}, vv);
         }
         // ================= stage 3: BDF corrector, x1 = half_ends (parked in H), x2 = X -> ends^(n+1)
         {
            const double *c = cf + SC_BDF, *pw = cf + SC_BDF_POW;
            const double alpha = pw[0], gamma = c[0];
            double w = 0.0, up1 = row_first ? 0.0 : up_h, up2 = row_first ? 0.0 : up_x;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double2 h = *reinterpret_cast<const double2 *>(th + swz(j0 + jj));
               const double2 x = ldx(j0 + jj);
               double b = fma(c[4], h.y, fma(c[3], h.x, c[2]));
               double u = fma(c[12], h.y, fma(c[11], h.x, c[10]));
               if (CORR)
               {
                  b = fma(c[5], xs[jj], b);
                  u = fma(c[13], xs[jj], u);
               }
               b = fma(c[9], up2, fma(c[8], up1, fma(c[7], x.y, fma(c[6], x.x, b))));
               u = fma(c[17], up2, fma(c[16], up1, fma(c[15], x.y, fma(c[14], x.x, u))));
               U[jj] = u;
               w = fma(alpha, w, b);
               wl[jj] = w;
               up1 = h.y;
               up2 = x.y;
            }
            const double cin = scan(std::integral_constant<int, 3>{}, pw[CPT - 1], 6.0, c[1], w);
            double v = cin;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double vn = fma(pw[jj], cin, wl[jj]);
               *reinterpret_cast<double2 *>(th + swz(j0 + jj)) = make_double2(fma(gamma, v, U[jj]), vn);
               vv[jj] = vn;
               v = vn;
            }
            if (valid != WT) publish_ragged(std::integral_constant<int, 3>{}, vv);
            fence_proxy_async(); // generic-proxy writes -> visible to the TMA store
            __syncwarp();
            if (lane == 0)
            {
               tma_store_3d(&maps.m[a.iout], 0, sub_start >> 3, ti.row, th); // cells past the row's end are clipped
               bulk_commit();
            }
            __syncwarp(); // as below: no single-lane block at the end of a region
         }
      }
      if (NW > 1)
      {
         __syncwarp();
         if (lane == 0) st_flag(&s_done[warp], seq + 1);
         __syncwarp(); // reconverge BEFORE the back edge (a single-lane block that ends the loop body makes the whole loop divergent for ptxas)
      }
   }
   if (lane == 0) bulk_wait_read<0>();
   if (warp == 0 && lane == 0 && q_start < total_tiles) atomicAdd(a.ticket, 1ull); // every CTA makes exactly one draw past the end
}

template <bool CORR, int CPT, int NW>
static cudaError_t launch_warp_one(const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st, int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   int ctas_per_sm = plan_slot->load(std::memory_order_acquire);
   auto kern = step_warp_kernel<CORR, CPT, NW>;
   const size_t smem = (size_t)NW * 3 * 32 * CPT * 16 + 1024; // per warp 2 X + H/OUT, + alignment slack
   if (ctas_per_sm <= 0)
   {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      int occ = 0;
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 32 * NW, smem);
      if (e != cudaSuccess) return e;
      ctas_per_sm = occ > 0 ? occ : 1;
      plan_slot->store(ctas_per_sm, std::memory_order_release);
   }
   const char *cap = getenv("RT_CUDA_STEP_CTAS"); // resident CTAs per SM (experiments)
   if (cap && atoi(cap) >= 1 && atoi(cap) < ctas_per_sm) ctas_per_sm = atoi(cap);
   const long long total = (long long)a.n_chains * a.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, 32 * NW, smem, st>>>(maps, a);
   return cudaGetLastError();
}

bool step_chain_supported(int nw) { return nw == 1 || nw == 2 || nw == 4 || nw == 8; }

cudaError_t launch_step_warp(bool corr, int cpt, int chain_warps, const SweepMaps &maps, const StepArgs &a, int num_sms, cudaStream_t st,
                             int *grid_out)
{
   if (a.ticket_batch < 1 || a.ticket_batch > 8 || !a.coef_step) return cudaErrorInvalidValue;
#define RT_STEP_CASE(C, P, W) \
   if (corr == C && cpt == P && chain_warps == W) return launch_warp_one<C, P, W>(maps, a, num_sms, st, grid_out);
   RT_STEP_CASE(true, 8, 1) RT_STEP_CASE(true, 8, 2) RT_STEP_CASE(true, 8, 4) RT_STEP_CASE(true, 8, 8)
   RT_STEP_CASE(false, 8, 1) RT_STEP_CASE(false, 8, 2) RT_STEP_CASE(false, 8, 4) RT_STEP_CASE(false, 8, 8)
   RT_STEP_CASE(true, 4, 1) RT_STEP_CASE(true, 4, 2) RT_STEP_CASE(true, 4, 4) RT_STEP_CASE(true, 4, 8)
   RT_STEP_CASE(false, 4, 1) RT_STEP_CASE(false, 4, 2) RT_STEP_CASE(false, 4, 4) RT_STEP_CASE(false, 4, 8)
#undef RT_STEP_CASE
   return cudaErrorInvalidValue;
}
} // namespace rtk
