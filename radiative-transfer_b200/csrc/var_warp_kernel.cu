// var_warp_kernel.cu -- the fused BDF2 step for CELL-DEPENDENT material fields (var_step_kernel.cu: same stage algebra, same
// four look-back chains per tile) built from WARP tiles like the uniform step (step_warp_kernel.cu).  EXTENSION (SURVEY
// 8(f).2 / 8(f).4); oracle: the cell-dependent restatement in oracle/rt_oracle.cpp.
//
// Profile of the CTA-wide version (profiles/r01_v9_var_step_ncu_digest.txt): 12 CTA barriers per tile (12 % of the stall
// samples, and behind them the look-back that only thread 0 does), ~330 thread instructions per cell and step of which
// 145 are fp64 -- the fp64 pipe alone bounds this path at ~0.64 of the HBM peak.  This version
//   * has no CTA barrier in its tile loop: a warp owns a sub-tile of 32 lanes x 4 cells (2 KB of state, 1 KB per material
//     table) with its own TMA pipeline, the NW warps of a CTA take NW consecutive sub-tiles of one row and pass the
//     recurrence carry from warp to warp through shared memory ({value, sequence} words); only warp 0 looks back to the
//     previous CTA tile through global memory;
//   * runs a SEEDED pair scan: the carry is folded into lane 0's chunk before the warp scan of the pairs
//     (A = product of the chunk's alphas, W = its zero-inflow outflow), so the scan delivers true values and there is no
//     fix-up pass; the A products of the BE kind are scanned once and serve stages 0 and 2;
//   * pass 2 is parallel over the lane's cells: v_j = P_j cin + w_j with the running products P_j and local values w_j
//     that pass 1 forms anyway (no serial chain);
//   * forms 1/det of the BE kind ONCE per cell for stages 0 and 2 (the reciprocal is the expensive part of the per-cell
//     constants) and re-reads X from shared memory instead of holding it in registers;
//   * parks only the stage result that is this row's half_ends (stage 0 for mu>0 rows, stage 1 for mu<0 rows),
//     thread-private; the stage-3 result overwrites it in place and leaves by TMA store;
//   * cells past the end of a row are the identity element (alpha = 1, beta = 0), so a ragged sub-tile publishes its
//     inclusive value like a full one.
// Convergence rules as in step_warp_kernel.cu: warp index through a shuffle, polls of 16-byte words leave through a vote,
// single-lane blocks are followed by __syncwarp before a back edge.
#include <atomic>
#include <cstdlib>
#include <type_traits>

#include "rt_common.cuh"
#include "step_kernel.cuh"
#include "sweep_device.cuh"
#include "warp_tile.cuh"
#include "sweep_kernel.cuh"
#include "var_device.cuh"
#include "var_kernel.cuh"

#ifndef VAR_INV_REUSE
#define VAR_INV_REUSE 1 // 1/det of the BE kind from stage 0 serves stage 2 (+3 % in this kernel, -3 % in the CTA-wide one)
#endif
#ifndef VAR_WARP_REGS
#define VAR_WARP_REGS 168 // register budget per thread the resident-CTA target is derived from
#endif
#ifndef RT_SPIN_SHIFT
#define RT_SPIN_SHIFT 22
#endif

namespace rtk
{
namespace vw
{
constexpr long long SPIN_W = 1ll << RT_SPIN_SHIFT, SPIN_S = 1ll << (RT_SPIN_SHIFT + 3);
__device__ __forceinline__ void mbar_wait_vote(unsigned long long *bar, unsigned parity)
{
   long long spins = 0;
   while (__any_sync(0xffffffffu, !mbar_try_wait(bar, parity)))
      if (++spins > SPIN_W) __trap();
}
constexpr int CPT = 4, WT = 32 * CPT, WT_BYTES = WT * 16, MAT_BYTES = WT * 8;
template <bool CORR>
constexpr int warp_smem_bytes()
{
   return 3 * WT_BYTES + (CORR ? 5 : 2) * MAT_BYTES; // 2 X + H/OUT + one slot of material tables: 11 KB / 8 KB
}
template <bool CORR, int NW>
constexpr int min_blocks()
{
   const int by_smem = (227 * 1024) / (NW * warp_smem_bytes<CORR>() + 3072), by_regs = 65536 / (32 * NW * VAR_WARP_REGS);
   return by_smem < by_regs ? by_smem : by_regs;
}
} // namespace vw

template <bool CORR, int NW>
__global__ void __launch_bounds__(32 * NW, vw::min_blocks<CORR, NW>())
var_warp_kernel(const __grid_constant__ SweepMaps maps, const __grid_constant__ MatMaps mats, const VarStepArgs va)
{
   using namespace vw;
   constexpr int TILE = WT * NW; // the CTA's tile: NW consecutive sub-tiles of one row
   constexpr int NT = CORR ? 5 : 2;
   const StepArgs &a = va.s;

   // dynamic shared memory, 1024-byte aligned by hand: per warp [X slot 0][X slot 1][H / OUT][NT material tables]
   extern __shared__ unsigned char smem_raw[];
   unsigned char *const tiles = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
   __shared__ __align__(16) double2 s_halo[NW][2];     // X at the cell upstream of the sub-tile
   __shared__ __align__(16) double2 s_carry[4][4][NW]; // [tile sequence & 3][stage][warp]: inclusive value of a sub-tile
   __shared__ __align__(16) WarpTile s_desc[16];
   __shared__ __align__(8) unsigned long long s_full[NW][2], s_mfull[NW];
   __shared__ int s_decoded;  // ring entries decoded by warp 0 so far
   __shared__ int s_done[NW]; // tiles each warp has finished or sat out

   const int lane = threadIdx.x & 31, warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
   const unsigned flag_incl = a.epoch * 2u + 2u;
   const long long total_tiles = (long long)a.n_chains * a.tiles_per_chain;
   const int half_m = a.M >> 1, dm = a.reflective_left ? half_m : a.M;
   const int batch = a.ticket_batch; // <= 8
   constexpr int WARP_SMEM = 3 * WT_BYTES + (CORR ? 5 : 2) * MAT_BYTES; // = warp_smem_bytes<CORR>()
   unsigned char *const wt = tiles + warp * WARP_SMEM;
   unsigned char *const th = wt + 2 * WT_BYTES;
   unsigned char *const tm = wt + 3 * WT_BYTES;

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
         const WarpTile d = decode_warp_tile(a, q_start + lane, total_tiles, half_m, dm);
         s_desc[(dec + lane) & 15] = d;
      }
      if (q_start + batch > total_tiles) ended = true;
      dec += batch;
      __syncwarp();
      if (lane == 0)
      {
         __threadfence_block();
         st_flag(&s_decoded, dec);
      }
   };

   // ---- lane 0 of each warp: loads of the warp's sub-tile of tile d (X into pipeline buffer b; the material tables have
   // ONE slot, refilled as soon as the tile before has turned them into registers)
   auto issue_x = [&](const WarpTile &d, int sub_start, int b) {
      const bool halo = sub_start > 0;
      mbar_expect_tx(&s_full[warp][b], (unsigned)WT_BYTES + (halo ? 16u : 0u));
      tma_load_3d(wt + b * WT_BYTES, &maps.m[a.ix], 0, sub_start >> 3, d.row, &s_full[warp][b]);
      if (halo) bulk_load(&s_halo[warp][b], a.x + (size_t)d.row * a.Np + sub_start - 1, 16u, &s_full[warp][b]);
   };
   auto issue_mat = [&](const WarpTile &d, int sub_start) {
      const int gl = d.row / a.M, trow = ((d.tf >> 1) & 1) * va.Gl + gl; // the table copy in this row's sweep order
      mbar_expect_tx(&s_mfull[warp], (unsigned)(NT * MAT_BYTES));
#pragma unroll
      for (int t = 0; t < NT; ++t) tma_load_3d(tm + t * MAT_BYTES, &mats.m[t], 0, sub_start >> 4, trow, &s_mfull[warp]);
   };

   if (lane == 0)
   {
      mbar_init(&s_full[warp][0], 1);
      mbar_init(&s_full[warp][1], 1);
      mbar_init(&s_mfull[warp], 1);
      fence_proxy_async();
   }
   for (int i = threadIdx.x; i < 4 * 4 * NW; i += 32 * NW) (&s_carry[0][0][0])[i] = make_double2(0.0, 0.0);
   if (threadIdx.x == 0) s_decoded = 0;
   if (threadIdx.x < NW) s_done[threadIdx.x] = 0;
   __syncthreads(); // the only CTA-wide barrier

   // byte offsets of the lane's cells in the swizzled tile images: lane constants, formed once (the shift / xor chains in
   // front of every shared-memory access were 4 % of the instructions and sat on the critical path of the loads)
   const int j0 = lane * CPT;
   int ox[CPT], om[CPT / 2];
#pragma unroll
   for (int jj = 0; jj < CPT; ++jj) ox[jj] = swz(j0 + jj);
#pragma unroll
   for (int pp = 0; pp < CPT / 2; ++pp) om[pp] = swz((j0 >> 1) + pp);
   const int oup = swz(lane > 0 ? j0 - 1 : 0);

   unsigned phase = 0, mphase = 0;
   int issued = 0;   // tiles whose X loads this warp has issued (or skipped: sub-tile past the end of the row)
   int mat_for = -1; // tile sequence number whose material tables are in (or on their way into) the warp's slot
   for (int seq = 0;; ++seq)
   {
      int avail;
      if (warp == 0)
      {
         while (!ended && dec < seq + 3) draw();
         avail = dec;
         // back-pressure: at most two tiles ahead of the SLOWEST warp (bounds the carry and descriptor rings); every warp
         // counts its own tiles (a warp may sit a tile out)
         long long spins = 0;
         while (__any_sync(0xffffffffu, ld_flag(&s_done[lane % NW]) < seq - 2))
            if (++spins > SPIN_S) __trap();
      }
      else
      {
         long long spins = 0;
         while (__any_sync(0xffffffffu, (avail = ld_flag(&s_decoded)) <= seq))
            if (++spins > SPIN_S) __trap();
         __threadfence_block();
      }
      const WarpTile ti = s_desc[seq & 15];
      if (ti.row < 0) break;
      // warp 0: probe the predecessor tile's four look-back words first of all (lane k holds the word of stage k)
      double2 *const status = a.tile_status + 4 * ((size_t)ti.chain * a.tiles_per_chain + ti.p);
      double2 wpre = make_double2(0.0, 0.0);
      if (warp == 0 && ti.p > 0 && lane < 4) wpre = ld_status(status - 4 + lane);
      __syncwarp(); // every lane is done with the X slot and halo that are refilled now
      while (issued <= seq + 1 && issued < avail)
      {
         if (lane == 0)
         {
            const WarpTile d = s_desc[issued & 15];
            const int sub = (d.tf >> 2) * TILE + warp * WT;
            if (d.row >= 0 && sub < a.N) issue_x(d, sub, issued & 1);
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
         const bool is_last = valid_total <= (warp + 1) * WT;
         const bool first_sub = (warp == 0);
         const bool full = valid == WT;
         double2 *const carry_slot = &s_carry[seq & 3][0][0];
         const int gl = ti.row / a.M, dir = ti.row - gl * a.M;
         const double mu = va.mu[dir];

         if (mat_for != seq) // first tile, or the tile before was sat out
         {
            if (lane == 0) issue_mat(ti, sub_start);
            mat_for = seq;
            __syncwarp();
         }
         mbar_wait_vote(&s_full[warp][buf], (phase >> buf) & 1u); // sub-tile and halo cell have landed
         phase ^= 1u << buf;
         mbar_wait_vote(&s_mfull[warp], mphase); // and the material tables
         mphase ^= 1u;

         const unsigned char *const sx = wt + buf * WT_BYTES;
         const bool row_first = (sub_start == 0) && (lane == 0);
         auto ldx = [&](int jj) { return *reinterpret_cast<const double2 *>(sx + ox[jj]); };
         double up_x = 0.0; // X outflow node of the cell upstream of this lane's first cell
         if (lane == 0) { if (sub_start > 0) up_x = s_halo[warp][buf].y; }
         else up_x = reinterpret_cast<const double2 *>(sx + oup)->y;

         // the cell's material enters every stage through sigma and the source terms qa, qb only, and those do not depend
         // on the stage kind (same tau = dt/2): read the staged tables ONCE per tile, keep 3 values per cell
         double sigc[CPT], qac[CPT], qbc[CPT];
         {
            const RowConst<KIND_BE, CORR> rq(mu, va.dx, va.tau, va.dt, va.beta);
#pragma unroll
            for (int pp = 0; pp < CPT / 2; ++pp)
            {
               double2 m[5];
#pragma unroll
               for (int t = 0; t < 5; ++t)
                  m[t] = (t < NT) ? *reinterpret_cast<const double2 *>(tm + t * MAT_BYTES + om[pp]) : make_double2(0.0, 0.0);
               sigc[2 * pp] = m[0].x; sigc[2 * pp + 1] = m[0].y;
               rq.source_terms(m[1].x, m[2].x, m[3].x, m[4].x, qac[2 * pp], qbc[2 * pp]);
               rq.source_terms(m[1].y, m[2].y, m[3].y, m[4].y, qac[2 * pp + 1], qbc[2 * pp + 1]);
            }
         }
         __syncwarp(); // every lane has read the tables: the slot is free for the next tile's
         if (seq + 1 < avail)
         {
            const WarpTile dn = s_desc[(seq + 1) & 15];
            const int subn = (dn.tf >> 2) * TILE + warp * WT;
            if (dn.row >= 0 && subn < a.N)
            {
               if (lane == 0) issue_mat(dn, subn);
               mat_for = seq + 1;
               __syncwarp();
            }
         }

         // ---- carry into the sub-tile for stage k (every lane runs the same loads: the warp stays converged)
         auto carry_in = [&](auto kc, const double kbc) -> double {
            constexpr int k = decltype(kc)::value;
            if (first_sub)
            {
               double2 w;
               w.x = __shfl_sync(0xffffffffu, wpre.x, k);
               w.y = __shfl_sync(0xffffffffu, wpre.y, k);
               if (ti.p == 0) return kbc * va.bdry[(size_t)dir + (size_t)a.M * (va.G0 + gl)];
               long long spins = 0;
               while (__any_sync(0xffffffffu, (int)(status_flag(w) - flag_incl) < 0))
               {
                  w = ld_status(status - 4 + k);
                  if (++spins > SPIN_W) __trap();
               }
               return junction ? kbc * w.x : w.x;
            }
            const double2 *src = carry_slot + k * NW + (warp - 1);
            const int *seq_word = reinterpret_cast<const int *>(src) + 2;
            long long spins = 0;
            while (__any_sync(0xffffffffu, ld_flag(seq_word) != seq + 1))
               if (++spins > SPIN_S) __trap();
            double2 w = lds_word(src);
            while (__any_sync(0xffffffffu, (int)__double_as_longlong(w.y) != seq + 1))
            {
               w = lds_word(src);
               if (++spins > SPIN_S) __trap();
            }
            return w.x;
         };
         // ---- seeded warp scan of the pairs of one stage: (A, W) = product of the lane's alphas and its outflow with zero
         // inflow; returns the true inflow of the lane's chunk and publishes the sub-tile's inclusive value
         auto scan = [&](auto kc, const double kbc, double A, double W) -> double {
            constexpr int k = decltype(kc)::value;
            const double ctile = carry_in(kc, kbc);
            const double seeded = fma(A, ctile, W);
            if (lane == 0) W = seeded;
#pragma unroll
            for (int s = 0; s < 5; ++s)
            {
               const double Wu = __shfl_up_sync(0xffffffffu, W, 1 << s);
               const double Wn = fma(A, Wu, W);
               if (s < 4)
               {
                  const double Au = __shfl_up_sync(0xffffffffu, A, 1 << s);
                  const double An = A * Au;
                  if (lane >= (1 << s)) A = An;
               }
               if (lane >= (1 << s)) W = Wn;
            }
            double cin = __shfl_up_sync(0xffffffffu, W, 1);
            if (lane == 0) cin = ctile;
            if (lane == 31)
            {
               if (!is_last) sts_word(carry_slot + k * NW + warp, W, seq + 1);
               else st_status(status + k, W, flag_incl);
            }
            return cin;
         };

         // ---- pass 1 of one stage: per-cell constants and right-hand sides; the lane's pair (A, W) and, per cell, the
         // running product P_j, the local value w_j (zero inflow), gamma_j and U_j for pass 2.  Cells past the end of the
         // row are the identity element.  REUSE: 1/det of the BE kind comes from stage 0.
         double ap[CPT], wl[CPT], gam[CPT], U[CPT], invbe[CPT];
         auto pass1m = [&](auto kindc, auto maskc, auto reusec, const double2 (&x1)[CPT], const double2 (&x2)[CPT], const double (&xs)[CPT],
                           double up1, double up2, double &A, double &W) {
            constexpr int KIND = decltype(kindc)::value;
            constexpr bool MASKED = decltype(maskc)::value, REUSE = decltype(reusec)::value && VAR_INV_REUSE;
            const RowConst<KIND, CORR> rc(mu, va.dx, va.tau, va.dt, va.beta);
            A = 1.0; W = 0.0;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double inv = (KIND == KIND_BE && REUSE) ? invbe[jj] : rc.inv_det(sigc[jj]);
               if (KIND == KIND_BE && !REUSE) invbe[jj] = inv;
               const RowCoef c = rc.cell_from_inv(sigc[jj], inv, qac[jj], qbc[jj]);
               double2 X1[1] = {x1[jj]}, X2[1] = {x2[jj]};
               double XS[1] = {xs[jj]}, w1[1], Uc[1];
               stage_pass1<KIND, CORR, 1>(c, X1, X2, XS, up1, up2, row_first && jj == 0, w1, Uc);
               const bool ok = !MASKED || (j0 + jj) < valid;
               const double al = ok ? c.alpha : 1.0;
               gam[jj] = ok ? c.gamma : 0.0;
               U[jj] = ok ? Uc[0] : 0.0;
               W = fma(al, W, ok ? w1[0] : 0.0);
               A *= al;
               ap[jj] = A;
               wl[jj] = W;
               if (KIND == KIND_BDF) up1 = x1[jj].y;
               if (KIND != KIND_BE) up2 = x2[jj].y;
            }
         };
         auto pass1 = [&](auto kindc, auto reusec, const double2 (&x1)[CPT], const double2 (&x2)[CPT], const double (&xs)[CPT], double up1,
                          double up2, double &A, double &W) {
            if (full) pass1m(kindc, std::false_type{}, reusec, x1, x2, xs, up1, up2, A, W);
            else pass1m(kindc, std::true_type{}, reusec, x1, x2, xs, up1, up2, A, W);
         };
         // ---- pass 2: true values of the lane's cells from the true inflow of its chunk
         auto pass2 = [&](const double cin, double2 (&r)[CPT]) {
            double v = cin;
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               const double vn = fma(ap[jj], cin, wl[jj]);
               r[jj] = make_double2(fma(gam[jj], v, U[jj]), vn);
               v = vn;
            }
         };

         double2 r[CPT], xr[CPT];
         double xsum[CPT], A, W, up_h = 0.0;
         using KBE = std::integral_constant<int, KIND_BE>;
         using KCN = std::integral_constant<int, KIND_CN>;
         using KBDF = std::integral_constant<int, KIND_BDF>;

         // ================= stage 0: BE predictor, x1 = X (ends^n)
         {
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               xr[jj] = ldx(jj);
               xsum[jj] = xr[jj].x + xr[jj].y;
            }
            pass1(KBE{}, std::false_type{}, xr, xr, xsum, 0.0, 0.0, A, W);
            const double cin = scan(std::integral_constant<int, 0>{}, 1.0, A, W);
            if (sgn) up_h = cin;
            if (lane == 0) bulk_wait_read<0>(); // the previous tile's TMA store has read H / OUT
            __syncwarp();
            pass2(cin, r);
            if (sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + ox[jj]) = r[jj];
            }
         }
         // ================= stage 1: CN corrector, x1 = stage-0 result, upstream X
         {
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               xr[jj] = ldx(jj);
               xsum[jj] = r[jj].x + r[jj].y;
            }
            pass1(KCN{}, std::false_type{}, r, xr, xsum, 0.0, up_x, A, W);
            const double cin = scan(std::integral_constant<int, 1>{}, 2.0, A, W);
            if (!sgn) up_h = cin;
            pass2(cin, r);
            if (!sgn)
            {
#pragma unroll
               for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + ox[jj]) = r[jj];
            }
         }
         // ================= stage 2: BE predictor, x1 = stage-1 result; only u+v survives (the BDF stage's source): left out
         // when there is no correction source to read it
         if (CORR)
         {
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
            pass1(KBE{}, std::true_type{}, r, r, xsum, 0.0, 0.0, A, W);
            const double cin = scan(std::integral_constant<int, 2>{}, 1.0, A, W);
            pass2(cin, r);
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj) xsum[jj] = r[jj].x + r[jj].y;
         }
         // ================= stage 3: BDF corrector, x1 = half_ends (parked in H), x2 = X -> ends^(n+1)
         {
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj)
            {
               r[jj] = *reinterpret_cast<const double2 *>(th + ox[jj]);
               xr[jj] = ldx(jj);
            }
            pass1(KBDF{}, std::false_type{}, r, xr, xsum, up_h, up_x, A, W);
            const double cin = scan(std::integral_constant<int, 3>{}, 6.0, A, W);
            pass2(cin, r);
#pragma unroll
            for (int jj = 0; jj < CPT; ++jj) *reinterpret_cast<double2 *>(th + ox[jj]) = r[jj];
            fence_proxy_async(); // generic-proxy writes -> visible to the TMA store
            __syncwarp();
            if (lane == 0)
            {
               tma_store_3d(&maps.m[a.iout], 0, sub_start >> 3, ti.row, th); // cells past the row's end are clipped
               bulk_commit();
            }
            __syncwarp();
         }
      }
      __syncwarp();
      if (lane == 0) st_flag(&s_done[warp], seq + 1);
      __syncwarp(); // reconverge BEFORE the back edge
   }
   if (lane == 0) bulk_wait_read<0>();
   if (warp == 0 && lane == 0 && q_start < total_tiles) atomicAdd(a.ticket, 1ull); // every CTA makes exactly one draw past the end
}

// ------------------------------------------------------------------------------------------ host side
template <bool CORR, int NW>
static cudaError_t var_warp_launch_one(const SweepMaps &maps, const MatMaps &mats, const VarStepArgs &va, int num_sms, cudaStream_t st,
                                       int *grid_out)
{
   static DevicePlanCache plans;
   std::atomic<int> *plan_slot = nullptr;
   if (cudaError_t e0 = device_plan_slot(plans, &plan_slot)) return e0;
   int ctas_per_sm = plan_slot->load(std::memory_order_acquire);
   auto kern = var_warp_kernel<CORR, NW>;
   const size_t smem = (size_t)NW * vw::warp_smem_bytes<CORR>() + 1024; // + alignment slack
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
   const long long total = (long long)va.s.n_chains * va.s.tiles_per_chain;
   long long grid = (long long)num_sms * ctas_per_sm;
   if (grid > total) grid = total;
   if (grid < 1) grid = 1;
   if (grid_out) *grid_out = (int)grid;
   kern<<<(unsigned)grid, 32 * NW, smem, st>>>(maps, mats, va);
   return cudaGetLastError();
}

bool var_warp_chain_supported(int nw) { return nw == 2 || nw == 4 || nw == 8; }
int var_warp_subtile_cells() { return vw::WT; }

cudaError_t launch_var_warp(bool corr, int chain_warps, const SweepMaps &maps, const MatMaps &mats, const VarStepArgs &va, int num_sms,
                            cudaStream_t st, int *grid_out)
{
   if (va.s.ticket_batch < 1 || va.s.ticket_batch > 8) return cudaErrorInvalidValue;
#define RT_VW_CASE(C, W) \
   if (corr == C && chain_warps == W) return var_warp_launch_one<C, W>(maps, mats, va, num_sms, st, grid_out);
   RT_VW_CASE(true, 2) RT_VW_CASE(true, 4) RT_VW_CASE(true, 8) RT_VW_CASE(false, 2) RT_VW_CASE(false, 4) RT_VW_CASE(false, 8)
#undef RT_VW_CASE
   return cudaErrorInvalidValue;
}
} // namespace rtk
