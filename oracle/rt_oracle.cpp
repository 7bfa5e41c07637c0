// oracle/rt_oracle.cpp -- TEST INFRASTRUCTURE (oracle-2).
//
// Plain C++ CPU restatement of the reference's time-dependent multigroup S_N slab solve
// (reference src/solver.cpp, src/correction.cpp, src/Planck.cpp, src/GLQuad.cpp,
// include/Constants.h).  It exists so that parity tests can (a) snapshot the state after any
// number of stage sweeps -- the reference has no per-step hook -- and (b) run at sizes where the
// reference's per-cell `half_ends = ends` copy (O(N^2)) is intractable.
//
// Parity status: PINNED.  Checked bit-for-bit against the unmodified reference compiled in place
// (oracle/_ref/ref_dump) by tests/golden/make_golden.py and tests/test_oracle_vs_golden.py.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use this library; the product path (librt_cuda.so) never routes through it.
#include "rt_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "planck_port.inc"
#include "correction_port.inc"
#include "sweep_port.inc"

struct rto_solver
{
   rto_params p;
   std::vector<double> mu, wt;
   rto::GaussRule gauss;
   rto::GroupTerms gt;
   rto::Sweep sw;
   int stages_done;
   bool var;            // EXTENSION: cell-dependent material fields
   rto::CellTables ct;
   rto_solver() : gauss(rto::PLANCK_ACCURACY), stages_done(0), var(false) {}
};

// per-cell group terms (extension): the reference's G-length Correction pipeline evaluated with the cell's rho, T, kappa
static void rto_cell_tables(rto_solver *s)
{
   const int N = s->p.N, G = s->p.G;
   rto::CellTables &ct = s->ct;
   const size_t n = (size_t)G * N;
   ct.B.assign(n, 0.0); ct.dBdT.assign(n, 0.0); ct.dEB.assign(n, 0.0); ct.cor1.assign(n, 0.0); ct.cor2.assign(n, 0.0); ct.cor3.assign(n, 0.0);
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(s->p.num_threads > 0 ? s->p.num_threads : 1)
#endif
   for (int c = 0; c < N; c++)
   {
      rto::GroupTerms g;
      g.setup(G, s->gt.e_edge.data(), &ct.kappa[(size_t)G * c], ct.rho[c]);
      g.per_stage(s->gauss, ct.T[c]);
      for (int k = 0; k < G; k++)
      {
         const size_t o = (size_t)k + (size_t)G * c;
         ct.B[o] = g.B[k]; ct.dBdT[o] = g.dBdT[k]; ct.dEB[o] = g.dEB[k]; ct.cor1[o] = g.cor1[k]; ct.cor2[o] = g.cor2[k]; ct.cor3[o] = g.cor3[k];
      }
   }
}

// EXTENSION, PARITY UNPINNED -- the reference has no material-temperature update; this
// restatement is the definition the device path is checked against (see rto_params::cv in rt_oracle.h).
static void rto_temperature_update(rto_solver *s, bool first_step)
{
   const rto_params &p = s->p;
   const rto::Sweep &sw = s->sw;
   rto::CellTables &ct = s->ct;
   const int M = p.M, N = p.N, G = p.G;
   double W = 0.;
   for (int m = 0; m < M; m++) W += s->wt[m];
   ct.xA.assign(N, 0.0); ct.xE.assign(N, 0.0); ct.xD.assign(N, 0.0);
   std::vector<double> Tn(N);
   for (int c = 0; c < N; c++)
   {
      double A = 0., E = 0., D = 0.;
      for (int g = 0; g < G; g++)
      {
         const size_t o = (size_t)g + (size_t)G * c;
         const double sig = ct.rho[c] * ct.kappa[o];
         double phi = 0.;
         for (int m = 0; m < M; m++) phi += s->wt[m] * sw.psi[sw.ip(m, g, c)];
         A += sig * phi;
         E += W * (sig * ct.B[o]);
         D += W * (sig * ct.dBdT[o]);
      }
      ct.xA[c] = A; ct.xE[c] = E; ct.xD[c] = D;
      const double rcv = ct.rho[c] * p.cv, T = ct.T[c];
      if (first_step || p.ts_method != 3) Tn[c] = T + p.dt * (A - E) / (rcv + p.dt * D);
      else Tn[c] = (rcv * (2. * T - 0.5 * ct.T_prev[c]) + p.dt * (A - E + D * T)) / (1.5 * rcv + p.dt * D);
   }
   ct.T_prev = ct.T;
   ct.T = Tn;
   s->p.T_cell = ct.T.data();
   rto_cell_tables(s);
}

extern "C" {

// GLQuad::build, GLQuad.cpp:4-44 (Newton on the Legendre recurrence, tolerance 1e-12, truncated PI)
void rto_glquad(int M, double norm, double *mu, double *wt)
{
   const double x1 = -1.0, x2 = 1.0;
   const double xm = 0.5 * (x2 + x1), xl = 0.5 * (x2 - x1);
   const double dnp = static_cast<double>(M);
   const int m = (M + 1) / 2;
   for (int i = 1; i <= m; ++i)
   {
      double di = static_cast<double>(i);
      double z1, pp, z = std::cos(rto::PI_TRUNC * (di - 0.25) / (dnp + 0.5));
      do
      {
         double p1 = 1.0, p2 = 0.0;
         for (int j = 1; j <= M; ++j)
         {
            double dj = static_cast<double>(j);
            double p3 = p2;
            p2 = p1;
            p1 = ((2.0 * dj - 1.0) * z * p2 - (dj - 1.0) * p3) / dj;
         }
         pp = dnp * (z * p1 - p2) / (z * z - 1.0);
         z1 = z;
         z = z1 - p1 / pp;
      } while (std::fabs(z - z1) > 1.0e-12);
      mu[i - 1] = xm - xl * z;
      mu[M - i] = xm + xl * z;
      wt[i - 1] = norm * xl / ((1.0 - z * z) * pp * pp);
      wt[M - i] = wt[i - 1];
   }
}

// Solver::generate_group_edges, solver.cpp:6-19.  For G == 1 the reference's
// `assert(logfac = 1.)` is an assignment; logfac is unused then anyway.
void rto_group_edges(int G, double efirst, double elast, double *e_edge)
{
   double logfac = (std::log(elast) - std::log(efirst)) / (G - 1.0);
   logfac = std::exp(logfac);
   e_edge[0] = 0.0;
   e_edge[1] = efirst;
   for (int g = 1; g < G; g++) e_edge[g + 1] = e_edge[g] * logfac;
}

void rto_planck(double T, int G, const double *e_edge, double *B, double *dBdT)
{
   rto::GaussRule q(rto::PLANCK_ACCURACY);
   std::fill(B, B + G, 0.0);
   std::fill(dBdT, dBdT + G, 0.0);
   rto::planck_groups(q, T, G, e_edge, B, dBdT);
}

// Correction::generate_multigroup_opacities, correction.cpp:66-97 (present in the reference but never called; second
// statement in tests/test_planckian.cpp:267-349): Planck-weighted group opacities normalised so that
// sum_g kappa_g B_g = kappa_grey a c T^4.  B is in jk units (after the unit change of correction.cpp:31-35).
void rto_multigroup_opacities(double T, double kappa_grey, int G, const double *e_edge, const double *B, double *kappa,
                              double *emis_spec, double *ukappa, double *ckappa)
{
   using namespace rto;
   std::vector<double> uk(G), ck(G);
   const double fourpi = 4.0 * PI_TRUNC;
   for (int g = 0; g < G; g++)
   {
      const double tmp1 = 2.0 * fourpi * K_JPK * T / (std::pow(H_KEV_SH, 3) * std::pow(C_LIGHT, 2));
      const double tmp2 = std::exp(-e_edge[g] / T) - std::exp(-e_edge[g + 1] / T);
      const double e_ave = 0.5 * (e_edge[g] + e_edge[g + 1]);
      uk[g] = tmp1 * tmp2 / B[g];
      ck[g] = (1.0 - std::exp(-e_ave / T)) / std::pow(e_ave, 3);
   }
   double emis_tot = 0.0;
   const double acT4 = RAD_A * C_LIGHT * std::pow(T, 4);
   for (int g = 0; g < G; g++) emis_tot += uk[g] * B[g];
   const double nfac = acT4 / emis_tot;
   for (int g = 0; g < G; g++)
   {
      const double k = kappa_grey * nfac * uk[g];
      if (kappa) kappa[g] = k;
      if (emis_spec) emis_spec[g] = k * B[g];
      if (ukappa) ukappa[g] = uk[g];
      if (ckappa) ckappa[g] = ck[g];
   }
}

rto_solver *rto_create(const rto_params *p)
{
   rto_solver *s = new rto_solver();
   s->p = *p;
   const int M = p->M, N = p->N, G = p->G;
   s->mu.assign(p->mu, p->mu + M);
   s->wt.assign(p->wt, p->wt + M);
   s->p.mu = s->mu.data();
   s->p.wt = s->wt.data();
   s->gt.setup(G, p->e_edge, p->kappa, p->rho);
   s->p.e_edge = s->gt.e_edge.data();
   s->p.kappa = s->gt.kappa.data();
   rto::Sweep &sw = s->sw;
   sw.M = M; sw.N = N; sw.G = G; sw.dx = p->dx; sw.dt = p->dt;
   const size_t n3 = (size_t)M * G * N;
   sw.ends.assign(2 * n3, 0.0);
   sw.prev.assign(2 * n3, 0.0);
   sw.half.assign(2 * n3, 0.0);
   sw.psi.assign(n3, 0.0);
   sw.psi_source.assign((size_t)M * G, 0.0);
   if (p->psi_source) std::copy(p->psi_source, p->psi_source + (size_t)M * G, sw.psi_source.begin());
   s->p.psi_source = sw.psi_source.data();
   s->var = p->rho_cell || p->T_cell || p->kappa_cell || p->cv > 0.;
   if (s->var)
   {
      rto::CellTables &ct = s->ct;
      ct.rho.assign(N, p->rho); ct.T.assign(N, p->T); ct.kappa.assign((size_t)G * N, 0.0);
      if (p->rho_cell) std::copy(p->rho_cell, p->rho_cell + N, ct.rho.begin());
      if (p->T_cell) std::copy(p->T_cell, p->T_cell + N, ct.T.begin());
      for (int c = 0; c < N; c++)
         for (int g = 0; g < G; g++) ct.kappa[(size_t)g + (size_t)G * c] = p->kappa_cell ? p->kappa_cell[(size_t)g + (size_t)G * c] : p->kappa[g];
      s->p.rho_cell = ct.rho.data(); s->p.T_cell = ct.T.data(); s->p.kappa_cell = ct.kappa.data();
      rto_cell_tables(s);
   }
   // Correction ctor computes the Planck integrals once (correction.cpp:318-319); the solver then
   // initialises psi = ends(L) = ends(R) = B_g (solver.cpp:166-181).
   rto::planck_groups(s->gauss, p->T, G, s->gt.e_edge.data(), s->gt.B.data(), s->gt.dBdT.data());
   for (int c = 0; c < N; c++)
      for (int g = 0; g < G; g++)
         for (int i = 0; i < M; i++)
         {
            double v = s->var ? s->ct.B[(size_t)g + (size_t)G * c] : s->gt.B[g];
            sw.psi[sw.ip(i, g, c)] = v;
            sw.ends[sw.ie(i, g, c, 0)] = v;
            sw.ends[sw.ie(i, g, c, 1)] = v;
         }
   return s;
}

void rto_destroy(rto_solver *s) { delete s; }

int rto_total_stages(const rto_solver *s) { return s->p.max_timesteps * (s->p.ts_method == 3 ? 4 : 1); }
int rto_stages_done(const rto_solver *s) { return s->stages_done; }

void rto_run_stages(rto_solver *s, int n)
{
   const rto_params &p = s->p;
   rto::Sweep &sw = s->sw;
   rto::GroupTerms &gt = s->gt;
   const int M = p.M, G = p.G;
   const double beta = p.V / rto::C_LIGHT;
   for (int k = 0; k < n; k++)
   {
      const int it = s->stages_done;
      if (it == 0 && p.use_mg_equilib)
      {
         // Solver::computeEquilibriumSources, solver.cpp:287-315
         gt.per_stage(s->gauss, p.T);
         for (int i = 0; i < M; i++)
            for (int g = 0; g < G; g++)
            {
               double Bg = gt.B[g], dEBg = gt.dEB[g];
               if (s->var)
               {
                  // extension: the boundary cell's material state (cell 0 feeds mu > 0, cell N-1 feeds mu < 0)
                  const size_t o = (size_t)g + (size_t)G * (s->mu[i] < 0. ? p.N - 1 : 0);
                  Bg = s->ct.B[o]; dEBg = s->ct.dEB[o];
               }
               double val = 4 * Bg - dEBg;
               double mult = s->mu[i] * p.V / rto::C_LIGHT;
               val *= mult;
               val += Bg;
               sw.psi_source[i + (size_t)M * g] = val;
            }
      }
      gt.per_stage(s->gauss, p.T);               // solver.cpp:608
      int kind;
      double tau;
      bool snapshot = false;
      if (p.ts_method != 3 || it % 4 == 0) sw.prev = sw.ends; // solver.cpp:620-625
      if (p.ts_method == 1) { kind = 0; tau = p.dt; }
      else if (p.ts_method == 2) { kind = 1; tau = p.dt; }
      else
      {
         const int ind = it % 4;                 // solver.cpp:723-750
         kind = (ind == 0 || ind == 2) ? 0 : (ind == 1 ? 1 : 2);
         tau = p.dt / 2.;
         snapshot = (ind == 1);
      }
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(p.num_threads > 0 ? p.num_threads : 1)
#endif
      for (int g = 0; g < G; g++)
         sw.sweep_group(g, kind, tau, snapshot, s->mu.data(), p.bc_left, p.bc_right, gt.rho[g], gt.kappa[g], gt.B[g],
                        p.use_correction != 0, gt.cor1[g], gt.cor2[g], gt.cor3[g], beta, s->var ? &s->ct : nullptr);
      s->stages_done++;
      const int per_step = p.ts_method == 3 ? 4 : 1;
      if (p.cv > 0. && s->stages_done % per_step == 0) rto_temperature_update(s, s->stages_done == per_step);
   }
}

void rto_get_psi(const rto_solver *s, double *psi) { std::copy(s->sw.psi.begin(), s->sw.psi.end(), psi); }

void rto_get_ends(const rto_solver *s, int which, double *out)
{
   const std::vector<double> &v = which == 0 ? s->sw.ends : (which == 1 ? s->sw.prev : s->sw.half);
   std::copy(v.begin(), v.end(), out);
}

// solver.cpp:191-237 (ascending-m accumulation)
void rto_moments(const rto_solver *s, double *phi, double *phi_plus, double *F)
{
   const rto::Sweep &sw = s->sw;
   const int M = sw.M, N = sw.N, G = sw.G;
   for (int g = 0; g < G; g++)
      for (int c = 0; c < N; c++)
      {
         double a = 0., b = 0., f = 0.;
         for (int m = 0; m < M; m++) a += s->wt[m] * sw.psi[sw.ip(m, g, c)];
         for (int m = M / 2; m < M; m++) b += s->wt[m] * sw.psi[sw.ip(m, g, c)];
         for (int m = 0; m < M; m++) f += s->mu[m] * s->wt[m] * sw.psi[sw.ip(m, g, c)];
         if (phi) phi[g + (size_t)G * c] = a;
         if (phi_plus) phi_plus[g + (size_t)G * c] = b;
         if (F) F[g + (size_t)G * c] = f;
      }
}

// solver.cpp:240-284
void rto_balance(const rto_solver *s, double *balance)
{
   const rto::Sweep &sw = s->sw;
   const int M = sw.M, N = sw.N, G = sw.G;
   const double ac = rto::RAD_A * rto::C_LIGHT;
   std::vector<double> phi((size_t)G * N);
   rto_moments(s, phi.data(), nullptr, nullptr);
   for (int g = 0; g < G; g++)
   {
      double j_half_minus = 0., j_half_plus = 0., jN_half_minus = 0., jN_half_plus = 0., a = 0., src = 0.;
      for (int m = 0; m < M; m++)
      {
         double mu = s->mu[m];
         if (mu < 0.)
         {
            j_half_minus -= sw.ends[sw.ie(m, g, 0, 0)] * mu * s->wt[m];
            jN_half_minus -= sw.ends[sw.ie(m, g, N - 1, 0)] * mu * s->wt[m];
         }
         else
         {
            j_half_plus += sw.ends[sw.ie(m, g, 0, 1)] * mu * s->wt[m];
            jN_half_plus += sw.ends[sw.ie(m, g, N - 1, 1)] * mu * s->wt[m];
         }
      }
      for (int c = 0; c < N; c++)
      {
         double rho = s->gt.rho[g], kap = s->gt.kappa[g], T = s->p.T;
         if (s->var) { rho = s->ct.rho[c]; kap = s->ct.kappa[(size_t)g + (size_t)G * c]; T = s->ct.T[c]; }
         a += rho * kap * phi[g + (size_t)G * c] * sw.dx;
         src += rho * kap * ac * std::pow(T, 4) * sw.dx;
      }
      double sources = j_half_plus + jN_half_minus + src;
      double sinks = jN_half_plus + j_half_minus + a;
      balance[g] = std::fabs(sinks - sources) / sources;
   }
}

// solver.cpp:826-850
void rto_group_ends(const rto_solver *s, double *left, double *right)
{
   const rto::Sweep &sw = s->sw;
   for (int g = 0; g < sw.G; g++)
   {
      double l = 0., r = 0.;
      for (int m = 0; m < sw.M; m++)
      {
         if (s->mu[m] < 0.) l += sw.ends[sw.ie(m, g, 0, 0)];
         else r += sw.ends[sw.ie(m, g, sw.N - 1, 1)];
      }
      left[g] = l / (s->gt.de_ave[g] * rto::C_LIGHT);
      right[g] = r / (s->gt.de_ave[g] * rto::C_LIGHT);
   }
}

void rto_get_group_terms(const rto_solver *s, double *B, double *dBdT, double *dEB, double *dsigEdE, double *dkapEB,
                         double *kappa_edge)
{
   const rto::GroupTerms &gt = s->gt;
   if (B) std::copy(gt.B.begin(), gt.B.end(), B);
   if (dBdT) std::copy(gt.dBdT.begin(), gt.dBdT.end(), dBdT);
   if (dEB) std::copy(gt.dEB.begin(), gt.dEB.end(), dEB);
   if (dsigEdE) std::copy(gt.dsigEdE.begin(), gt.dsigEdE.end(), dsigEdE);
   if (dkapEB) std::copy(gt.dkapEB.begin(), gt.dkapEB.end(), dkapEB);
   if (kappa_edge) std::copy(gt.kappa_edge.begin(), gt.kappa_edge.end(), kappa_edge);
}

void rto_get_psi_source(const rto_solver *s, double *out)
{
   std::copy(s->sw.psi_source.begin(), s->sw.psi_source.end(), out);
}

void rto_get_temperature(const rto_solver *s, double *T, double *A, double *E, double *D)
{
   const int N = s->p.N;
   for (int c = 0; c < N; c++)
   {
      if (T) T[c] = s->var ? s->ct.T[c] : s->p.T;
      if (A) A[c] = s->ct.xA.empty() ? 0. : s->ct.xA[c];
      if (E) E[c] = s->ct.xE.empty() ? 0. : s->ct.xE[c];
      if (D) D[c] = s->ct.xD.empty() ? 0. : s->ct.xD[c];
   }
}

// correction.cpp:39-63 and 100-122
int rto_validate(const rto_solver *s, double *out3)
{
   const rto::GroupTerms &gt = s->gt;
   const double ac = rto::RAD_A * rto::C_LIGHT, T = s->p.T;
   double bsum = 0., dbsum = 0., emis = 0.;
   for (int g = 0; g < gt.G; g++) { bsum += gt.B[g]; dbsum += gt.dBdT[g]; }
   for (int g = 0; g < gt.G; g++) emis += gt.kappa[g] * gt.B[g];
   if (out3) { out3[0] = bsum; out3[1] = dbsum; out3[2] = emis; }
   double acT4 = ac * std::pow(T, 4), dacT4 = 4.0 * ac * std::pow(T, 3);
   bool ok = !(std::fabs(acT4 - bsum) > rto::VALIDATION_TOL || std::fabs(dacT4 - dbsum) > rto::VALIDATION_TOL);
   ok = ok && !(std::fabs(emis - s->p.kappa_grey * acT4) > rto::VALIDATION_TOL);
   return ok ? 1 : 0;
}

} // extern "C"
