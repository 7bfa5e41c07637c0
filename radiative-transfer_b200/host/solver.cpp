// host/solver.cpp -- rt::Solver (reference include/solver.h:76-98) implemented on the C ABI of librt_cuda.so.
//
// Replaces the reference's src/solver.cpp + src/correction.cpp + src/Planck.cpp in a build of the reference's
// drivers: link main.cc / test_gray.cpp with this file, the reference's own ParameterHandler.cpp, param.cpp and
// GLQuad.cpp, and -lrt_cuda.  Every method forwards to one entry point of include/rt_cuda.h; the only arithmetic
// done here is the O(G) energy grid (solver.cpp:6-43), which is an INPUT of the device path.
#include "solver.h"

#include "rt_cuda.h"

#include <cassert>
#include <cmath>
#include <cstdlib>

namespace rt
{
// ------------------------------------------------------------------------------------------------
// energy grid (solver.cpp:6-43): same operations in the same order, so e_edge is bit-identical
void Solver::generate_group_edges()
{
   logfac = exp((log(elast) - log(efirst)) / (num_groups - 1.0));
   if (num_groups == 1) logfac = 1.; // the reference's assert(logfac = 1.) is an assignment
   e_edge(0) = 0.0;
   e_edge(1) = efirst;
   for (int g = 2; g <= num_groups; g++) e_edge(g) = e_edge(g - 1) * logfac;
}

void Solver::generate_group_averages()
{
   for (int g = 0; g < num_groups; g++)
   {
      e_ave(g) = 0.5 * (e_edge(g) + e_edge(g + 1));
      de_ave(g) = e_edge(g + 1) - e_edge(g);
   }
}

void Solver::fill_energy_bound_arrays()
{
   for (int g = 0; g < num_groups; g++)
   {
      energy_discretization(g, 0) = e_edge(g);
      energy_discretization(g, 1) = e_edge(g + 1);
   }
}

// ------------------------------------------------------------------------------------------------
void Solver::die(const char *where) const
{
   // the reference's fatal paths are assert / exit(1) (ParameterHandler.cpp:146-149, solver.cpp:609-612)
   cerr << where << ": " << rt_last_error() << endl;
   exit(1);
}

Solver::Solver(ParameterHandler &parameter_handler,
               Eigen::Tensor<double, 3> &psi_mat,
               Eigen::Ref<Eigen::MatrixXd> phi,
               Eigen::Ref<Eigen::MatrixXd> F)
   : ph(parameter_handler), psi_mat_ref(psi_mat), phi_ref(phi), F_ref(F), handle(nullptr)
{
   cout << "Solver constructor.\n";
   M = ph.get_M();
   N = ph.get_N();
   num_groups = ph.get_G();
   efirst = ph.get_efirst();
   elast = ph.get_elast();
   logfac = 1.;

   // boundary sources are only fetched when some BC is a source; otherwise zero (solver.cpp:64-73)
   psi_source.resize(M, num_groups);
   psi_source.setConstant(0.);
   const bool have_src = ph.get_bc_left_indicator() == 1 || ph.get_bc_right_indicator() == 1;
   if (have_src) ph.get_psi_source(psi_source);

   // direction quadrature: the reference's own GLQuad (solver.cpp:76-78)
   GLQuad quad(M, Constants::FOUR_PI);
   m_mu = quad.mu();
   m_wt = quad.wt();
   cout << setw(16) << left << "Mu" << setw(16) << left << "Wt" << endl;
   for (int i = 0; i < M; ++i) cout << showpos << setw(16) << left << m_mu(i) << setw(16) << left << m_wt(i) << endl;
   cout << noshowpos << endl;

   // energy discretisation (solver.cpp:82-104)
   e_edge.resize(num_groups + 1);
   e_ave.resize(num_groups);
   de_ave.resize(num_groups);
   energy_discretization.resize(num_groups, 2);
   if (ph.get_have_group_bounds()) ph.get_group_bounds(e_edge);
   else generate_group_edges();
   generate_group_averages();
   fill_energy_bound_arrays();
   cout << left << setw(13) << "Group Index" << left << setw(16) << "Average Energy" << left << setw(14) << "Upper Energy"
        << left << setw(13) << "Group Width" << endl;
   for (int g = 0; g < num_groups; ++g)
      cout << left << setw(13) << g << left << setw(16) << e_ave(g) << left << setw(14) << e_edge(g + 1) << left << setw(13)
           << de_ave(g) << endl;
   cout << "\n" << endl;

   // opacities (solver.cpp:145-152)
   kappa_vec.resize(num_groups);
   if (ph.get_have_group_absorption_opacities()) ph.get_group_kappa(kappa_vec);
   else kappa_vec.setConstant(ph.get_kappa_grey());

   phi_plus.resize(num_groups, N);
   balance.resize(num_groups);
   left_ends.resize(num_groups);
   right_ends.resize(num_groups);

   // marshal into plain contiguous arrays for the C ABI
   std::vector<double> mu(M), wt(M), edge(num_groups + 1), kap(num_groups), src((size_t)M * num_groups);
   for (int i = 0; i < M; ++i) { mu[i] = m_mu(i); wt[i] = m_wt(i); }
   for (int g = 0; g <= num_groups; ++g) edge[g] = e_edge(g);
   for (int g = 0; g < num_groups; ++g) kap[g] = kappa_vec(g);
   for (int g = 0; g < num_groups; ++g)
      for (int i = 0; i < M; ++i) src[i + (size_t)M * g] = psi_source(i, g);

   rt_params p;
   p.M = M; p.N = N; p.G = num_groups;
   p.group_begin = 0; p.group_count = num_groups;
   p.ts_method = ph.get_ts_method();
   p.max_timesteps = ph.get_max_timesteps();
   p.bc_left = ph.get_bc_left_indicator();
   p.bc_right = ph.get_bc_right_indicator();
   p.use_correction = ph.get_use_correction() ? 1 : 0;
   p.use_mg_equilib = ph.get_use_mg_equilib() ? 1 : 0;
   p.validate = ph.get_validation() ? 1 : 0;
   // GPUs: RT_CUDA_DEVICES = a count ("8": devices 0..7) or a list ("0,2,5"); default: the one device RT_CUDA_DEVICE (or 0).
   // The energy groups are sharded over them inside the library (rt_multi_*, one host thread per device).
   std::vector<int> devices;
   if (const char *dl = getenv("RT_CUDA_DEVICES"))
   {
      const std::string str(dl);
      if (str.find(',') == std::string::npos) { for (int k = 0; k < atoi(dl); ++k) devices.push_back(k); }
      else
      {
         size_t pos = 0;
         while (pos <= str.size())
         {
            const size_t e = str.find(',', pos);
            const std::string tok = str.substr(pos, e == std::string::npos ? std::string::npos : e - pos);
            if (!tok.empty()) devices.push_back(atoi(tok.c_str()));
            if (e == std::string::npos) break;
            pos = e + 1;
         }
      }
   }
   if (devices.empty())
   {
      const char *dev = getenv("RT_CUDA_DEVICE");
      devices.push_back(dev ? atoi(dev) : 0);
   }
   if ((int)devices.size() > num_groups) devices.resize(num_groups); // at least one group per shard
   p.device = devices[0];
   p.groups_per_batch = 0;
   p.dx = ph.get_dx(); p.dt = ph.get_dt(); p.rho = ph.get_rho(); p.T = ph.get_T(); p.V = ph.get_V();
   p.kappa_grey = ph.get_kappa_grey();
   p.mu = mu.data(); p.wt = wt.data(); p.e_edge = edge.data(); p.kappa = kap.data();
   p.psi_source = have_src ? src.data() : nullptr;
   p.rho_cell = p.T_cell = p.kappa_cell = nullptr; // the reference's slab is uniform (solver.cpp:150-157)
   p.cv = 0.0;                                     // and its temperature never changes
   if (rt_multi_create(&p, (int)devices.size(), devices.data(), &handle)) die("Solver");
   if (devices.size() > 1)
   {
      cout << "energy groups sharded over " << devices.size() << " GPUs:";
      for (int d : devices) cout << " " << d;
      cout << endl;
   }

   // initial condition psi = B_g, evaluated on the device (solver.cpp:166-181)
   pull_psi();
   std::vector<double> B(num_groups);
   if (rt_multi_get_group_terms(handle, B.data(), nullptr, nullptr, nullptr, nullptr, nullptr)) die("Solver");
   cout << "B:";
   for (int g = 0; g < num_groups; ++g) cout << " " << B[g];
   cout << endl;
   if (const char *pc = getenv("RT_CUDA_PRINT_CORRECTION"))
      if (atoi(pc) != 0) print_correction();
   cout << "end solver constructor\n";
}

Solver::~Solver()
{
   if (handle) rt_multi_destroy(handle);
   handle = nullptr;
}

// psi of the current state into the caller's tensor (reference index order == the C ABI's)
void Solver::pull_psi()
{
   int shards = 0;
   if (rt_multi_shards(handle, &shards, nullptr, nullptr, nullptr)) die("Solver");
   for (int k = 0; k < shards; ++k)
   {
      rt_plan plan;
      if (rt_multi_get_plan(handle, k, &plan)) die("Solver");
      if (!plan.resident) return; // batched solves keep psi on the device; moments remain available
   }
   const size_t n = (size_t)M * num_groups * N;
   stage_buf.resize(n);
   if (rt_multi_get_psi(handle, stage_buf.data())) die("Solver");
   size_t k = 0;
   for (int c = 0; c < N; ++c)
      for (int g = 0; g < num_groups; ++g)
         for (int i = 0; i < M; ++i) psi_mat_ref(i, g, c) = stage_buf[k++];
}

void Solver::pull_moment(int which, Eigen::Ref<Eigen::MatrixXd> dst)
{
   const size_t n = (size_t)num_groups * N;
   stage_buf.resize(n);
   double *b = stage_buf.data();
   if (rt_multi_moments(handle, which == 0 ? b : nullptr, which == 1 ? b : nullptr, which == 2 ? b : nullptr)) die("Solver");
   size_t k = 0;
   for (int c = 0; c < N; ++c)
      for (int g = 0; g < num_groups; ++g) dst(g, c) = b[k++];
}

void Solver::solve()
{
   // the whole stage loop of solver.cpp:590-823 runs on the device; validation failure is fatal like the assert
   if (rt_multi_solve(handle)) die("Solver::solve");
   int done = 0;
   rt_multi_stages_done(handle, &done);
   cout << "============= stage sweeps: " << done << " =============" << endl;
   pull_psi();
}

void Solver::compute_angle_integrated_intensity() { pull_moment(0, phi_ref); }

void Solver::compute_positive_angle_integrated_intensity() { pull_moment(1, phi_plus); }

void Solver::compute_radiative_flux() { pull_moment(2, F_ref); }

void Solver::compute_balance()
{
   stage_buf.resize(num_groups);
   if (rt_multi_balance(handle, stage_buf.data())) die("Solver::compute_balance");
   for (int g = 0; g < num_groups; ++g)
   {
      balance(g) = stage_buf[g];
      cout << "balance at (" << g << "): " << balance(g) << endl;
   }
}

void Solver::print_correction(std::ostream &os)
{
   size_t n = 0;
   if (rt_multi_correction_print(handle, nullptr, 0, &n)) die("Solver::print_correction");
   std::vector<char> text(n + 1);
   if (rt_multi_correction_print(handle, text.data(), text.size(), &n)) die("Solver::print_correction");
   os << text.data();
}

void Solver::compute_group_ends()
{
   std::vector<double> l(num_groups), r(num_groups);
   if (rt_multi_group_ends(handle, l.data(), r.data())) die("Solver::compute_group_ends");
   for (int g = 0; g < num_groups; ++g) { left_ends(g) = l[g]; right_ends(g) = r[g]; }
}

void Solver::get_ends(const string side, Eigen::Ref<Eigen::VectorXd> group_ends)
{
   assert((side == "left" || side == "right") && "Invalid option for 'side'.");
   if (side == "left") group_ends = left_ends;
   else group_ends = right_ends;
}

} // namespace rt
