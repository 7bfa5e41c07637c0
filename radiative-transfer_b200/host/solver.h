// host/solver.h -- drop-in replacement for the reference's include/solver.h on top of librt_cuda.so.
//
// The PUBLIC section is the reference's rt::Solver surface (reference include/solver.h:76-98): same
// constructor, same method names, same argument types, same ownership rule (the caller owns psi_mat,
// phi and F; the solver writes into them when solve() / compute_*() return).  The reference's drivers
// (src/main.cc:100-133, tests/test_gray.cpp:68-79) compile against this header unchanged; ParameterHandler
// and GLQuad are the reference's own host classes and are used as they are.
//
// The PRIVATE section is different on purpose: no Eigen state tensors, no 2x2 scratch, no Correction
// object -- the state lives in HBM behind an rt_solver handle (include/rt_cuda.h) and every number is
// produced by the CUDA kernels.  There is no CPU fallback: if the handle cannot be created the
// constructor prints the library's message and exits like the reference's own fatal paths
// (ParameterHandler.cpp:146-149).
#ifndef SOLVER
#define SOLVER

#include "constants.h"
#include "ParameterHandler.h"
#include "GLQuad.h"

#include <Eigen/Dense>
#include <unsupported/Eigen/CXX11/Tensor>
#include <iomanip>
#include <iostream>
#include <string>
#include <vector>

struct rt_multi; // include/rt_cuda.h: the multi-device handle (one shard per GPU of RT_CUDA_DEVICES)

using namespace std;

namespace rt
{
class Solver
{
private:
   ParameterHandler &ph;
   int M, N, num_groups;
   double efirst, elast, logfac;

   // caller-owned result buffers (solver.h:30-33 in the reference)
   Eigen::Tensor<double, 3> &psi_mat_ref;
   Eigen::Ref<Eigen::MatrixXd> phi_ref;
   Eigen::Ref<Eigen::MatrixXd> F_ref;

   // host-side setup that stays on the CPU: quadrature, energy grid, opacity table, boundary sources
   Eigen::VectorXd m_mu, m_wt;
   Eigen::VectorXd e_edge, e_ave, de_ave;
   Eigen::MatrixXd energy_discretization;
   Eigen::VectorXd kappa_vec;
   Eigen::MatrixXd psi_source;

   // results owned by the solver
   Eigen::MatrixXd phi_plus;
   Eigen::VectorXd balance, left_ends, right_ends;

   rt_multi *handle;
   std::vector<double> stage_buf; // contiguous staging between the C ABI and Eigen objects

   void die(const char *where) const;
   void pull_psi();
   void pull_moment(int which, Eigen::Ref<Eigen::MatrixXd> dst);

public:
   void generate_group_edges();
   void generate_group_averages();
   void fill_energy_bound_arrays();
   Solver(ParameterHandler &parameter_handler,
          Eigen::Tensor<double, 3> &psi_mat,
          Eigen::Ref<Eigen::MatrixXd> phi,
          Eigen::Ref<Eigen::MatrixXd> F);
   ~Solver();

   void compute_angle_integrated_intensity();
   void compute_positive_angle_integrated_intensity();
   void compute_radiative_flux();
   void compute_balance();
   void get_balance(Eigen::Ref<Eigen::VectorXd> balance) { balance = this->balance; }
   void get_phi_plus(Eigen::Ref<Eigen::MatrixXd> phi_plus) { phi_plus = this->phi_plus; }

   void solve();

   void get_e_ave(Eigen::Ref<Eigen::VectorXd> e_ave) { e_ave = this->e_ave; }
   void compute_group_ends();
   void get_ends(const string side, Eigen::Ref<Eigen::VectorXd> group_ends);

   // additions (not in the reference): the handle for callers that want timing / per-step hooks
   rt_multi *c_handle() const { return handle; }
   // Correction::Print() (correction.cpp:404-603) -- the reference's Solver keeps its Correction private and never calls
   // Print; here the report is formatted from the device's tables.  RT_CUDA_PRINT_CORRECTION=1 prints it from the
   // constructor, so that an unmodified driver can show it.
   void print_correction(std::ostream &os = std::cout);
};
} // End namespace rt
#endif // SOLVER
