// oracle/ref_dump.cpp -- TEST INFRASTRUCTURE (oracle-1 harness), not product code.
//
// Drives the UNMODIFIED reference solver (compiled in place from
// /root/reference/src/*.cpp by oracle/Makefile) exactly the way the reference's
// own driver does (reference src/main.cc:79-133) and dumps every buffer that the
// parity tests compare, as raw little-endian fp64 records, because the
// reference's CSV writers keep only 6 significant digits.
//
// usage: ref_dump <file.prm> <out.bin>      (cwd must contain ../prm/ for the
//        ref_dump <file.prm> --print        the reference's own Correction::Print() (correction.cpp:404-603) after one
//                                            compute_correction, between two marker lines on stdout: the golden text of
//                                            tests/golden/correction_print_*.txt (no reference driver ever calls it)
//
//                                            group-bounds / kappa data files,
//                                            reference src/ParameterHandler.cpp:141,172)
//
// record format: char name[32]; int64 ndim; int64 dims[ndim]; double data[prod(dims)]
// (column-major / first index fastest, i.e. the reference's own Eigen order).
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

// The reference keeps ends/prev_ends/half_ends, B, dEB, mu, wt ... private and
// offers no getter (reference include/solver.h:21-60).  The harness needs them
// for the 1e-9 comparison, so access control is lifted for the two class headers.
#define private public
#include "solver.h"
#undef private

static FILE *g_out = nullptr;

static void put(const char *name, const double *p, std::vector<int64_t> dims)
{
   char nm[32];
   std::memset(nm, 0, sizeof nm);
   std::strncpy(nm, name, 31);
   std::fwrite(nm, 1, 32, g_out);
   int64_t nd = (int64_t)dims.size(), n = 1;
   std::fwrite(&nd, 8, 1, g_out);
   for (auto d : dims) { std::fwrite(&d, 8, 1, g_out); n *= d; }
   std::fwrite(p, 8, (size_t)n, g_out);
}

static void put_scalar(const char *name, double v) { put(name, &v, {1}); }

int main(int argc, char **argv)
{
   if (argc != 3) { std::fprintf(stderr, "usage: ref_dump <file.prm> <out.bin> | ref_dump <file.prm> --print\n"); return 2; }
   ParameterHandler ph(argv[1]);
   const int M = ph.get_M(), N = ph.get_N(), G = ph.get_G();

   Eigen::Tensor<double, 3> psi(M, G, N);
   psi.setConstant(0.);
   Eigen::MatrixXd phi(G, N), F(G, N), phi_plus(G, N);
   Eigen::VectorXd balance(G), e_ave(G), left_ends(G), right_ends(G);

   rt::Solver solver(ph, psi, phi, F);

   if (std::strcmp(argv[2], "--print") == 0)
   {
      solver.correction->compute_correction(psi); // fills the edge opacities, differences and cor1..3 (correction.cpp:372-401)
      std::cout << "@@@ Correction::Print begin" << std::endl;
      solver.correction->Print();
      std::cout << "@@@ Correction::Print end" << std::endl;
      return 0;
   }

   g_out = std::fopen(argv[2], "wb");
   if (!g_out) { std::perror("open out"); return 2; }

   put("mu", solver.m_mu.data(), {M});
   put("wt", solver.m_wt.data(), {M});
   put("e_edge", solver.e_edge.data(), {G + 1});
   put("kappa", solver.kappa_vec.data(), {G});
   put("B_init", solver.B.data(), {G});
   put("psi_init", psi.data(), {M, G, N});

   auto t0 = std::chrono::steady_clock::now();
   solver.solve();
   auto t1 = std::chrono::steady_clock::now();

   solver.compute_angle_integrated_intensity();
   solver.compute_radiative_flux();
   solver.compute_balance();
   solver.get_balance(balance);
   solver.compute_positive_angle_integrated_intensity();
   solver.get_phi_plus(phi_plus);
   solver.get_e_ave(e_ave);
   solver.compute_group_ends();
   solver.get_ends("left", left_ends);
   solver.get_ends("right", right_ends);

   put_scalar("solve_seconds", std::chrono::duration<double>(t1 - t0).count());
   put("psi", psi.data(), {M, G, N});
   put("ends", solver.ends.data(), {M, G, N, 2});
   put("prev_ends", solver.prev_ends.data(), {M, G, N, 2});
   put("half_ends", solver.half_ends.data(), {M, G, N, 2});
   put("phi", phi.data(), {G, N});
   put("F", F.data(), {G, N});
   put("phi_plus", phi_plus.data(), {G, N});
   put("balance", balance.data(), {G});
   put("e_ave", e_ave.data(), {G});
   put("left_ends", left_ends.data(), {G});
   put("right_ends", right_ends.data(), {G});
   put("psi_source", solver.psi_source.data(), {M, G});
   put("B", solver.correction->B.data(), {G});
   put("dBdT", solver.correction->dBdT.data(), {G});
   put("dEB", solver.correction->dEB.data(), {G});
   put("dsigEdE", solver.correction->dsigEdE.data(), {G});
   put("dkapEB", solver.correction->dkapEB.data(), {G});
   put("kappa_edge", solver.correction->kappa_edge.data(), {G + 1});
   put("cor1", solver.correction->cor1.data(), {G, N});
   put("cor2", solver.correction->cor2.data(), {G, N});
   put("cor3", solver.correction->cor3.data(), {G, N});
   if (solver.correction->total_correction.size() == (std::ptrdiff_t)M * G * N)
      put("total_correction", solver.correction->total_correction.data(), {M, G, N});
   std::fclose(g_out);
   return 0;
}
