// rt_multi.cu -- ONE process, SEVERAL GPUs: the multi-device handle of include/rt_cuda.h (rt_multi_*).
//
// The reference's drivers construct one rt::Solver in one process (src/main.cc:100-113, tests/test_gray.cpp:68-79), so
// for the C++ drop-in to use more than one B200 the sharding has to live under the C ABI ("the library owns
// streams, device buffers and NCCL communicators (single-process multi-GPU, one communicator per device)").
// A multi handle owns one rt_solver per device: contiguous blocks of energy groups (they never couple in the
// reference's solve), every shard driven by its own host thread.  Only the material-temperature extension
// (rt_params::cv) exchanges data: its group sums go through ncclAllReduce on communicators made here (ncclCommInitAll).
// Results are gathered on the host into the reference's index order over ALL owned groups.
// Built on the public C ABI plus the three helpers of rt_internal.h.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "rt_internal.h"

struct rt_multi
{
   rt_params p;                 // the whole problem as the caller described it
   std::vector<rt_solver *> h;  // one handle per shard
   std::vector<int> device, g0, gn;
   std::vector<void *> comm;    // NCCL communicators (temperature update on more than one shard), else empty
   int G_owned = 0;             // groups this multi handle owns (p.group_count)
   // per-step history (rt_multi_set_step_history) on more than one shard: every shard fills its own rows, the caller's
   // energy array is their sum
   double *hist_E = nullptr;
   int hist_cap = 0;
   std::vector<std::vector<double>> hist_part;
};

// the caller's energy rows = the shards' shares added in ascending shard order
static void combine_history(rt_multi *m)
{
   if (!m->hist_E || m->hist_part.empty()) return;
   const size_t n = (size_t)m->hist_cap * m->p.N;
   for (size_t i = 0; i < n; ++i)
   {
      double acc = 0.0;
      for (const auto &part : m->hist_part) acc += part[i];
      m->hist_E[i] = acc;
   }
}

static int mfail(const std::string &m) { return rt_internal_fail(m.c_str()); }

// run fn(shard) on one host thread per shard; the first failure's message becomes the caller's rt_last_error()
template <typename F>
static int for_each_shard(rt_multi *m, F fn)
{
   const int n = (int)m->h.size();
   std::vector<int> rc(n, 0);
   std::vector<std::string> err(n);
   auto body = [&](int k) {
      rc[k] = fn(k);
      if (rc[k]) err[k] = rt_last_error(); // (thread-local in the worker: copy it out)
   };
   if (n == 1) body(0);
   else
   {
      std::vector<std::thread> th;
      for (int k = 0; k < n; ++k) th.emplace_back(body, k);
      for (auto &t : th) t.join();
   }
   for (int k = 0; k < n; ++k)
      if (rc[k]) return mfail("shard " + std::to_string(k) + " (device " + std::to_string(m->device[k]) + ", groups [" + std::to_string(m->g0[k]) +
                              "," + std::to_string(m->g0[k] + m->gn[k]) + ")): " + err[k]);
   return 0;
}

extern "C" int rt_multi_destroy(rt_multi *m)
{
   if (!m) return 0;
   for (rt_solver *s : m->h)
      if (s) rt_destroy(s);
   for (size_t k = 0; k < m->comm.size(); ++k)
      if (m->comm[k])
      {
         cudaSetDevice(m->device[k]);
         rt_internal_nccl_destroy(m->comm[k]);
      }
   delete m;
   return 0;
}

extern "C" int rt_multi_create(const rt_params *p, int ndev, const int *devices, rt_multi **out)
{
   if (!out) return mfail("rt_multi_create: null out");
   *out = nullptr;
   if (!p || ndev < 1 || !devices) return mfail("rt_multi_create: null params / empty device list");
   if (p->group_count < ndev) return mfail("rt_multi_create: " + std::to_string(p->group_count) + " groups cannot be sharded over " + std::to_string(ndev) + " devices");
   rt_multi *m = new rt_multi();
   m->p = *p;
   m->G_owned = p->group_count;
   const int base = p->group_count / ndev, extra = p->group_count % ndev; // contiguous blocks, the first `extra` shards get one more
   for (int k = 0; k < ndev; ++k)
   {
      m->device.push_back(devices[k]);
      m->g0.push_back(p->group_begin + k * base + std::min(k, extra));
      m->gn.push_back(base + (k < extra ? 1 : 0));
   }
   m->h.assign(ndev, nullptr);
   const bool coupled = p->cv > 0.0 && ndev > 1;
   if (coupled)
   {
      for (int a = 0; a < ndev; ++a)
         for (int b = a + 1; b < ndev; ++b)
            if (devices[a] == devices[b])
            {
               rt_multi_destroy(m);
               return mfail("rt_multi_create: the temperature update (cv > 0) sums over the shards with NCCL, which needs one DISTINCT device per shard");
            }
      m->comm.assign(ndev, nullptr);
      if (rt_internal_nccl_init_all(m->comm.data(), ndev, devices))
      {
         const std::string e = rt_last_error();
         rt_multi_destroy(m);
         return mfail(e);
      }
   }
   // the shards are created one after the other (each rt_create sizes its batches from the free memory of ITS device;
   // two shards on one device must not plan against the same free bytes)
   for (int k = 0; k < ndev; ++k)
   {
      rt_params q = *p;
      q.device = devices[k];
      q.group_begin = m->g0[k];
      q.group_count = m->gn[k];
      if (rt_create(&q, &m->h[k]))
      {
         const std::string e = rt_last_error();
         rt_multi_destroy(m);
         return mfail("rt_multi_create: shard " + std::to_string(k) + " on device " + std::to_string(devices[k]) + ": " + e);
      }
      if (coupled && rt_set_allreduce(m->h[k], rt_internal_nccl_hook, m->comm[k]))
      {
         const std::string e = rt_last_error();
         rt_multi_destroy(m);
         return mfail(e);
      }
   }
   *out = m;
   return 0;
}

extern "C" int rt_multi_shards(const rt_multi *m, int *n, int *group_begin, int *group_count, int *device)
{
   if (!m || !n) return mfail("rt_multi_shards: null argument");
   *n = (int)m->h.size();
   for (int k = 0; k < *n; ++k)
   {
      if (group_begin) group_begin[k] = m->g0[k];
      if (group_count) group_count[k] = m->gn[k];
      if (device) device[k] = m->device[k];
   }
   return 0;
}

extern "C" int rt_multi_solve(rt_multi *m)
{
   if (!m) return mfail("rt_multi_solve: null handle");
   if (for_each_shard(m, [&](int k) { return rt_solve(m->h[k]); })) return 1;
   combine_history(m);
   return 0;
}

extern "C" int rt_multi_run_stages(rt_multi *m, int n)
{
   if (!m) return mfail("rt_multi_run_stages: null handle");
   if (for_each_shard(m, [&](int k) { return rt_run_stages(m->h[k], n); })) return 1;
   combine_history(m);
   return 0;
}

extern "C" int rt_multi_set_step_history(rt_multi *m, double *energy, double *temperature, int capacity_steps)
{
   if (!m) return mfail("rt_multi_set_step_history: null handle");
   const int ns = (int)m->h.size();
   m->hist_E = nullptr;
   m->hist_cap = 0;
   m->hist_part.clear();
   if (ns == 1) return rt_set_step_history(m->h[0], energy, temperature, capacity_steps);
   if (energy)
   {
      if (capacity_steps < 1) return mfail("rt_multi_set_step_history: capacity_steps < 1");
      m->hist_E = energy;
      m->hist_cap = capacity_steps;
      m->hist_part.assign(ns, std::vector<double>((size_t)capacity_steps * m->p.N, 0.0));
   }
   return for_each_shard(m, [&](int k) {
      return rt_set_step_history(m->h[k], energy ? m->hist_part[k].data() : nullptr, k == 0 ? temperature : nullptr, capacity_steps);
   });
}

extern "C" int rt_multi_reset(rt_multi *m)
{
   if (!m) return mfail("rt_multi_reset: null handle");
   return for_each_shard(m, [&](int k) { return rt_reset(m->h[k]); });
}

extern "C" int rt_multi_stages_done(const rt_multi *m, int *n)
{
   if (!m || !n) return mfail("rt_multi_stages_done: null argument");
   return rt_stages_done(m->h[0], n);
}

// device time of exactly `timed_steps` steps: every shard is timed on its own device with CUDA events, the slowest counts
extern "C" int rt_multi_bench_steps(rt_multi *m, int warmup_steps, int timed_steps, double *device_ms)
{
   if (!m || !device_ms) return mfail("rt_multi_bench_steps: null argument");
   std::vector<double> ms(m->h.size(), 0.0);
   if (for_each_shard(m, [&](int k) { return rt_bench_steps(m->h[k], warmup_steps, timed_steps, &ms[k]); })) return 1;
   *device_ms = *std::max_element(ms.begin(), ms.end());
   return 0;
}

extern "C" int rt_multi_get_timing(const rt_multi *m, int shard, rt_timing *t)
{
   if (!m || !t || shard < 0 || shard >= (int)m->h.size()) return mfail("rt_multi_get_timing: bad argument");
   return rt_get_timing(m->h[shard], t);
}

extern "C" int rt_multi_get_plan(const rt_multi *m, int shard, rt_plan *plan)
{
   if (!m || !plan || shard < 0 || shard >= (int)m->h.size()) return mfail("rt_multi_get_plan: bad argument");
   return rt_get_plan(m->h[shard], plan);
}

// ---- results in the reference's index order over ALL owned groups (include/rt_cuda.h):
//   phi/F(g,c) at g + G*c ; psi(i,g,c) at i + M*(g + G*c) ; ends(i,g,c,s) at i + M*(g + G*(c + N*s))     G = owned groups
// A shard returns the same layout with its own group count: rows of `unit * Gl` values per (c[,s]) go to their group offset.
static void scatter_rows(const double *src, double *dst, size_t rows, size_t unit, int gl, int G, int goff)
{
   for (size_t r = 0; r < rows; ++r) std::memcpy(dst + unit * ((size_t)G * r + goff), src + unit * (size_t)gl * r, unit * gl * sizeof(double));
}

extern "C" int rt_multi_moments(rt_multi *m, double *phi, double *phi_plus, double *F)
{
   if (!m) return mfail("rt_multi_moments: null handle");
   const size_t N = m->p.N;
   return for_each_shard(m, [&](int k) -> int {
      const size_t n = (size_t)m->gn[k] * N;
      std::vector<double> a(phi ? n : 0), b(phi_plus ? n : 0), c(F ? n : 0);
      if (rt_moments(m->h[k], phi ? a.data() : nullptr, phi_plus ? b.data() : nullptr, F ? c.data() : nullptr)) return 1;
      const int goff = m->g0[k] - m->p.group_begin;
      if (phi) scatter_rows(a.data(), phi, N, 1, m->gn[k], m->G_owned, goff);
      if (phi_plus) scatter_rows(b.data(), phi_plus, N, 1, m->gn[k], m->G_owned, goff);
      if (F) scatter_rows(c.data(), F, N, 1, m->gn[k], m->G_owned, goff);
      return 0;
   });
}

extern "C" int rt_multi_get_psi(rt_multi *m, double *psi)
{
   if (!m || !psi) return mfail("rt_multi_get_psi: null argument");
   const size_t N = m->p.N, M = m->p.M;
   return for_each_shard(m, [&](int k) -> int {
      std::vector<double> a(M * m->gn[k] * N);
      if (rt_get_psi(m->h[k], a.data())) return 1;
      scatter_rows(a.data(), psi, N, M, m->gn[k], m->G_owned, m->g0[k] - m->p.group_begin);
      return 0;
   });
}

extern "C" int rt_multi_get_ends(rt_multi *m, int which, double *ends)
{
   if (!m || !ends) return mfail("rt_multi_get_ends: null argument");
   const size_t N = m->p.N, M = m->p.M;
   return for_each_shard(m, [&](int k) -> int {
      std::vector<double> a(2 * M * m->gn[k] * N);
      if (rt_get_ends(m->h[k], which, a.data())) return 1;
      scatter_rows(a.data(), ends, 2 * N, M, m->gn[k], m->G_owned, m->g0[k] - m->p.group_begin);
      return 0;
   });
}

extern "C" int rt_multi_balance(rt_multi *m, double *balance)
{
   if (!m || !balance) return mfail("rt_multi_balance: null argument");
   return for_each_shard(m, [&](int k) { return rt_balance(m->h[k], balance + (m->g0[k] - m->p.group_begin)); });
}

extern "C" int rt_multi_group_ends(rt_multi *m, double *left, double *right)
{
   if (!m) return mfail("rt_multi_group_ends: null handle");
   return for_each_shard(m, [&](int k) {
      const int goff = m->g0[k] - m->p.group_begin;
      return rt_group_ends(m->h[k], left ? left + goff : nullptr, right ? right + goff : nullptr);
   });
}

// G-length tables are replicated on every shard (all G groups): shard 0 answers
extern "C" int rt_multi_get_group_terms(rt_multi *m, double *B, double *dBdT, double *dEB, double *dsigEdE, double *dkapEB, double *kappa_edge)
{
   if (!m) return mfail("rt_multi_get_group_terms: null handle");
   return rt_get_group_terms(m->h[0], B, dBdT, dEB, dsigEdE, dkapEB, kappa_edge);
}

extern "C" int rt_multi_validate(rt_multi *m, double *sums3, int *ok)
{
   if (!m) return mfail("rt_multi_validate: null handle");
   return rt_validate(m->h[0], sums3, ok);
}

// the per-group tables hold all G groups on every shard: shard 0 answers
extern "C" int rt_multi_correction_print(rt_multi *m, char *text, size_t capacity, size_t *length)
{
   if (!m) return mfail("rt_multi_correction_print: null handle");
   return rt_correction_print(m->h[0], text, capacity, length);
}

// T(c) is identical on every shard after the all-reduce: shard 0 answers
extern "C" int rt_multi_get_temperature(rt_multi *m, double *T, double *sums3)
{
   if (!m) return mfail("rt_multi_get_temperature: null handle");
   return rt_get_temperature(m->h[0], T, sums3);
}
