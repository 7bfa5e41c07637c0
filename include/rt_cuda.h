/* include/rt_cuda.h -- C ABI of librt_cuda.so: the B200 (sm_100a) implementation of the
 * radiative-transfer miniapp's hot path (time-dependent multigroup S_N slab solve).
 *
 * The reference has no FFI/plugin boundary; its hot layer is the C++ class rt::Solver
 * (reference include/solver.h:76-98).  This ABI sits directly UNDER that class surface: the
 * drop-in rt::Solver shipped in radiative-transfer_b200/host/ (and the Python mirror in
 * radiative-transfer_b200/solver.py) forwards each public method to one entry point below.
 * Plain C types only; all arrays are fp64; every pointer argument is a HOST pointer owned by
 * the caller; results are written in the REFERENCE's own index order (Eigen column-major):
 *
 *     psi(i,g,c)      at i + M*(g + Gl*c)
 *     ends(i,g,c,s)   at i + M*(g + Gl*(c + N*s))    s=0 left node, s=1 right node
 *     phi/F(g,c)      at g + Gl*c
 *     psi_source(i,g) at i + M*g                     (g global, 0..G-1)
 *
 * where Gl = group_count is the number of groups THIS handle owns (energy groups are the
 * sharding unit across GPUs and the batching unit inside one GPU; they never couple).
 *
 * Every function returns 0 on success, non-zero on failure; rt_last_error() gives the message.
 * There is no CPU fallback: rt_create fails if no CUDA device is usable.
 * Thread-safety: a handle must be used from one host thread at a time (like rt::Solver).
 */
#ifndef RT_CUDA_H
#define RT_CUDA_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 5

typedef struct rt_params
{
   /* sizes: replaces ParameterHandler::get_M/get_N/get_G (reference solver.cpp:57-59) */
   int M;                /* directions, even                                             */
   int N;                /* cells                                                        */
   int G;                /* energy groups of the whole problem                           */
   int group_begin;      /* first group owned by this handle (0 for a single-GPU solve)  */
   int group_count;      /* number of groups owned (G for a single-GPU solve)            */
   /* ParameterHandler getters used in the hot loop (solver.cpp:338,620,641,667,706) */
   int ts_method;        /* 1 BE, 2 CN, 3 BDF2                                           */
   int max_timesteps;
   int bc_left;          /* 0 vacuum (falls through to source), 1 source, 2 reflective */
   int bc_right;         /* 0 vacuum, 1 source, 2 reflective (=0 inflow, solver.cpp:647-652)      */
   int use_correction;
   int use_mg_equilib;
   int validate;         /* include_validation: rt_solve fails if the sums are off (solver.cpp:609-612) */
   int device;           /* CUDA device ordinal                                          */
   int groups_per_batch; /* 0 = automatic: all owned groups if they fit in HBM, else the largest batch that fits */
   double dx, dt, rho, T, V, kappa_grey;
   const double *mu;         /* M    GLQuad nodes, ascending (solver.cpp:76-78)           */
   const double *wt;         /* M    GLQuad weights                                       */
   const double *e_edge;     /* G+1  group edges (solver.cpp:96-100)                      */
   const double *kappa;      /* G    group opacities (solver.cpp:145-152)                 */
   const double *psi_source; /* M*G  boundary sources or NULL (=0) (solver.cpp:67-73)     */
   /* EXTENSION (ABI version 2).  The reference's slab is uniform: rho, T, kappa_g are filled with
    * constants and the spatial variation is a TODO (solver.cpp:150-157, correction.h:91).  All three NULL = that uniform
    * slab, the pinned path.  Any non-NULL switches the solve to the cell-dependent kernels (csrc/var_kernel.cu; missing
    * fields are filled with the scalars above): per-stage launches only, all owned groups resident, validate must be 0.
    * Equilibrium boundary sources use the state of the boundary cell (cell 0 for mu > 0, cell N-1 for mu < 0).
    * Parity: checked against the oracle's cell-dependent restatement, which is bit-identical to the reference-pinned
    * uniform oracle for uniform arrays; beyond uniform inputs there is no reference to compare with. */
   const double *rho_cell;   /* N    or NULL */
   const double *T_cell;     /* N    or NULL */
   const double *kappa_cell; /* G*N at g + G*c, or NULL */
   /* EXTENSION (ABI version 3): material-temperature update.  PARITY UNPINNED -- the reference has no
    * temperature update at all (its T is a scalar that never changes); the definition is the oracle's
    * restatement (oracle/rt_oracle.cpp, rto_temperature_update), against which the device path is checked at 1e-9.
    * cv > 0 (specific heat) switches it on, and with it the cell-dependent kernels.  After every time step (BDF2: 4 stage
    * sweeps), per cell c with W = sum_m w_m:
    *    A = sum_g sigma_g(c) phi_g(c)     E = W sum_g sigma_g(c) B_g(T(c))     D = W sum_g sigma_g(c) dB_g/dT(T(c))
    *    first step or ts_method != 3:  T' = T + dt (A - E) / (rho cv + dt D)                       (backward Euler, linearised)
    *    later BDF2 steps:              T' = [rho cv (2T - T_prev/2) + dt (A - E + D T)] / (3/2 rho cv + dt D)
    * then the per-cell Planck integrals and correction terms are re-evaluated at T'.  The three group sums are the ONE
    * place where energy groups couple: a handle that owns a group shard forms partial sums and hands the 3*N device
    * buffer to the all-reduce hook (rt_set_allreduce) before the update -- NCCL over NVLink in the multi-GPU runs. */
   double cv;
} rt_params;

typedef struct rt_solver rt_solver;

typedef struct rt_timing
{
   double sweep_ms[3];        /* device time spent in BE / CN / BDF stage-sweep launches (CUDA events) */
   long long sweep_launches[3];
   double moments_ms;
   long long stages;          /* stage sweeps executed (per batch, summed)                */
   long long unknown_stages;  /* M * groups * N summed over executed stage sweeps         */
   long long kernel_launches; /* all kernels launched by the library since rt_create      */
   long long h2d_bytes, d2h_bytes;
   double step_ms;            /* device time spent in fused BDF2-step launches (4 stage sweeps each)  */
   long long step_launches;
   /* material-temperature update (extension): moments + group sums + all-reduce hook + T update ; Planck + per-cell tables */
   double tupdate_ms, tables_ms;
   long long tupdate_count;
} rt_timing;

typedef struct rt_plan
{
   int groups_per_batch, n_batches, resident; /* resident = 1 when one batch holds every owned group */
   int tile_cells, tiles_per_row, sweep_grid, sweep_block; /* grid / block of the most recent sweep or step launch */
   size_t bytes_per_group, state_bytes, free_bytes, total_bytes;
   int fused;            /* 1: whole BDF2 steps run as one fused launch (RT_CUDA_FUSED=0 turns it off)          */
   int step_tile_cells;  /* cells per ticket of the fused step kernel (sub-tile x chained warps)                 */
   int step_block;       /* threads per CTA of the fused step kernel                                             */
   int step_chain;       /* warps chained per CTA tile (1 for the CTA-wide tile shapes)                          */
   int time_levels;      /* time-level buffers allocated: 2 for the fused step, 4 once a step has run stage by stage (ABI 4) */
   int step_kernel;      /* fused step kernel a whole step would use now: 1 = four scalar scans (plain steps), 4 = one 4-vector
                          * scan that also sums phi (per-step output on); RT_CUDA_STEP_KERNEL=1|4 forces one                */
   int var_scan_steps;   /* cell-dependent fused step: warp-scan steps that are run, 5 = all; 2 when no row of the solver can
                          * feel more (small time steps; RT_CUDA_VAR_SCAN_STEPS=5 forces all) (ABI 5)                       */
} rt_plan;

const char *rt_last_error(void);
int rt_abi_version(void);

/* rt::Solver::Solver (solver.cpp:46-188): uploads the tables, evaluates the Planck integrals and
 * per-group correction terms on the device, initialises psi = ends = B_g. */
int rt_create(const rt_params *p, rt_solver **out);
int rt_destroy(rt_solver *s);
int rt_get_plan(const rt_solver *s, rt_plan *plan);

/* rt::Solver::solve (solver.cpp:590-823): all max_timesteps*(ts_method==3?4:1) stage sweeps.
 * Works in resident and in batched mode. */
int rt_solve(rt_solver *s);
/* n further stage sweeps (resident mode only) -- the per-step hook the parity tests use. */
int rt_run_stages(rt_solver *s, int n);
int rt_stages_done(const rt_solver *s, int *n);
int rt_total_stages(const rt_solver *s, int *n);
/* re-initialise psi = ends = B_g and the stage counter (new run on the same tables) */
int rt_reset(rt_solver *s);

/* caller-owned result buffers (solver.h:30-33; main.cc:88-98) */
int rt_get_psi(rt_solver *s, double *psi);                 /* M*Gl*N ; resident mode               */
int rt_get_ends(rt_solver *s, int which, double *ends);    /* which: 0 ends, 1 prev_ends, 2 half_ends ; M*Gl*N*2 ; resident mode */
/* compute_angle_integrated_intensity / compute_positive_... / compute_radiative_flux
 * (solver.cpp:191-237); each Gl*N, any pointer may be NULL */
int rt_moments(rt_solver *s, double *phi, double *phi_plus, double *F);
/* same three moments in the device-native group-major order g*N + c (no layout conversion) */
int rt_moments_group_major(rt_solver *s, double *phi, double *phi_plus, double *F);
int rt_balance(rt_solver *s, double *balance);             /* Gl ; compute_balance, solver.cpp:240-284 */
int rt_group_ends(rt_solver *s, double *left, double *right); /* Gl each ; compute_group_ends, solver.cpp:826-850 */
/* per-group terms of the Correction class for ALL G groups (correction.cpp:25-36,125-249,328-340):
 * B, dBdT, dEB, dsigEdE, dkapEB (G each), kappa_edge (G+1); any pointer may be NULL */
int rt_get_group_terms(rt_solver *s, double *B, double *dBdT, double *dEB, double *dsigEdE, double *dkapEB,
                       double *kappa_edge);
int rt_get_psi_source(rt_solver *s, double *psi_source);   /* M*G */
/* Correction::Print (correction.cpp:404-603): the class's text report -- energy grid, Planck integrals, opacities, edge
 * opacities, dEB, dsigEdE, cor1..3 and the zero-sum check of dkapEB -- formatted from the device's tables with the
 * reference's titles, column widths and number format (6 significant digits).  No reference driver calls Print; the
 * drop-in offers it as Solver::print_correction.  text may be NULL (then only *length is set); the text is
 * NUL-terminated, *length excludes the terminator; a buffer that is too small is an error.  The three columns the
 * reference never fills (ukappa, ckappa, emis_spec: generate_multigroup_opacities is commented out, correction.cpp:376)
 * print 0. */
int rt_correction_print(rt_solver *s, char *text, size_t capacity, size_t *length);
/* Correction::validate_correction (correction.cpp:39-63,100-122): sums3 = {sum B, sum dBdT, sum kappa*B};
 * *ok = 1 when all three identities hold to 1e-6 */
int rt_validate(rt_solver *s, double *sums3, int *ok);

/* Material-temperature update (extension, rt_params::cv).  rt_get_temperature: the current T(c), N doubles; sums3 (may be
 * NULL) receives the group sums A, E, D of the most recent update, 3*N doubles, AFTER the all-reduce.
 * rt_set_allreduce: `fn(buf, count, stream, user)` must sum the `count` fp64 values at DEVICE pointer `buf` element-wise
 * over all handles that share the problem (in place, result on every rank) and order that work after what `stream`
 * (a cudaStream_t) holds and before what is submitted to it next; non-zero return aborts the solve.  NULL = the handle
 * owns every group.  A group shard calls it TWICE per time step: 3*N values (the group sums A, E, D) before the update and
 * 2*N values after it (per-cell sums of the shard's Planck integrals at T': the last group is the grey remainder over all
 * groups, Planck.cpp:58-76, so the integrals are sharded like the sweeps and only these sums travel).  The library itself does not link NCCL: the Python host passes torch.distributed's all_reduce, a
 * C++ host its ncclAllReduce. */
typedef int (*rt_allreduce_fn)(void *device_buf, size_t count, void *cuda_stream, void *user);
int rt_set_allreduce(rt_solver *s, rt_allreduce_fn fn, void *user);
int rt_get_temperature(rt_solver *s, double *T, double *sums3);

/* Native NCCL for that exchange ("the library owns streams, device buffers and NCCL communicators"; the
 * reference itself has no collective).  For hosts that run ONE PROCESS PER GPU: rank 0 calls
 * rt_nccl_unique_id (128 bytes, ncclUniqueId), the host distributes the id (MPI_Bcast, torch.distributed.broadcast ...),
 * every rank calls rt_set_nccl on its handle after rt_create: the library creates the communicator on the handle's device
 * and sums the group sums with ncclAllReduce on its own stream -- no host callback in the time loop.  libnccl.so.2 is
 * loaded with dlopen on first use (RT_CUDA_NCCL_LIB overrides the name); a process that never asks needs no NCCL.
 * The communicator belongs to the handle (rt_destroy releases it); rt_set_allreduce replaces it with the host's own sum. */
int rt_nccl_unique_id(void *id128);
int rt_set_nccl(rt_solver *s, const void *id128, int nranks, int rank);

/* ---- ONE process, SEVERAL GPUs (csrc/rt_multi.cu).  The reference's drivers construct one rt::Solver in one process
 * (src/main.cc:100-113, tests/test_gray.cpp:68-79); for that host to use more than one B200 the sharding lives here
 * ("the library owns streams, device buffers and NCCL communicators (single-process multi-GPU)").
 * rt_multi_create: `p` describes the whole problem (p->device is ignored); the groups [group_begin, group_begin+group_count)
 * are cut into ndev contiguous blocks, shard k on devices[k], every shard driven by its own host thread.  A device may be
 * listed twice (two shards on one GPU) unless cv > 0: the temperature update sums over the shards with ncclAllReduce on
 * communicators created here (ncclCommInitAll), one distinct device per shard.  Results come back in the reference's
 * index order over ALL owned groups (G = p->group_count in the layouts at the top of this file).  The drop-in
 * rt::Solver (host/solver.cpp) takes the device list from RT_CUDA_DEVICES. */
typedef struct rt_multi rt_multi;
int rt_multi_create(const rt_params *p, int ndev, const int *devices, rt_multi **out);
int rt_multi_destroy(rt_multi *m);
int rt_multi_shards(const rt_multi *m, int *n, int *group_begin, int *group_count, int *device); /* arrays of n, may be NULL */
int rt_multi_solve(rt_multi *m);                              /* rt_solve on every shard, concurrently                    */
int rt_multi_run_stages(rt_multi *m, int n);
int rt_multi_reset(rt_multi *m);
int rt_multi_stages_done(const rt_multi *m, int *n);
int rt_multi_get_psi(rt_multi *m, double *psi);               /* M*G*N                                                    */
int rt_multi_get_ends(rt_multi *m, int which, double *ends);  /* M*G*N*2                                                  */
int rt_multi_moments(rt_multi *m, double *phi, double *phi_plus, double *F); /* G*N each, any may be NULL               */
int rt_multi_balance(rt_multi *m, double *balance);           /* G                                                        */
int rt_multi_group_ends(rt_multi *m, double *left, double *right);
int rt_multi_get_group_terms(rt_multi *m, double *B, double *dBdT, double *dEB, double *dsigEdE, double *dkapEB, double *kappa_edge);
int rt_multi_validate(rt_multi *m, double *sums3, int *ok);
int rt_multi_get_temperature(rt_multi *m, double *T, double *sums3);
int rt_multi_correction_print(rt_multi *m, char *text, size_t capacity, size_t *length);
/* energy: summed over the shards (ascending); temperature: identical on all shards, shard 0 answers */
int rt_multi_set_step_history(rt_multi *m, double *energy, double *temperature, int capacity_steps);
int rt_multi_bench_steps(rt_multi *m, int warmup_steps, int timed_steps, double *device_ms); /* slowest shard's device time */
int rt_multi_get_timing(const rt_multi *m, int shard, rt_timing *t);
int rt_multi_get_plan(const rt_multi *m, int shard, rt_plan *plan);

/* After every completed time step (BDF2: 4 stage sweeps) copy phi (Gl*N, group-major g*N + c when
 * group_major != 0, reference order otherwise) into `phi_host`; NULL switches it off.  Used by the
 * end-to-end measurement: the device->host read of each step's result is inside rt_solve. */
int rt_set_step_output(rt_solver *s, double *phi_host, int group_major);

/* Measurement: for every batch of owned groups re-initialise, run `warmup_steps` untimed time steps, then time
 * exactly `timed_steps` time steps (BDF2: 4 stage sweeps each) with CUDA events on the launching stream.
 * *device_ms is the sum over batches; rt_timing afterwards covers the timed regions only. */
int rt_bench_steps(rt_solver *s, int warmup_steps, int timed_steps, double *device_ms);
/* Per-step history: the per-step quantities the port's acceptance names besides the scalar intensity (no counterpart in
 * the reference, which keeps nothing per step).  After every completed time step k (0-based, counted from the initial state,
 * k < capacity_steps):
 *   energy[k*N + c]      = (1/c_light) * sum over the groups THIS handle owns of phi_g(c): radiation energy density in the
 *                          reference's normalisation of phi (solver.cpp:191-204; equilibrium value a T^4)
 *   temperature[k*N + c] = T(c) after the update of step k (needs rt_params::cv > 0)
 * Either pointer may be NULL; both NULL switches the history off.  The rows stay on the device while the steps run (group
 * batches add their share) and are copied into the caller's arrays when rt_run_stages / rt_solve returns. */
int rt_set_step_history(rt_solver *s, double *energy, double *temperature, int capacity_steps);

/* pinned host memory for caller-owned result buffers (cudaMallocHost / cudaFreeHost) */
int rt_host_alloc(void **p, size_t bytes);
int rt_host_free(void *p);

int rt_get_timing(const rt_solver *s, rt_timing *t);
int rt_reset_timing(rt_solver *s);

/* Standalone device evaluation of the Planck group integrals (Planck.cpp:50-77 + the unit change of
 * correction.cpp:25-36) for `ncell` temperatures: B,dBdT are [ncell][G] (g fastest). */
int rt_planck_groups(int device, int G, const double *e_edge, int ncell, const double *T, double *B, double *dBdT,
                     double *kernel_ms);

/* NO COUNTERPART IN THE REFERENCE.  rt_planck_groups for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it): same inputs,
 * same outputs (HOST pointers; B, dBdT are [ncell][G], g fastest), same kernels, but it calls no cudaMalloc / cudaFree /
 * allocating library routine.  Its device scratch (copies of e_edge and T, the kernel's two group-major tables, the
 * staging of the layout conversion) is ONE static device array of RT_NOALLOC_MAX_DOUBLES doubles (8 MiB) that every GPU
 * which loads the library holds until the process exits.  Limit: 3*G*ncell + G + 1 + ncell doubles (each part rounded up
 * to 32) must fit, e.g. 256 groups x 1024 temperatures; larger requests fail with a message and nothing is computed.
 * All calls share the array and it keeps the last call's contents (the tables are cleared at the start of every call):
 * calls that overlap on one GPU -- two host threads, two streams -- RACE; serialise them.  The call is synchronous. */
#define RT_NOALLOC_MAX_DOUBLES (1 << 20)
int rt_planck_groups_noalloc(int device, int G, const double *e_edge, int ncell, const double *T, double *B, double *dBdT);

/* Correction::generate_multigroup_opacities (correction.cpp:66-97; kept by the reference but never called -- second
 * statement tests/test_planckian.cpp:267-349): Planck-weighted group opacities at temperature T normalised so that
 * sum_g kappa_g B_g = kappa_grey a c T^4.  Outputs are G long, any may be NULL. */
int rt_multigroup_opacities(int device, int G, const double *e_edge, double T, double kappa_grey, double *kappa,
                            double *emis_spec, double *ukappa, double *ckappa);

#ifdef __cplusplus
}
#endif
#endif /* RT_CUDA_H */
