"""Python mirror of the reference's rt::Solver class surface (reference include/solver.h:76-98),
implemented on librt_cuda.so.  Same method names, argument meaning and error behaviour; the caller
owns psi_mat / phi / F exactly as in reference src/main.cc:88-113.

    ph = prm.load("prm/single_group.prm")
    psi = np.zeros((ph.M, ph.G, ph.N), order="F"); phi = np.zeros((ph.G, ph.N), order="F"); F = phi.copy()
    s = Solver(ph, psi, phi, F)
    s.solve(); s.compute_angle_integrated_intensity(); s.compute_radiative_flux(); s.compute_balance()

Arrays use the reference's index order (Eigen column-major == numpy Fortran order).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import capi, grid
from .prm import Parameters


def _f(shape):
    return np.zeros(shape, dtype=np.float64, order="F")


class Solver:
    def __init__(self, parameter_handler: Parameters, psi_mat: Optional[np.ndarray] = None,
                 phi: Optional[np.ndarray] = None, F: Optional[np.ndarray] = None, *, device: int = 0,
                 group_begin: int = 0, group_count: Optional[int] = None, groups_per_batch: int = 0,
                 fetch_psi: bool = True, dist=None, devices=None):
        ph = self.ph = parameter_handler
        self.M, self.N, self.num_groups = ph.M, ph.N, ph.G
        self.group_begin = group_begin
        self.group_count = ph.G - group_begin if group_count is None else group_count
        Gl = self.group_count
        # solver.cpp:76-78 quadrature, 96-104 energy grid, 145-152 opacities
        self.m_mu, self.m_wt = grid.gauss_legendre(ph.M, grid.FOUR_PI)
        if ph.have_group_bounds:
            self.e_edge = np.ascontiguousarray(ph.group_bounds, dtype=np.float64)
        else:
            self.e_edge = self.generate_group_edges()
        self.e_ave, self.de_ave = self.generate_group_averages()
        if ph.have_group_absorption_opacities:
            self.kappa_vec = np.ascontiguousarray(ph.group_kappa, dtype=np.float64)
        else:
            self.kappa_vec = np.full(ph.G, ph.kappa_grey, dtype=np.float64)
        # solver.cpp:67-73: psi_source is only fetched when some BC is a source; otherwise it is 0
        src = None
        if ph.bc_left_indicator == 1 or ph.bc_right_indicator == 1:
            src = np.asfortranarray(np.asarray(ph.psi_source, dtype=np.float64).reshape(ph.M, ph.G))
        self._src = src
        self._fetch_psi = fetch_psi

        def own(a, shape, name):
            if a is None:
                return _f(shape)
            if a.shape != shape or a.dtype != np.float64 or not a.flags.f_contiguous:
                raise ValueError(f"{name} must be a float64 Fortran-ordered array of shape {shape}")
            return a

        self.psi_mat_ref = own(psi_mat, (ph.M, Gl, ph.N), "psi_mat") if fetch_psi else psi_mat
        self.phi_ref = own(phi, (Gl, ph.N), "phi")
        self.F_ref = own(F, (Gl, ph.N), "F")
        self.phi_plus = _f((Gl, ph.N))
        self.balance = np.zeros(Gl)
        self.left_ends = np.zeros(Gl)
        self.right_ends = np.zeros(Gl)

        p = capi.RtParams()
        p.M, p.N, p.G = ph.M, ph.N, ph.G
        p.group_begin, p.group_count = group_begin, Gl
        p.ts_method, p.max_timesteps = ph.ts_method, ph.max_timesteps
        p.bc_left, p.bc_right = ph.bc_left_indicator, ph.bc_right_indicator
        p.use_correction, p.use_mg_equilib = int(ph.use_correction), int(ph.use_mg_equilib)
        p.validate = int(ph.include_validation)
        p.device, p.groups_per_batch = device, groups_per_batch
        p.dx, p.dt, p.rho, p.T, p.V, p.kappa_grey = ph.dx, ph.dt, ph.rho, ph.T, ph.V, ph.kappa_grey
        p.mu, p.wt = capi.ptr(self.m_mu), capi.ptr(self.m_wt)
        p.e_edge, p.kappa = capi.ptr(self.e_edge), capi.ptr(self.kappa_vec)
        p.psi_source = capi.ptr(src) if src is not None else None
        # extension: cell-dependent material fields; all None = the reference's uniform slab
        self._cell_fields = []
        for name, shape in (("rho_cell", (ph.N,)), ("T_cell", (ph.N,)), ("kappa_cell", (ph.G, ph.N))):
            a = getattr(ph, name, None)
            if a is not None:
                a = np.asfortranarray(np.asarray(a, dtype=np.float64).reshape(shape))  # kappa(g, c) at g + G*c
                self._cell_fields.append(a)
                setattr(p, name, capi.ptr(a))
        # extension: material-temperature update when the specific heat is given
        p.cv = float(getattr(ph, "cv", 0.0) or 0.0)
        self._h = C.c_void_p()
        # devices=[d0, d1, ...]: ONE process, several GPUs -- the library shards the owned groups itself (rt_multi_*,
        # csrc/rt_multi.cu), the way the C++ drop-in does with RT_CUDA_DEVICES; results come back over all owned groups
        self._multi = devices is not None
        if self._multi:
            devs = (C.c_int * len(devices))(*[int(d) for d in devices])
            capi.check(capi.lib().rt_multi_create(C.byref(p), len(devices), devs, C.byref(self._h)))
            self._n_shards = len(devices)
        else:
            capi.check(capi.lib().rt_create(C.byref(p), C.byref(self._h)))
        # a group shard of a temperature-coupled solve: the group sums go through torch.distributed (NCCL) every step
        self._allreduce = None
        self.exchange = "none (this handle owns every group)"
        if self._multi and p.cv > 0.0 and len(devices) > 1:
            self.exchange = "ncclAllReduce inside the library (rt_multi: one communicator per device, one process)"
        elif p.cv > 0.0 and dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
            from . import sharding
            import os
            if os.environ.get("RT_CUDA_ALLREDUCE", "nccl") == "torch":
                # the host's own sum: torch.distributed.all_reduce on the library's buffer, called back from the time loop
                self._allreduce = sharding.make_allreduce_hook(dist, device)
                capi.check(capi.lib().rt_set_allreduce(self._h, self._allreduce, None))
                self.exchange = "torch.distributed all_reduce through the rt_set_allreduce callback"
            else:
                # native: the library creates its own NCCL communicator and calls ncclAllReduce on its stream
                capi.check(capi.lib().rt_set_nccl(self._h, sharding.broadcast_nccl_id(dist, device), dist.get_world_size(), dist.get_rank()))
                self.exchange = "ncclAllReduce inside the library (rt_set_nccl)"
        self.resident = all(self.plan(k).resident == 1 for k in range(self._n_shards)) if self._multi else self.plan().resident == 1
        if fetch_psi and self.resident:
            self._pull_psi()  # initial condition psi = B_g (solver.cpp:166-181)

    def _f(self, name):
        """the entry point `rt_<name>` of a single handle, or `rt_multi_<name>` of a multi-device handle"""
        if self._multi:
            fn = getattr(capi.lib(), "rt_multi_" + name, None)
            if fn is None:
                raise capi.RtError(f"rt_{name} is not available on a multi-device handle (devices=[...])")
            return fn
        return getattr(capi.lib(), "rt_" + name)

    # ------------------------------------------------------------------ lifetime
    def close(self):
        if getattr(self, "_h", None):
            self._f("destroy")(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ------------------------------------------------------------------ reference public surface
    def generate_group_edges(self):
        """solver.cpp:6-19"""
        return grid.group_edges(self.ph.G, self.ph.efirst, self.ph.elast)

    def generate_group_averages(self):
        """solver.cpp:22-32"""
        return grid.group_averages(self.e_edge)

    def fill_energy_bound_arrays(self):
        """solver.cpp:35-43: (G,2) array of [lower, upper] edges"""
        return np.asfortranarray(np.stack([self.e_edge[:-1], self.e_edge[1:]], axis=1))

    def solve(self):
        """solver.cpp:590-823"""
        capi.check(self._f("solve")(self._h))
        if self._fetch_psi and self.resident:
            self._pull_psi()

    def compute_angle_integrated_intensity(self):
        capi.check(self._f("moments")(self._h, capi.ptr(self.phi_ref), None, None))

    def compute_positive_angle_integrated_intensity(self):
        capi.check(self._f("moments")(self._h, None, capi.ptr(self.phi_plus), None))

    def compute_radiative_flux(self):
        capi.check(self._f("moments")(self._h, None, None, capi.ptr(self.F_ref)))

    def compute_balance(self):
        capi.check(self._f("balance")(self._h, capi.ptr(self.balance)))

    def get_balance(self):
        return self.balance.copy()

    def get_phi_plus(self):
        return self.phi_plus.copy(order="F")

    def get_e_ave(self):
        return self.e_ave.copy()

    def compute_group_ends(self):
        capi.check(self._f("group_ends")(self._h, capi.ptr(self.left_ends), capi.ptr(self.right_ends)))

    def get_ends(self, side: str):
        assert side in ("left", "right"), "Invalid option for 'side'."
        return (self.left_ends if side == "left" else self.right_ends).copy()

    # ------------------------------------------------------------------ output files (main.cc:117-133)
    def write_outputs(self, directory: str = ".") -> None:
        """The 8 files the reference's driver writes after solve() (src/main.cc:117-133: phi, phi_plus, psi, x, F, e_ave,
        left_ends, right_ends `.csv`), same layout (one matrix row per line, blank-separated, ostream default precision
        = 6 significant digits; psi as M rows x (G*N) columns, Eigen::Tensor's flattening) so that
        scripts/plot_results.py reads them unchanged.  Calls the compute_* methods itself, like main.cc:105-129."""
        import os

        def fmt(v):
            return "%g" % v

        def dump(name, a2d):
            with open(os.path.join(directory, name), "w") as fh:
                fh.write("\n".join(" ".join(fmt(v) for v in row) for row in a2d) + "\n")

        self.compute_angle_integrated_intensity()
        self.compute_radiative_flux()
        self.compute_balance()
        self.compute_positive_angle_integrated_intensity()
        self.compute_group_ends()
        if self._fetch_psi and self.resident:
            self._pull_psi()
        col = lambda v: np.asarray(v, dtype=np.float64).reshape(-1, 1)
        dump("phi.csv", self.phi_ref)
        dump("phi_plus.csv", self.phi_plus)
        if self.psi_mat_ref is not None and self.resident:
            dump("psi.csv", self.psi_mat_ref.reshape(self.M, self.group_count * self.N, order="F"))
        dump("x.csv", col((np.arange(self.N) + 0.5) * self.ph.dx))
        dump("F.csv", self.F_ref)
        dump("e_ave.csv", col(self.e_ave[self.group_begin:self.group_begin + self.group_count]))
        dump("left_ends.csv", col(self.left_ends))
        dump("right_ends.csv", col(self.right_ends))

    # ------------------------------------------------------------------ extras used by tests / bench
    # ---- per-step snapshot stream: parity "at each step" inspectable without re-running ---------
    SNAPSHOT_MAGIC = b"RTSNAP01"

    def solve_with_snapshots(self, path: str, every: int = 1, with_ends: bool = False) -> int:
        """Run the remaining time steps like solve() and append one record per `every` time steps to the binary file
        `path`: header {magic, M, G, N, group_begin, group_count, stages per step, with_ends} (int64 each after the
        8-byte magic), then per record {time step, stages done} (int64) + phi, phi_plus, F as [group][cell] float64
        (+ ends in the reference's (i, g, cell, node) order when with_ends).  Needs every owned group resident
        (a batched solve runs its batches one after the other through all time steps).  Returns the record count."""
        if not self.resident:
            raise capi.RtError("solve_with_snapshots: the solve is group-batched; snapshots need a resident state")
        per_step = 4 if self.ph.ts_method == 3 else 1
        n = 0
        with open(path, "wb") as f:
            f.write(self.SNAPSHOT_MAGIC)
            np.array([self.M, self.num_groups, self.N, self.group_begin, self.group_count, per_step, int(with_ends)],
                     dtype=np.int64).tofile(f)
            while self.stages_done < self.total_stages:
                todo = min(every * per_step, self.total_stages - self.stages_done)
                capi.check(self._f("run_stages")(self._h, int(todo)))
                phi, pp, F = self.moments_group_major()
                np.array([self.stages_done // per_step, self.stages_done], dtype=np.int64).tofile(f)
                for a in (phi, pp, F):
                    np.ascontiguousarray(a).tofile(f)
                if with_ends:
                    np.asfortranarray(self.ends(0)).ravel(order="F").tofile(f)
                n += 1
        if self._fetch_psi:
            self._pull_psi()
        return n

    @classmethod
    def read_snapshots(cls, path: str):
        """Generator over the records of a snapshot stream: dicts with step, stages, phi, phi_plus, F (+ ends)."""
        with open(path, "rb") as f:
            if f.read(8) != cls.SNAPSHOT_MAGIC:
                raise ValueError(f"{path}: not a snapshot stream")
            M, G, N, g0, gc, per_step, with_ends = (int(v) for v in np.fromfile(f, dtype=np.int64, count=7))
            while True:
                head = np.fromfile(f, dtype=np.int64, count=2)
                if head.size < 2:
                    return
                rec = {"step": int(head[0]), "stages": int(head[1]), "group_begin": g0}
                for k in ("phi", "phi_plus", "F"):
                    rec[k] = np.fromfile(f, dtype=np.float64, count=gc * N).reshape(gc, N)
                if with_ends:
                    rec["ends"] = np.fromfile(f, dtype=np.float64, count=M * gc * N * 2).reshape((M, gc, N, 2), order="F")
                yield rec

    def _pull_psi(self):
        capi.check(self._f("get_psi")(self._h, capi.ptr(self.psi_mat_ref)))

    def run_stages(self, n: int):
        capi.check(self._f("run_stages")(self._h, int(n)))
        if self._fetch_psi and self.resident:
            self._pull_psi()

    def reset(self):
        capi.check(self._f("reset")(self._h))

    @property
    def stages_done(self) -> int:
        n = C.c_int()
        capi.check(self._f("stages_done")(self._h, C.byref(n)))
        return n.value

    @property
    def total_stages(self) -> int:
        if self._multi:
            return self.ph.max_timesteps * (4 if self.ph.ts_method == 3 else 1)
        n = C.c_int()
        capi.check(capi.lib().rt_total_stages(self._h, C.byref(n)))
        return n.value

    def ends(self, which: int = 0):
        a = _f((self.M, self.group_count, self.N, 2))
        capi.check(self._f("get_ends")(self._h, which, capi.ptr(a)))
        return a

    def moments(self):
        phi, pp, F = (_f((self.group_count, self.N)) for _ in range(3))
        capi.check(self._f("moments")(self._h, capi.ptr(phi), capi.ptr(pp), capi.ptr(F)))
        return phi, pp, F

    def moments_group_major(self):
        if self._multi:
            return tuple(np.ascontiguousarray(a) for a in self.moments())
        phi, pp, F = (np.zeros((self.group_count, self.N)) for _ in range(3))
        capi.check(capi.lib().rt_moments_group_major(self._h, capi.ptr(phi), capi.ptr(pp), capi.ptr(F)))
        return phi, pp, F

    def group_terms(self):
        G = self.num_groups
        out = {k: np.zeros(G) for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB")}
        out["kappa_edge"] = np.zeros(G + 1)
        capi.check(self._f("get_group_terms")(
            self._h, *[capi.ptr(out[k]) for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB", "kappa_edge")]))
        return out

    def set_step_history(self, capacity_steps: int, temperature: bool = False):
        """record, for every time step k < capacity_steps, the radiation energy density (1/c) sum_g phi_g(c) -- summed over
        the groups this solver owns: under torchrun each rank holds its share -- and, with cv > 0, the material temperature;
        the arrays ([capacity_steps, N], returned and kept as step_energy / step_temperature) are filled when solve() /
        run_stages() returns"""
        self.step_energy = np.zeros((capacity_steps, self.N))
        self.step_temperature = np.zeros((capacity_steps, self.N)) if temperature else None
        capi.check(self._f("set_step_history")(self._h, capi.ptr(self.step_energy),
                                                capi.ptr(self.step_temperature) if temperature else None, capacity_steps))
        return self.step_energy, self.step_temperature

    def correction_report(self) -> str:
        """the text of the reference's Correction::Print() (correction.cpp:404-603) from the device's tables"""
        n = C.c_size_t()
        capi.check(self._f("correction_print")(self._h, None, 0, C.byref(n)))
        buf = C.create_string_buffer(n.value + 1)
        capi.check(self._f("correction_print")(self._h, buf, n.value + 1, C.byref(n)))
        return buf.value.decode()

    def print_correction(self, file=None):
        print(self.correction_report(), end="", file=file)

    def temperature(self, with_sums: bool = False):
        """extension: the material temperature T(c); with_sums also the group sums A, E, D (absorbed power, emitted
        power, emission derivative; after the all-reduce) of the most recent temperature update"""
        T = np.zeros(self.N)
        sums = np.zeros((3, self.N)) if with_sums else None
        capi.check(self._f("get_temperature")(self._h, capi.ptr(T), capi.ptr(sums) if with_sums else None))
        return (T, sums[0], sums[1], sums[2]) if with_sums else T

    def psi_source(self):
        a = _f((self.M, self.num_groups))
        capi.check(self._f("get_psi_source")(self._h, capi.ptr(a)))
        return a

    def validate(self):
        sums = np.zeros(3)
        ok = C.c_int()
        capi.check(self._f("validate")(self._h, capi.ptr(sums), C.byref(ok)))
        return bool(ok.value), sums

    def set_step_output(self, phi_host_ptr, group_major: bool = True):
        capi.check(self._f("set_step_output")(self._h, phi_host_ptr, int(group_major)))

    def bench_steps(self, warmup_steps: int, timed_steps: int) -> float:
        """device milliseconds of exactly `timed_steps` time steps of all owned groups (summed over batches)"""
        ms = C.c_double()
        capi.check(self._f("bench_steps")(self._h, warmup_steps, timed_steps, C.cast(C.byref(ms), C.POINTER(C.c_double))))
        return ms.value

    def timing(self, shard: int = 0) -> capi.RtTiming:
        t = capi.RtTiming()
        if self._multi:
            capi.check(capi.lib().rt_multi_get_timing(self._h, int(shard), C.byref(t)))
        else:
            capi.check(capi.lib().rt_get_timing(self._h, C.byref(t)))
        return t

    def reset_timing(self):
        capi.check(self._f("reset_timing")(self._h))

    def plan(self, shard: int = 0) -> capi.RtPlan:
        pl = capi.RtPlan()
        if self._multi:
            capi.check(capi.lib().rt_multi_get_plan(self._h, int(shard), C.byref(pl)))
        else:
            capi.check(capi.lib().rt_get_plan(self._h, C.byref(pl)))
        return pl

    def shards(self):
        """[(first group, groups, device)] of a multi-device handle"""
        n = C.c_int()
        capi.check(capi.lib().rt_multi_shards(self._h, C.byref(n), None, None, None))
        g0, gn, dev = ((C.c_int * n.value)() for _ in range(3))
        capi.check(capi.lib().rt_multi_shards(self._h, C.byref(n), g0, gn, dev))
        return [(g0[k], gn[k], dev[k]) for k in range(n.value)]


def planck_groups(e_edge, T, device: int = 0):
    """Device Planck group integrals for an array of temperatures: returns (B, dBdT, kernel_ms), B[cell, g]."""
    e_edge = np.ascontiguousarray(e_edge, dtype=np.float64)
    T = np.ascontiguousarray(np.atleast_1d(T), dtype=np.float64)
    G = e_edge.size - 1
    B = np.zeros((T.size, G))
    dB = np.zeros((T.size, G))
    ms = C.c_double()
    capi.check(capi.lib().rt_planck_groups(device, G, capi.ptr(e_edge), T.size, capi.ptr(T), capi.ptr(B), capi.ptr(dB),
                                           C.cast(C.byref(ms), C.POINTER(C.c_double))))
    return B, dB, ms.value


def multigroup_opacities(e_edge, T: float, kappa_grey: float, device: int = 0):
    """Device version of Correction::generate_multigroup_opacities (correction.cpp:66-97): dict(kappa, emis_spec,
    ukappa, ckappa), each G long."""
    e_edge = np.ascontiguousarray(e_edge, dtype=np.float64)
    G = e_edge.size - 1
    out = {k: np.zeros(G) for k in ("kappa", "emis_spec", "ukappa", "ckappa")}
    capi.check(capi.lib().rt_multigroup_opacities(device, G, capi.ptr(e_edge), float(T), float(kappa_grey),
                                                  *[capi.ptr(out[k]) for k in ("kappa", "emis_spec", "ukappa", "ckappa")]))
    return out
