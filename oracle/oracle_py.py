"""oracle/oracle_py.py -- TEST INFRASTRUCTURE.

ctypes access to the CPU oracles:
  * oracle-2  librt_oracle.so  (rt_oracle.cpp, the CPU restatement)          -> class Oracle
  * oracle-1  _ref/ref_dump    (the UNMODIFIED reference compiled in place)  -> run_reference()

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  The product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import struct
import subprocess
import tempfile
from typing import Dict, Optional

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "librt_oracle.so")
REF_DUMP = os.path.join(HERE, "_ref", "ref_dump")

_dp = C.POINTER(C.c_double)


class _Params(C.Structure):
    _fields_ = [
        ("M", C.c_int), ("N", C.c_int), ("G", C.c_int),
        ("ts_method", C.c_int), ("max_timesteps", C.c_int),
        ("bc_left", C.c_int), ("bc_right", C.c_int),
        ("use_correction", C.c_int), ("use_mg_equilib", C.c_int), ("num_threads", C.c_int),
        ("dx", C.c_double), ("dt", C.c_double), ("rho", C.c_double), ("T", C.c_double),
        ("V", C.c_double), ("kappa_grey", C.c_double),
        ("mu", _dp), ("wt", _dp), ("e_edge", _dp), ("kappa", _dp), ("psi_source", _dp),
        ("rho_cell", _dp), ("T_cell", _dp), ("kappa_cell", _dp),
        ("cv", C.c_double),
    ]


_lib = None


def build(force: bool = False) -> None:
    if force or not os.path.exists(LIB_PATH):
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB_PATH)
        L.rto_create.restype = C.c_void_p
        L.rto_create.argtypes = [C.POINTER(_Params)]
        for name in ("rto_destroy",):
            getattr(L, name).argtypes = [C.c_void_p]
        L.rto_total_stages.argtypes = [C.c_void_p]
        L.rto_stages_done.argtypes = [C.c_void_p]
        L.rto_run_stages.argtypes = [C.c_void_p, C.c_int]
        L.rto_get_psi.argtypes = [C.c_void_p, _dp]
        L.rto_get_ends.argtypes = [C.c_void_p, C.c_int, _dp]
        L.rto_moments.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.rto_balance.argtypes = [C.c_void_p, _dp]
        L.rto_group_ends.argtypes = [C.c_void_p, _dp, _dp]
        L.rto_get_group_terms.argtypes = [C.c_void_p] + [_dp] * 6
        L.rto_get_psi_source.argtypes = [C.c_void_p, _dp]
        L.rto_validate.argtypes = [C.c_void_p, _dp]
        L.rto_get_temperature.argtypes = [C.c_void_p, _dp, _dp, _dp, _dp]
        L.rto_glquad.argtypes = [C.c_int, C.c_double, _dp, _dp]
        L.rto_group_edges.argtypes = [C.c_int, C.c_double, C.c_double, _dp]
        L.rto_planck.argtypes = [C.c_double, C.c_int, _dp, _dp, _dp]
        L.rto_multigroup_opacities.argtypes = [C.c_double, C.c_double, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(_dp)


def glquad(M: int, norm: float):
    mu, wt = np.zeros(M), np.zeros(M)
    lib().rto_glquad(M, norm, _p(mu), _p(wt))
    return mu, wt


def group_edges(G: int, efirst: float, elast: float):
    e = np.zeros(G + 1)
    lib().rto_group_edges(G, efirst, elast, _p(e))
    return e


def planck(T: float, e_edge: np.ndarray):
    e_edge = np.ascontiguousarray(e_edge, dtype=np.float64)
    G = e_edge.size - 1
    B, dB = np.zeros(G), np.zeros(G)
    lib().rto_planck(T, G, _p(e_edge), _p(B), _p(dB))
    return B, dB


def multigroup_opacities(T: float, kappa_grey: float, e_edge: np.ndarray, B: np.ndarray):
    """correction.cpp:66-97: returns dict(kappa, emis_spec, ukappa, ckappa)"""
    e_edge = np.ascontiguousarray(e_edge, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    G = B.size
    out = {k: np.zeros(G) for k in ("kappa", "emis_spec", "ukappa", "ckappa")}
    lib().rto_multigroup_opacities(T, kappa_grey, G, _p(e_edge), _p(B), *[_p(out[k]) for k in ("kappa", "emis_spec", "ukappa", "ckappa")])
    return out


PLANCKIAN = os.path.join(HERE, "_ref", "test_planckian")


def run_test_planckian(M, X, N, G, efirst, elast, T, kappa_grey):
    """Drive the reference's interactive tests/test_planckian.cpp (compiled unmodified, oracle/_ref/test_planckian) through
    stdin and parse the tables it prints (6 significant digits)."""
    inp = "\n".join(str(v) for v in (M, X, N, G, efirst, elast, T, kappa_grey)) + "\n"
    txt = subprocess.run([PLANCKIAN], input=inp, capture_output=True, text=True, timeout=120).stdout
    lines = txt.splitlines()

    def table(header_word, ncols, rows, first="Group"):
        for i, ln in enumerate(lines):
            f = ln.split()
            if f and f[0] == first and header_word in ln:
                out = []
                for r in lines[i + 2:i + 2 + rows]:
                    out.append([float(x) for x in r.split()[:ncols]])
                return np.array(out)
        raise KeyError(header_word)

    t_planck = table("dBdT", 5, G)
    t_uk = table("Planck-Averaged Opacity", 4, G)
    t_fin = table("Final Opacity", 4, G)
    t_deb = table("dEB", 3, G)
    t_edge = table("Opacity", 3, G + 1, first="Edge")
    t_dkde = table("dkapEdE", 3, G)
    t_dkeb = table("dkapEB", 3, G)
    t_cor = table("cor1", 5, G)
    return {"B": t_planck[:, 3], "dBdT": t_planck[:, 4], "ukappa": t_uk[:, 2], "ckappa": t_uk[:, 3], "kappa": t_fin[:, 2],
            "emis_spec": t_fin[:, 3], "dEB": t_deb[:, 2], "kappa_edge": t_edge[:, 2], "dkapEdE": t_dkde[:, 2],
            "dkapEB": t_dkeb[:, 2], "cor1": t_cor[:, 2], "cor2": t_cor[:, 3], "cor3": t_cor[:, 4], "e_ave": t_uk[:, 1]}


class Oracle:
    """oracle-2 solver.  Arrays come back in the REFERENCE's index order, exposed as numpy
    arrays indexed [i, g, c(, s)] (Fortran order views)."""

    def __init__(self, *, M, N, G, ts_method, max_timesteps, bc_left, bc_right, use_correction, use_mg_equilib,
                 dx, dt, rho, T, V, kappa_grey, mu, wt, e_edge, kappa, psi_source=None, num_threads=1,
                 rho_cell=None, T_cell=None, kappa_cell=None, cv=0.0):
        """rho_cell[N], T_cell[N], kappa_cell[G, N]: the cell-dependent material extension; all None =
        the reference's uniform slab.  cv > 0: material-temperature update after every time step (parity
        unpinned -- the reference has none)"""
        self.M, self.N, self.G = M, N, G
        self._keep = [np.ascontiguousarray(a, dtype=np.float64) for a in (mu, wt, e_edge, kappa)]
        ps = None
        if psi_source is not None:
            ps = np.asfortranarray(np.asarray(psi_source, dtype=np.float64).reshape(M, G))
            self._keep.append(ps)
        prm = _Params(M, N, G, ts_method, max_timesteps, bc_left, bc_right, int(bool(use_correction)),
                      int(bool(use_mg_equilib)), num_threads, dx, dt, rho, T, V, kappa_grey,
                      _p(self._keep[0]), _p(self._keep[1]), _p(self._keep[2]), _p(self._keep[3]),
                      ps.ctypes.data_as(_dp) if ps is not None else None)
        for name, a, shape in (("rho_cell", rho_cell, (N,)), ("T_cell", T_cell, (N,)), ("kappa_cell", kappa_cell, (G, N))):
            if a is not None:
                a = np.asfortranarray(np.asarray(a, dtype=np.float64).reshape(shape))
                self._keep.append(a)
                setattr(prm, name, a.ctypes.data_as(_dp))
        prm.cv = float(cv)
        self._h = lib().rto_create(C.byref(prm))

    def close(self):
        if self._h:
            lib().rto_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def total_stages(self):
        return lib().rto_total_stages(self._h)

    @property
    def stages_done(self):
        return lib().rto_stages_done(self._h)

    def run_stages(self, n):
        lib().rto_run_stages(self._h, int(n))

    def solve(self):
        self.run_stages(self.total_stages - self.stages_done)

    def psi(self):
        a = np.zeros((self.M, self.G, self.N), order="F")
        lib().rto_get_psi(self._h, _p(a))
        return a

    def ends(self, which=0):
        a = np.zeros((self.M, self.G, self.N, 2), order="F")
        lib().rto_get_ends(self._h, which, _p(a))
        return a

    def temperature(self, with_sums=False):
        """T(c); with_sums also returns the group sums A, E, D of the last temperature update"""
        T, A, E, D = (np.zeros(self.N) for _ in range(4))
        lib().rto_get_temperature(self._h, _p(T), _p(A), _p(E), _p(D))
        return (T, A, E, D) if with_sums else T

    def moments(self):
        phi, pp, F = (np.zeros((self.G, self.N), order="F") for _ in range(3))
        lib().rto_moments(self._h, _p(phi), _p(pp), _p(F))
        return phi, pp, F

    def balance(self):
        b = np.zeros(self.G)
        lib().rto_balance(self._h, _p(b))
        return b

    def group_ends(self):
        l, r = np.zeros(self.G), np.zeros(self.G)
        lib().rto_group_ends(self._h, _p(l), _p(r))
        return l, r

    def group_terms(self):
        G = self.G
        out = {k: np.zeros(G) for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB")}
        out["kappa_edge"] = np.zeros(G + 1)
        lib().rto_get_group_terms(self._h, *[_p(out[k]) for k in ("B", "dBdT", "dEB", "dsigEdE", "dkapEB", "kappa_edge")])
        return out

    def psi_source(self):
        a = np.zeros((self.M, self.G), order="F")
        lib().rto_get_psi_source(self._h, _p(a))
        return a

    def validate(self):
        s = np.zeros(3)
        ok = lib().rto_validate(self._h, _p(s))
        return bool(ok), s


def oracle_from_parameters(p, mu, wt, e_edge, kappa, num_threads=1) -> Oracle:
    """Build an Oracle from a radiative-transfer_b200.prm.Parameters the way Solver's ctor does
    (solver.cpp:57-73,96-100,145-152)."""
    need_src = p.bc_left_indicator == 1 or p.bc_right_indicator == 1
    return Oracle(M=p.M, N=p.N, G=p.G, ts_method=p.ts_method, max_timesteps=p.max_timesteps,
                  bc_left=p.bc_left_indicator, bc_right=p.bc_right_indicator, use_correction=p.use_correction,
                  use_mg_equilib=p.use_mg_equilib, dx=p.dx, dt=p.dt, rho=p.rho, T=p.T, V=p.V,
                  kappa_grey=p.kappa_grey, mu=mu, wt=wt, e_edge=e_edge, kappa=kappa,
                  psi_source=p.psi_source if need_src else None, num_threads=num_threads,
                  rho_cell=getattr(p, "rho_cell", None), T_cell=getattr(p, "T_cell", None),
                  kappa_cell=getattr(p, "kappa_cell", None), cv=getattr(p, "cv", 0.0))


# ---------------------------------------------------------------- oracle-1 (the reference itself)

def read_dump(path: str) -> Dict[str, np.ndarray]:
    """Parse the record stream written by oracle/ref_dump.cpp (Fortran-order arrays)."""
    out: Dict[str, np.ndarray] = {}
    with open(path, "rb") as fh:
        data = fh.read()
    off = 0
    while off < len(data):
        name = data[off:off + 32].split(b"\0", 1)[0].decode()
        off += 32
        (nd,) = struct.unpack_from("<q", data, off)
        off += 8
        dims = struct.unpack_from("<%dq" % nd, data, off)
        off += 8 * nd
        n = int(np.prod(dims))
        arr = np.frombuffer(data, dtype="<f8", count=n, offset=off).copy()
        off += 8 * n
        out[name] = arr.reshape(dims, order="F")
    return out


def reference_available() -> bool:
    return os.path.exists(REF_DUMP) and os.access(REF_DUMP, os.X_OK)


def run_reference(prm_path: str, data_dir: Optional[str] = None, timeout: Optional[float] = None) -> Dict[str, np.ndarray]:
    """Run the unmodified reference on one .prm and return its dumped buffers.

    The reference opens its table files as ../prm/<name> relative to the cwd, so a scratch
    directory <tmp>/run is used as cwd with <tmp>/prm holding copies of the tables found next
    to the .prm (or in data_dir)."""
    if not reference_available():
        raise RuntimeError("oracle/_ref/ref_dump not built (make -C oracle ref needs /root/reference)")
    src_dir = data_dir or os.path.dirname(os.path.abspath(prm_path))
    tmp = tempfile.mkdtemp(prefix="rtref_")
    try:
        os.makedirs(os.path.join(tmp, "run"))
        os.makedirs(os.path.join(tmp, "prm"))
        for f in os.listdir(src_dir):
            if f.endswith(".txt") or f.endswith(".prm"):
                shutil.copy(os.path.join(src_dir, f), os.path.join(tmp, "prm", f))
        local_prm = os.path.join(tmp, "prm", os.path.basename(prm_path))
        if not os.path.exists(local_prm):
            shutil.copy(prm_path, local_prm)
        out_bin = os.path.join(tmp, "run", "out.bin")
        with open(os.path.join(tmp, "run", "stdout.log"), "w") as log:
            subprocess.check_call([REF_DUMP, local_prm, out_bin], cwd=os.path.join(tmp, "run"), stdout=log,
                                  stderr=subprocess.STDOUT, timeout=timeout)
        return read_dump(out_bin)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
