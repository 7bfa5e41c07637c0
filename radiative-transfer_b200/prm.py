"""Host-side `.prm` parameter files: Python mirror of the reference's ParameterHandler.

Follows the reference's parsing rules exactly (reference include/param.h:61-75,
src/param.cpp:5-66, src/ParameterHandler.cpp:100-212) so that the same `.prm`
files drive the reference, the oracle and the CUDA path:

* lines whose first character is ``#`` are skipped; a line is split at its FIRST ``=``;
  the value is the rest of the line verbatim (trailing blanks/comments included);
  the first occurrence of a key wins (``std::map::insert``).
* ints/doubles are read with ``stoi``/``stod`` semantics: leading blanks skipped, the
  longest numeric prefix is converted, trailing text ignored.
* bools are true only if the value is EXACTLY one of ``yes Yes true True``.
* group-bounds / opacity tables are whitespace-separated doubles in ``../prm/<name>``
  relative to the working directory (ParameterHandler.cpp:141,172).  As a convenience
  the directory of the ``.prm`` file itself is searched when that path does not exist.
"""
from __future__ import annotations

import os
import re
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

_FLOAT_PREFIX = re.compile(
    r"\s*([+-]?(?:(?:\d+\.?\d*|\.\d+)(?:[eE][+-]?\d+)?|inf(?:inity)?|nan))", re.IGNORECASE
)
_INT_PREFIX = re.compile(r"\s*([+-]?\d+)")
_TRUE_STRINGS = ("yes", "Yes", "true", "True")


def _stod(text: str) -> float:
    m = _FLOAT_PREFIX.match(text)
    if not m:
        raise ValueError(f"stod: no conversion could be performed on {text!r}")
    return float(m.group(1))


def _stoi(text: str) -> int:
    m = _INT_PREFIX.match(text)
    if not m:
        raise ValueError(f"stoi: no conversion could be performed on {text!r}")
    return int(m.group(1))


def read_key_values(path: str) -> Dict[str, str]:
    """param.h:61-75 -- returns {} (and the caller falls back to defaults) if the file is missing."""
    params: Dict[str, str] = {}
    try:
        with open(path, "r") as fh:
            for raw in fh:
                line = raw.rstrip("\n")
                if line.startswith("#"):
                    continue
                idx = line.find("=")
                if idx != -1:
                    params.setdefault(line[:idx], line[idx + 1:])
    except OSError:
        print(f"Could not open file {path}")
    return params


def read_doubles(path: str) -> np.ndarray:
    """`while (fin >> d)` -- stops at the first token that is not a number."""
    out: List[float] = []
    with open(path, "r") as fh:
        for tok in fh.read().split():
            try:
                out.append(float(tok))
            except ValueError:
                break
    return np.asarray(out, dtype=np.float64)


@dataclass
class Parameters:
    """Same fields, defaults and derived values as the reference's ParameterHandler
    (include/ParameterHandler.h:17-61, src/ParameterHandler.cpp:102-211)."""

    M: int = 2
    G: int = 1
    efirst: float = 0.1
    elast: float = 10.0
    X: float = 1.0
    N: int = 100
    bc_left_indicator: int = 2
    bc_right_indicator: int = 1
    use_mg_equilib: bool = False
    psi_source: Optional[np.ndarray] = None  # (M, G)
    have_group_bounds: bool = False
    have_group_absorption_opacities: bool = False
    filename_group_bounds: str = "NA"
    filename_group_kappa: str = "NA"
    group_bounds: Optional[np.ndarray] = None  # G+1
    group_kappa: Optional[np.ndarray] = None   # G
    rho: float = 1.0
    kappa_grey: float = 1.0
    T: float = 1.0
    V: float = 0.0
    use_correction: bool = False
    ts_method: int = 3
    dt: float = 0.00001
    max_timesteps: int = 1000
    include_validation: bool = True
    source_file: str = field(default="", compare=False)
    # EXTENSION (the reference's slab is uniform, solver.cpp:150-157 TODO): material fields per cell.
    # All None = the reference's uniform slab; any given switches the solve to the cell-dependent kernels.
    rho_cell: Optional[np.ndarray] = field(default=None, compare=False)    # N
    T_cell: Optional[np.ndarray] = field(default=None, compare=False)      # N
    kappa_cell: Optional[np.ndarray] = field(default=None, compare=False)  # (G, N)
    # EXTENSION (the reference has no temperature update): specific heat of the material.
    # cv > 0 switches on the material-temperature update after every time step (include/rt_cuda.h, rt_params::cv).
    cv: float = 0.0

    @property
    def dx(self) -> float:
        return self.X / self.N

    def __post_init__(self):
        if self.psi_source is None:
            self.psi_source = np.zeros((self.M, self.G), dtype=np.float64)


def _find_table(name: str, prm_path: str) -> str:
    primary = os.path.join("..", "prm", name)
    if os.path.exists(primary):
        return primary
    alt = os.path.join(os.path.dirname(os.path.abspath(prm_path)), name)
    if os.path.exists(alt):
        return alt
    raise FileNotFoundError(f"There was a problem opening the input file {primary!r} (also tried {alt!r})")


def load(path: str) -> Parameters:
    kv = read_key_values(path)

    def geti(k, d):
        return _stoi(kv[k]) if k in kv else d

    def getd(k, d):
        return _stod(kv[k]) if k in kv else d

    def getb(k, d):
        return (kv[k] in _TRUE_STRINGS) if k in kv else d

    p = Parameters(M=geti("M", 2), G=geti("G", 1), source_file=path)
    p.efirst = getd("efirst", 0.1)
    p.elast = getd("elast", 10.0)
    p.X = getd("X", 1.0)
    p.N = geti("N", 100)
    p.bc_left_indicator = geti("bc_left_indicator", 2)
    p.bc_right_indicator = geti("bc_right_indicator", 1)
    p.use_mg_equilib = getb("use_mg_equilib", False)
    p.psi_source = np.zeros((p.M, p.G), dtype=np.float64)
    if not p.use_mg_equilib:
        counter = 0
        for tok in kv.get("psi_source", "no_sources_provided").split():
            try:
                val = float(tok)
            except ValueError:
                break
            p.psi_source[counter // p.G, counter % p.G] = val  # row-major fill, ParameterHandler.cpp:128-131
            counter += 1
    p.have_group_bounds = getb("have_group_bounds", False)
    if p.have_group_bounds:
        p.filename_group_bounds = kv.get("filename_group_bounds", "NA")
        p.group_bounds = read_doubles(_find_table(p.filename_group_bounds, path))
        if p.group_bounds.size != p.G + 1:
            raise AssertionError("Number of group bounds doesn't match the number of groups.")
    p.have_group_absorption_opacities = getb("have_group_absorption_opacities", False)
    if p.have_group_absorption_opacities:
        p.filename_group_kappa = kv.get("filename_group_kappa", "NA")
        p.group_kappa = read_doubles(_find_table(p.filename_group_kappa, path))
        if p.group_kappa.size != p.G:
            raise AssertionError("Number of group opacities doesn't match the number of groups.")
    p.rho = getd("rho", 1.0)
    p.kappa_grey = getd("kappa_grey", 1.0)
    p.T = getd("T", 1.0)
    p.V = getd("V", 0.0)
    p.use_correction = getb("use_correction", False)
    p.ts_method = geti("ts_method", 3)
    p.dt = getd("dt", 0.00001)
    p.max_timesteps = geti("max_timesteps", 1000)
    p.include_validation = getb("include_validation", True)
    return p


def dump(p: Parameters, path: str, extra_comment: str = "") -> None:
    """Write a `.prm` file the reference's parser reads back to the same values
    (17 significant digits; no trailing blanks after bools)."""
    lines = ["#####", "# generated parameter file " + extra_comment, "#####", ""]

    def d(v):
        return repr(float(v))

    lines += [f"M={p.M}", f"G={p.G}", f"efirst={d(p.efirst)}", f"elast={d(p.elast)}", f"X={d(p.X)}", f"N={p.N}"]
    lines += [f"bc_left_indicator={p.bc_left_indicator}", f"bc_right_indicator={p.bc_right_indicator}"]
    lines += ["use_mg_equilib=" + ("true" if p.use_mg_equilib else "false")]
    if not p.use_mg_equilib and p.psi_source is not None and np.any(p.psi_source != 0.0):
        lines += ["psi_source=" + " ".join(d(v) for v in np.asarray(p.psi_source).reshape(-1))]
    if p.have_group_bounds:
        lines += ["have_group_bounds=true", f"filename_group_bounds={p.filename_group_bounds}"]
    if p.have_group_absorption_opacities:
        lines += ["have_group_absorption_opacities=true", f"filename_group_kappa={p.filename_group_kappa}"]
    lines += [f"rho={d(p.rho)}", f"kappa_grey={d(p.kappa_grey)}", f"T={d(p.T)}", f"V={d(p.V)}"]
    lines += ["use_correction=" + ("true" if p.use_correction else "false")]
    lines += [f"ts_method={p.ts_method}", f"dt={d(p.dt)}", f"max_timesteps={p.max_timesteps}"]
    lines += ["include_validation=" + ("true" if p.include_validation else "false")]
    with open(path, "w") as fh:
        fh.write("\n".join(lines) + "\n")
