"""radiative-transfer_b200 -- B200 (sm_100a) implementation of the radiative-transfer miniapp's hot path.

The package directory name contains a hyphen; import it with
``importlib.import_module("radiative-transfer_b200")``.

Contents (only what the hot path needs):
  csrc/          hand-written fp64 CUDA kernels + the C ABI (include/rt_cuda.h) -> librt_cuda.so
  capi.py        ctypes binding of the C ABI (no CPU fallback)
  solver.py      Python mirror of the reference's rt::Solver class surface
  prm.py         `.prm` parameter files (mirror of the reference's ParameterHandler)
  grid.py        Gauss-Legendre angles and the energy-group grid (host-side setup)
  sharding.py    energy-group sharding across ranks and result gathering
  host/          C++ drop-in rt::Solver (same header surface as the reference) on the C ABI
"""
from . import capi, grid, prm  # noqa: F401
from .prm import Parameters  # noqa: F401
from .solver import Solver, multigroup_opacities, planck_groups  # noqa: F401

__all__ = ["capi", "grid", "prm", "Parameters", "Solver", "planck_groups", "multigroup_opacities"]
