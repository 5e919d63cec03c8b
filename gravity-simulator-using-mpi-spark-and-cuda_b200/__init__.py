"""gravsim_b200 -- ctypes binding of libgravsim.so, the B200-native all-pairs gravity engine.

The directory name (``gravity-simulator-using-mpi-spark-and-cuda_b200``) is not a valid Python
identifier; load it with ``load_package()`` from ``__graft_entry__`` / ``tests/conftest.py`` (which
registers it as ``gravsim_b200``).

Everything numerical happens inside the CUDA library.  There is no Python/NumPy/CPU fallback:
if the shared library is missing or no B200 is visible, calls raise.
"""
from .api import (  # noqa: F401
    GravSim, GravSimError, GsConfig, GsInfo, GsStateHeader, lib, lib_path, build,
    ic_reference_uniform, ic_plummer, ic_two_galaxy, plan_slice, plan_cta_range,
    plan_tile_owners, plan_tiles, plan_sym_row, plan_sym_unit, plan_sym_cta, nccl_unique_id, state_write, state_read,
    dump_positions_cuda_format, dump_positions_mpi_format, EXPORTED_SYMBOLS,
    INTEGRATOR_EULER, INTEGRATOR_LEAPFROG, G_SI, FLAG_TWO_PASS, FLAG_NO_UNIFORM_MASS, FLAG_OVERLAP, FLAG_NCCL_EXCHANGE, FLAG_COMPENSATED, FLAG_DETERMINISTIC, FLAG_FP64_FORCES, potential_energy_noalloc, NOALLOC_MAX_BODIES,
)
