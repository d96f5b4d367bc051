"""mpmc_b200 -- B200-native (sm_100a, fp64 CUDA) implementation of MPMC's per-step
energy hot path behind the reference's energy()/polar() interface.

Only what the path needs lives here:
  csrc/   hand-written CUDA kernels + the C ABI (include/mpmc_b200.h)
  shim/   the drop-in energy(system_t*) / polar(system_t*) symbols over that ABI
  api.py  host-side mirror of the reference interface (ctypes over the C ABI)
  systems.py  synthetic systems of BASELINE.json's configs
"""
from .api import (Counters, Energies, Engine, EngineError, Params, System, default_params, load_library,  # noqa: F401
                  ABI_SYMBOLS, LIB_PATH)
