"""engine-sim-cli_b200 — B200-native batched engine-sim renderer.

Host-side mirror (Python, over the C ABI in include/engine_sim_b200.h) of the reference's
simulator boundary: `Simulator` follows ISimulator / ICombustionEngine
(engine-sim-bridge/include/simulator/ISimulator.h:12-45), `BatchSimulator` is the batch
extension the offline renderer uses.  All arithmetic runs in the CUDA kernels of
libenginesim_b200.so; there is no CPU path, and importing the API without the built library
raises."""
from .api import (BatchSimulator, MultiBatchSimulator, Controls, EngineSimError, Simulator, lib, library_path, ESIM_SUCCESS,  # noqa: F401
                  fir_run, probe_peak, check_fast_arith, check_fast_arith_noalloc, NOALLOC_MAX_N, state_fields, params_size, compile_script, script_library_root, TELEMETRY_STRIDE)
