"""B200-native PI-ADMM consensus solver: host-side mirror of the reference's PI_ADMM_class
interface over the C ABI of csrc/libpiadmm.so (hand-written sm_100a CUDA)."""
from ._abi import (Params, Buffers, default_params, preset, inner_tight, load_library, PIADMMError,
                   DUAL_PLAIN, DUAL_PI, DUAL_PI_OLD)

__all__ = ["Params", "Buffers", "default_params", "preset", "inner_tight", "load_library",
           "PIADMMError", "DUAL_PLAIN", "DUAL_PI", "DUAL_PI_OLD"]
