"""radio_observer_b200 -- B200-native (sm_100a) spectral hot path of radio-observer.

The product is the CUDA library (csrc/, C ABI in include/ro_cuda.h) and the host
side that mirrors the reference's Backend/recorder interface.  There is no CPU
fallback: importing works without a GPU (tables, host logic), compute calls need
libro_cuda.so and a B200.
"""
from .capi import (  # noqa: F401
    RoError, WfConfig, DetRec, Waterfall, load_library, library_path,
    window_table, twiddle_table, IN_INT16, IN_FLOAT32, IN_COMPLEX64, PREC_FP32, PREC_FP64,
    TUNE_NO_FUSED, TUNE_NO_PAIR,
)

from . import archive  # noqa: F401,E402  (time-sharded archive reprocessing, host-side trigger state machine)
from . import fits, recorders  # noqa: F401,E402  (archive path: FITS writer, snapshot / bolid file planning)

__all__ = [
    "archive", "fits", "recorders",
    "RoError", "WfConfig", "DetRec", "Waterfall", "load_library", "library_path",
    "window_table", "twiddle_table", "IN_INT16", "IN_FLOAT32", "IN_COMPLEX64", "PREC_FP32", "PREC_FP64",
    "TUNE_NO_FUSED", "TUNE_NO_PAIR",
]
