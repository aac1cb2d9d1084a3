"""B200-native trace-ingest path of Accel-Sim (text .traceg -> warp-instruction
records -> coalesced memory transactions -> AccelWattch counter totals).

Only the CUDA library (csrc/libtraceingest.so, sm_100a) computes anything; this
package is the thin host binding around its C ABI (include/traceingest.h).
"""
from . import records  # noqa: F401
from ._ffi import EF, TraceIngestError, load_library  # noqa: F401
from .api import KernelTrace, trace_parser  # noqa: F401
