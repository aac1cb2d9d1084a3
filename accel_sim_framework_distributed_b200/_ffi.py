"""ctypes binding of the C ABI in include/traceingest.h.

The product library is csrc/libtraceingest.so, built for sm_100a by
`__graft_entry__.build()`.  There is no CPU implementation behind this binding: if the
shared library is missing, or no CUDA device is present, the calls fail loudly.
"""
import ctypes as C
import os

import numpy as np

from . import records as R

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.path.join(_HERE, "csrc", "libtraceingest.so")

STATUS_NAMES = {
    0: "TI_OK", 1: "TI_ERR_INVALID_ARG", 2: "TI_ERR_CUDA", 3: "TI_ERR_IO", 4: "TI_ERR_BINARY_VERSION",
    5: "TI_ERR_PARSE", 6: "TI_ERR_CAPACITY", 7: "TI_ERR_TOO_LARGE", 8: "TI_ERR_NO_DEVICE", 9: "TI_ERR_STATE",
}
TI_ERR_PARSE = 5

# TI_EF_* bits
EF = dict(LINE_TOO_LONG=1 << 0, BAD_NUMBER=1 << 1, BAD_REG=1 << 2, TOO_MANY_REGS=1 << 3, UNKNOWN_OPCODE=1 << 4,
          ZERO_WIDTH=1 << 5, BAD_STRUCTURE=1 << 6, FRAMING=1 << 7, TRUNCATED=1 << 8, BAD_WIDTH=1 << 9,
          ATOMIC_STRADDLE=1 << 10, DENSE_CHUNK=1 << 11, LONG_OPCODE=1 << 12, UNDEF_SPACE=1 << 13)


class KernelHdr(C.Structure):
    _fields_ = [
        ("kernel_name", C.c_char * 256), ("kernel_id", C.c_uint32), ("grid_dim", C.c_uint32 * 3),
        ("tb_dim", C.c_uint32 * 3), ("shmem", C.c_uint32), ("nregs", C.c_uint32), ("cuda_stream_id", C.c_uint64),
        ("binary_version", C.c_uint32), ("trace_version", C.c_uint32), ("nvbit_version", C.c_char * 64),
        ("shmem_base_addr", C.c_uint64), ("local_base_addr", C.c_uint64), ("body_offset", C.c_uint64),
    ]


class Counts(C.Structure):
    _fields_ = [
        ("n_inst", C.c_uint64), ("n_mem", C.c_uint64), ("n_txn", C.c_uint64), ("n_tb", C.c_uint64),
        ("n_warp", C.c_uint64), ("text_bytes", C.c_uint64), ("error_flags", C.c_uint32),
        ("first_error_line_off", C.c_uint32),
    ]


class View(C.Structure):
    _fields_ = [
        ("text", C.c_void_p), ("inst", C.c_void_p), ("addr_rows", C.c_void_p), ("txn_off", C.c_void_p),
        ("txn", C.c_void_p), ("tb", C.c_void_p), ("warp", C.c_void_p), ("counts", Counts),
    ]


class ViewCompact(C.Structure):
    """ti_view_compact: 16 B descriptors instead of base+stride rows / dense transactions."""
    _fields_ = [
        ("text", C.c_void_p), ("inst", C.c_void_p), ("desc", C.c_void_p), ("rows_full", C.c_void_p), ("txn_off", C.c_void_p),
        ("txn", C.c_void_p), ("tb", C.c_void_p), ("warp", C.c_void_p), ("counts", Counts), ("n_full", C.c_uint64),
        ("n_txn_stored", C.c_uint64),
    ]


class DevView(C.Structure):
    _fields_ = [
        ("text", C.c_void_p), ("inst", C.c_void_p), ("addr_rows", C.c_void_p), ("txn_off", C.c_void_p),
        ("txn", C.c_void_p), ("tb", C.c_void_p), ("warp", C.c_void_p), ("counters", C.c_void_p), ("counts", Counts),
    ]


class NoAllocIO(C.Structure):
    """ti_noalloc_io: caller-owned device buffers for ti_decode_noalloc."""
    _fields_ = [
        ("d_text", C.c_void_p), ("len", C.c_uint64), ("binary_version", C.c_uint32), ("trace_version", C.c_uint32),
        ("shmem_base_addr", C.c_uint64), ("local_base_addr", C.c_uint64),
        ("d_inst", C.c_void_p), ("cap_inst", C.c_uint64), ("d_rows", C.c_void_p), ("cap_mem", C.c_uint64),
        ("d_txn_off", C.c_void_p), ("d_txn", C.c_void_p), ("cap_txn", C.c_uint64),
        ("d_tb", C.c_void_p), ("cap_tb", C.c_uint64), ("d_warp", C.c_void_p), ("cap_warp", C.c_uint64),
        ("d_counters", C.c_void_p), ("d_status", C.c_void_p),
    ]


TI_NOALLOC_MAX_TEXT = 256 << 20


class LocalMap(C.Structure):
    _fields_ = [("num_shader", C.c_uint32), ("n_thread_per_shader", C.c_uint32), ("local_mem_map", C.c_uint32),
                ("padded_threads_per_cta", C.c_uint32), ("max_cta_per_shader", C.c_uint32), ("reserved", C.c_uint32),
                ("local_generic_start", C.c_uint64)]


TI_LOCAL_GENERIC_START = 0xC0000000 - 80 * (96 << 10) - 80 * 2048 * (1 << 14)


class DerivedStats(C.Structure):
    _fields_ = [("active_lane_distro", C.c_uint64 * 33), ("smem_bank_conflict", C.c_uint64), ("uncoalesced_gmem", C.c_uint64)]


class PipelineStats(C.Structure):
    _fields_ = [("regfile_reads", C.c_uint64), ("regfile_writes", C.c_uint64), ("non_rf_operands", C.c_uint64),
                ("decoded_insn", C.c_uint64), ("int_decoded_insn", C.c_uint64), ("fp_decoded_insn", C.c_uint64)]


class RegroupStats(C.Structure):
    _fields_ = [("n_lines", C.c_uint64), ("n_keys", C.c_uint64), ("out_bytes", C.c_uint64), ("n_tb_written", C.c_uint32),
                ("n_tb_empty", C.c_uint32), ("n_warp_empty", C.c_uint32), ("device_ms", C.c_float)]


class Command(C.Structure):
    _fields_ = [("kind", C.c_int32), ("line_off", C.c_uint32), ("line_len", C.c_uint32)]


# every symbol include/traceingest.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "ti_strerror", "ti_last_error", "ti_device_count", "ti_config_qv100", "ti_ctx_create", "ti_ctx_destroy",
    "ti_parse_kernel_header", "ti_kernel_open_file", "ti_kernel_open_buffer", "ti_kernel_open_device",
    "ti_kernel_close", "ti_kernel_upload", "ti_kernel_decode", "ti_kernel_sync", "ti_kernel_fetch",
    "ti_kernel_dev_view", "ti_kernel_counters", "ti_counters_add", "ti_counters_finalize", "ti_kernel_timing",
    "ti_plan_shards", "ti_parse_commandlist", "ti_parse_memcpy", "ti_decode_noalloc", "ti_kernel_decode_streamed",
    "ti_regroup_device", "ti_regroup_buffer", "ti_regroup_commandlist", "ti_regroup_buckets", "ti_kernel_coalesce_local", "ti_counters_derive", "ti_kernel_path_stats",
    "ti_kernel_decode_streamed_compact", "ti_expand_row", "ti_expand_txns",
    "ti_compress_rows", "ti_kernel_pipeline_stats", "ti_kernel_decode_window_compact", "ti_kernel_host_body", "ti_kernel_release_text", "ti_nccl_unique_id", "ti_nccl_comm_create", "ti_nccl_comm_destroy", "ti_allreduce_counters",
]


class TraceIngestError(RuntimeError):
    def __init__(self, status, detail):
        self.status = status
        super().__init__("%s: %s" % (STATUS_NAMES.get(status, status), detail))


def load_library(path=None):
    """dlopen the CUDA library.  No fallback: a missing .so is an error."""
    path = path or DEFAULT_LIB
    if not os.path.exists(path):
        raise FileNotFoundError(
            "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU implementation of this library." % path)
    lib = C.CDLL(path)
    V, P, S, U64 = C.c_void_p, C.POINTER, C.c_size_t, C.c_uint64
    sig = {
        "ti_strerror": (C.c_char_p, [C.c_int]),
        "ti_last_error": (C.c_char_p, []),
        "ti_device_count": (C.c_int, []),
        "ti_config_qv100": (None, [V]),
        "ti_ctx_create": (C.c_int, [V, P(V)]),
        "ti_ctx_destroy": (None, [V]),
        "ti_parse_kernel_header": (C.c_int, [V, S, P(KernelHdr)]),
        "ti_kernel_open_file": (C.c_int, [V, C.c_char_p, P(KernelHdr), P(V)]),
        "ti_kernel_open_buffer": (C.c_int, [V, V, S, P(KernelHdr), P(V)]),
        "ti_kernel_open_device": (C.c_int, [V, V, S, P(KernelHdr), P(V)]),
        "ti_kernel_close": (None, [V]),
        "ti_kernel_upload": (C.c_int, [V, S, S]),
        "ti_kernel_decode": (C.c_int, [V]),
        "ti_kernel_sync": (C.c_int, [V, P(Counts)]),
        "ti_kernel_fetch": (C.c_int, [V, P(View)]),
        "ti_kernel_dev_view": (C.c_int, [V, P(DevView)]),
        "ti_kernel_counters": (C.c_int, [V, V]),
        "ti_counters_add": (None, [V, V]),
        "ti_counters_finalize": (C.c_int, [V, V, V]),
        "ti_kernel_timing": (C.c_int, [V, P(C.c_float)]),
        "ti_plan_shards": (C.c_int, [V, S, C.c_int, P(S)]),
        "ti_parse_commandlist": (C.c_int, [V, S, P(Command), S, P(S)]),
        "ti_parse_memcpy": (C.c_int, [V, S, P(U64), P(U64)]),
        "ti_decode_noalloc": (C.c_int, [V, P(NoAllocIO), V]),
        "ti_kernel_decode_streamed": (C.c_int, [V, S, P(View)]),
        "ti_regroup_device": (C.c_int, [V, V, S, V, S, P(S), P(RegroupStats)]),
        "ti_regroup_buffer": (C.c_int, [V, V, S, P(V), P(S), P(RegroupStats)]),
        "ti_regroup_commandlist": (C.c_int, [C.c_char_p, S, C.c_char_p, S, P(S)]),
        "ti_regroup_buckets": (C.c_int, [V, P(V), P(V), P(U64), P(C.c_uint32)]),
        "ti_kernel_coalesce_local": (C.c_int, [V, P(LocalMap), V, S, V, S, V, P(S)]),
        "ti_counters_derive": (C.c_int, [V, P(DerivedStats)]),
        "ti_kernel_path_stats": (C.c_int, [V, P(C.c_uint32 * 4)]),
        "ti_kernel_decode_streamed_compact": (C.c_int, [V, S, P(ViewCompact)]),
        "ti_kernel_decode_window_compact": (C.c_int, [V, S, S, S, P(ViewCompact)]),
        "ti_kernel_host_body": (C.c_int, [V, P(V), P(S)]),
        "ti_kernel_release_text": (C.c_int, [V, S, S]),
        "ti_compress_rows": (C.c_int, [V, V, V, S, C.c_int, V, S, V, P(S)]),
        "ti_kernel_pipeline_stats": (C.c_int, [V, C.c_int, C.c_int, P(PipelineStats)]),
        "ti_expand_row": (None, [P(ViewCompact), U64, V]),
        "ti_expand_txns": (C.c_uint32, [P(ViewCompact), U64, V]),
        "ti_nccl_unique_id": (C.c_int, [V]),
        "ti_nccl_comm_create": (C.c_int, [V, V, C.c_int, C.c_int, P(V)]),
        "ti_nccl_comm_destroy": (None, [V]),
        "ti_allreduce_counters": (C.c_int, [V, V, V, V]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def np_view(ptr, dtype, n, shape=None):
    """Zero-copy numpy view of `n` records at a host pointer returned by the library."""
    dtype = np.dtype(dtype)
    if not n:
        return np.zeros((0,) + tuple(shape[1:]) if shape else (0,), dtype=dtype)
    buf = (C.c_uint8 * (n * dtype.itemsize)).from_address(ptr)
    a = np.frombuffer(buf, dtype=dtype, count=n)
    return a.reshape(shape) if shape else a


__all__ = ["load_library", "NoAllocIO", "PipelineStats", "RegroupStats", "DerivedStats", "LocalMap", "TI_LOCAL_GENERIC_START", "TI_NOALLOC_MAX_TEXT", "KernelHdr", "Counts", "View", "ViewCompact", "DevView", "Command", "TraceIngestError", "np_view",
           "SYMBOLS", "EF", "R"]
