"""Host-side mirror of the reference's trace_parser interface on top of the C ABI.

Names follow gpu-simulator/trace-parser/trace_parser.h:104-123 (`trace_parser`,
`parse_commandlist_file`, `parse_kernel_info`, `parse_memcpy_info`,
`get_next_threadblock_traces`, `kernel_finalizer`).  This Python layer exists for the
tests and the bench harness; the drop-in C++ shim with the exact reference signatures
is shim/trace_parser_b200.{h,cc}.  All compute happens in libtraceingest.so (CUDA).
"""
import ctypes as C
import os

import numpy as np

from . import _ffi
from . import records as R

COMMAND_TYPES = {1: "kernel_launch", 2: "cpu_gpu_mem_copy", 3: "gpu_cpu_mem_copy", 4: "ncclCommInitAll",
                 5: "ncclCommDestroy", 6: "ncclGroupStart", 7: "ncclGroupEnd", 8: "ncclAllReduce"}


class Decoded:
    """Host view of one decoded shard: numpy arrays over the library's pinned buffers
    (valid until the kernel object is closed or decoded again)."""

    def __init__(self, view, text_len):
        n = view.counts
        self.counts = {f: getattr(n, f) for f, _ in _ffi.Counts._fields_}
        self.inst = _ffi.np_view(view.inst, R.INST, n.n_inst)
        self.rows = _ffi.np_view(view.addr_rows, "<u8", n.n_mem * 32, shape=(n.n_mem, 32))
        self.txn_off = _ffi.np_view(view.txn_off, "<u4", n.n_mem + 1)
        self.txn = _ffi.np_view(view.txn, R.TXN, n.n_txn)
        self.tb = _ffi.np_view(view.tb, R.TB, n.n_tb)
        self.warp = _ffi.np_view(view.warp, R.WARP, n.n_warp)
        self.text = _ffi.np_view(view.text, "u1", text_len) if view.text else None
        self.counters = None
        self.power = None

    def opcode(self, i):
        r = self.inst[i]
        return bytes(self.text[int(r["opc_off"]):int(r["opc_off"]) + int(r["opcode_len"])])


class DecodedCompact:
    """Host view of ti_view_compact: descriptors instead of base+stride rows and dense transactions;
    expand_rows() / expand_txns() give what the full view holds (the library's own host expansion)."""

    def __init__(self, lib, view, text_len):
        n = view.counts
        self._lib, self._view = lib, view
        self.counts = {f: getattr(n, f) for f, _ in _ffi.Counts._fields_}
        self.n_full, self.n_txn_stored = view.n_full, view.n_txn_stored
        self.inst = _ffi.np_view(view.inst, R.INST, n.n_inst)
        self.desc = _ffi.np_view(view.desc, R.MEM_DESC, n.n_mem)
        self.rows_full = _ffi.np_view(view.rows_full, "<u8", view.n_full * 32, shape=(view.n_full, 32))
        self.txn_off = _ffi.np_view(view.txn_off, "<u4", n.n_mem + 1)
        self.txn = _ffi.np_view(view.txn, R.TXN, view.n_txn_stored)
        self.tb = _ffi.np_view(view.tb, R.TB, n.n_tb)
        self.warp = _ffi.np_view(view.warp, R.WARP, n.n_warp)
        self.text = _ffi.np_view(view.text, "u1", text_len) if view.text else None
        self.counters = None
        self.d2h_bytes = (32 * len(self.inst) + 16 * len(self.desc) + 256 * len(self.rows_full) + 4 * len(self.txn_off) +
                          32 * len(self.txn) + 32 * len(self.tb) + 16 * len(self.warp))

    def expand_rows(self, lo=0, hi=None):
        """ti_expand_row for memory instructions [lo, hi)."""
        hi = len(self.desc) if hi is None else hi
        out = np.zeros((hi - lo, 32), dtype="<u8")
        for r in range(lo, hi):
            self._lib.ti_expand_row(C.byref(self._view), r, out[r - lo].ctypes.data)
        return out

    def expand_txns(self, lo=0, hi=None):
        """ti_expand_txns for memory instructions [lo, hi): (records, CSR offsets)."""
        hi = len(self.desc) if hi is None else hi
        buf = np.zeros(64, dtype=R.TXN)
        parts, off = [], [0]
        for r in range(lo, hi):
            n = self._lib.ti_expand_txns(C.byref(self._view), r, buf.ctypes.data)
            parts.append(buf[:n].copy())
            off.append(off[-1] + n)
        return (np.concatenate(parts) if parts else np.zeros(0, dtype=R.TXN)), np.asarray(off, dtype=np.uint32)


class KernelTrace:
    """kernel_trace_t (trace_parser.h:81-102) plus the device-side decode state."""

    def __init__(self, parser, handle, hdr, body_len, keepalive=None):
        self._p, self._h, self.hdr, self.body_len = parser, handle, hdr, body_len
        self._keep = keepalive
        self._next_tb = 0
        self._decoded = None

    # -- kernel_trace_t fields
    kernel_name = property(lambda s: s.hdr.kernel_name.decode())
    kernel_id = property(lambda s: s.hdr.kernel_id)
    grid_dim = property(lambda s: tuple(s.hdr.grid_dim))
    tb_dim = property(lambda s: tuple(s.hdr.tb_dim))
    binary_version = property(lambda s: s.hdr.binary_version)
    trace_version = property(lambda s: s.hdr.trace_version)
    shmem_base_addr = property(lambda s: s.hdr.shmem_base_addr)
    local_base_addr = property(lambda s: s.hdr.local_base_addr)

    def upload(self, begin=0, end=None):
        self._p._ck(self._p.lib.ti_kernel_upload(self._h, begin, self.body_len if end is None else end))
        self._range = (begin, self.body_len if end is None else end)

    def decode(self):
        self._p._ck(self._p.lib.ti_kernel_decode(self._h))

    def sync(self, check=True):
        counts = _ffi.Counts()
        st = self._p.lib.ti_kernel_sync(self._h, C.byref(counts))
        if st != 0 and (check or st != _ffi.TI_ERR_PARSE):
            self._p._ck(st)
        return counts

    def fetch(self):
        view = _ffi.View()
        self._p._ck(self._p.lib.ti_kernel_fetch(self._h, C.byref(view)))
        b, e = getattr(self, "_range", (0, self.body_len))
        d = Decoded(view, e - b)
        cnt = np.zeros(1, dtype=R.COUNTERS)
        self._p._ck(self._p.lib.ti_kernel_counters(self._h, cnt.ctypes.data))
        d.counters = cnt[0]
        pw = np.zeros(1, dtype=R.POWER)
        self._p._ck(self._p.lib.ti_counters_finalize(self._p.ctx, cnt.ctypes.data, pw.ctypes.data))
        d.power = pw[0]
        return d

    def counters(self):
        cnt = np.zeros(1, dtype=R.COUNTERS)
        self._p._ck(self._p.lib.ti_kernel_counters(self._h, cnt.ctypes.data))
        return cnt[0]

    def dev_view(self):
        v = _ffi.DevView()
        self._p._ck(self._p.lib.ti_kernel_dev_view(self._h, C.byref(v)))
        return v

    def path_stats(self):
        """(staging overflows, early resolves, parking clashes, descriptor-list overflows) of the last decode."""
        out = (C.c_uint32 * 4)()
        self._p._ck(self._p.lib.ti_kernel_path_stats(self._h, C.byref(out)))
        return tuple(out)

    def pipeline_stats(self, clock_gated_reg_file=0, n_regfile_gating_group=4):
        """ti_kernel_pipeline_stats of the last one-piece decode: register-file and decode counters."""
        out = _ffi.PipelineStats()
        self._p._ck(self._p.lib.ti_kernel_pipeline_stats(self._h, clock_gated_reg_file, n_regfile_gating_group, C.byref(out)))
        return {f: getattr(out, f) for f, _ in _ffi.PipelineStats._fields_}

    def decode_ms(self):
        ms = C.c_float()
        self._p._ck(self._p.lib.ti_kernel_timing(self._h, C.byref(ms)))
        return ms.value

    def run(self, begin=0, end=None, check=True):
        """upload + decode + sync + fetch: the whole path for one shard."""
        self.upload(begin, end)
        self.decode()
        counts = self.sync(check=check)
        if counts.error_flags:
            return counts
        self._decoded = self.fetch()
        self._next_tb = 0
        return self._decoded

    def run_streamed(self, shard_bytes=0, check=True):
        """The whole host -> host path with copies and kernel overlapped (ti_kernel_decode_streamed)."""
        view = _ffi.View()
        st = self._p.lib.ti_kernel_decode_streamed(self._h, shard_bytes, C.byref(view))
        if st != 0 and (check or st != _ffi.TI_ERR_PARSE):
            self._p._ck(st)
        if st != 0:
            return None
        self._range = (0, self.body_len)
        d = Decoded(view, self.body_len)
        cnt = np.zeros(1, dtype=R.COUNTERS)
        self._p._ck(self._p.lib.ti_kernel_counters(self._h, cnt.ctypes.data))
        d.counters = cnt[0]
        pw = np.zeros(1, dtype=R.POWER)
        self._p._ck(self._p.lib.ti_counters_finalize(self._p.ctx, cnt.ctypes.data, pw.ctypes.data))
        d.power = pw[0]
        self._decoded = d
        self._next_tb = 0
        return d

    def run_streamed_compact(self, shard_bytes=0):
        """ti_kernel_decode_streamed_compact: the pipelined host -> host path with compact results."""
        view = _ffi.ViewCompact()
        self._p._ck(self._p.lib.ti_kernel_decode_streamed_compact(self._h, shard_bytes, C.byref(view)))
        d = DecodedCompact(self._p.lib, view, self.body_len)
        cnt = np.zeros(1, dtype=R.COUNTERS)
        self._p._ck(self._p.lib.ti_kernel_counters(self._h, cnt.ctypes.data))
        d.counters = cnt[0]
        return d

    def coalesce_local(self, local_map, issues):
        """issues: numpy array of dtype records.LOCAL_ISSUE (inst, sid, tid0).  Returns (txn, txn_off)."""
        issues = np.ascontiguousarray(issues, dtype=R.LOCAL_ISSUE)
        n = len(issues)
        off = np.zeros(n + 1, dtype=np.uint32)
        n_txn = C.c_size_t()
        lib = self._p.lib
        st = lib.ti_kernel_coalesce_local(self._h, C.byref(local_map), issues.ctypes.data, n, None, 0, off.ctypes.data, C.byref(n_txn))
        if st not in (0, 6):
            self._p._ck(st)
        txn = np.zeros(max(1, n_txn.value), dtype=R.TXN)
        self._p._ck(lib.ti_kernel_coalesce_local(self._h, C.byref(local_map), issues.ctypes.data, n, txn.ctypes.data, len(txn), off.ctypes.data, C.byref(n_txn)))
        return txn[:n_txn.value], off

    def close(self):
        if self._h:
            self._p.lib.ti_kernel_close(self._h)
            self._h = None


class trace_parser:  # noqa: N801 -- the reference's class name (trace_parser.h:104)
    def __init__(self, kernellist_filepath=None, config=None, lib_path=None, device=0):
        self.lib = _ffi.load_library(lib_path)
        self.kernellist_filename = kernellist_filepath
        cfg = R.qv100_config() if config is None else config
        self.config = np.array([cfg], dtype=R.CONFIG)
        self.config["device"] = device
        ctx = C.c_void_p()
        self._ck(self.lib.ti_ctx_create(self.config.ctypes.data, C.byref(ctx)))
        self.ctx = ctx

    def _ck(self, st):
        if st != 0:
            raise _ffi.TraceIngestError(st, self.lib.ti_last_error().decode())

    def close(self):
        if self.ctx:
            self.lib.ti_ctx_destroy(self.ctx)
            self.ctx = None

    # trace_parser.cc:220-284
    def parse_commandlist_file(self):
        try:
            text = open(self.kernellist_filename, "rb").read()
        except OSError:
            raise _ffi.TraceIngestError(3, "Unable to open file: %s" % self.kernellist_filename)
        n = C.c_size_t()
        self._ck(self.lib.ti_parse_commandlist(text, len(text), None, 0, C.byref(n)))
        cmds = (_ffi.Command * max(1, n.value))()
        self._ck(self.lib.ti_parse_commandlist(text, len(text), cmds, n.value, C.byref(n)))
        directory = os.path.dirname(self.kernellist_filename)
        out = []
        for c in cmds[:n.value]:
            line = text[c.line_off:c.line_off + c.line_len].decode()
            if c.kind == 1:
                line = (directory + "/" + line) if directory else ("/" + line if self.kernellist_filename.startswith("/") else line)
            out.append((line, c.kind))
        return out

    # trace_parser.cc:286-297
    def parse_memcpy_info(self, memcpy_command):
        a, n = C.c_uint64(), C.c_uint64()
        b = memcpy_command.encode()
        self._ck(self.lib.ti_parse_memcpy(b, len(b), C.byref(a), C.byref(n)))
        return a.value, n.value

    # trace_parser.cc:299-377
    def parse_kernel_info(self, kerneltraces_filepath):
        hdr = _ffi.KernelHdr()
        h = C.c_void_p()
        self._ck(self.lib.ti_kernel_open_file(self.ctx, kerneltraces_filepath.encode(), C.byref(hdr), C.byref(h)))
        return KernelTrace(self, h, hdr, os.path.getsize(kerneltraces_filepath) - hdr.body_offset)

    def parse_header(self, image):
        hdr = _ffi.KernelHdr()
        buf = np.frombuffer(image, dtype=np.uint8)
        self._ck(self.lib.ti_parse_kernel_header(buf.ctypes.data, len(buf), C.byref(hdr)))
        return hdr

    def open_buffer(self, body, hdr):
        """Kernel object over a caller-held host buffer (numpy uint8 array) with the body."""
        body = np.ascontiguousarray(body, dtype=np.uint8)
        h = C.c_void_p()
        self._ck(self.lib.ti_kernel_open_buffer(self.ctx, body.ctypes.data, len(body), C.byref(hdr), C.byref(h)))
        return KernelTrace(self, h, hdr, len(body), keepalive=body)

    def open_device(self, dev_ptr, nbytes, hdr, keepalive=None):
        h = C.c_void_p()
        self._ck(self.lib.ti_kernel_open_device(self.ctx, C.c_void_p(dev_ptr), nbytes, C.byref(hdr), C.byref(h)))
        return KernelTrace(self, h, hdr, nbytes, keepalive=keepalive)

    # trace_parser.cc:387-447: hands out thread blocks in file order
    def get_next_threadblock_traces(self, kernel):
        """Returns (tb_rec, [records of warp 0, records of warp 1, ...]) for the next
        thread block, or None after the last.  The kernel must have been decoded
        (KernelTrace.run)."""
        d = kernel._decoded
        if d is None:
            d = kernel.run()
        k = kernel._next_tb
        if k >= len(d.tb):
            return None
        kernel._next_tb += 1
        tb = d.tb[k]
        w_lo = int(tb["first_warp"])
        w_hi = int(d.tb[k + 1]["first_warp"]) if k + 1 < len(d.tb) else len(d.warp)
        warps = {}
        for w in d.warp[w_lo:w_hi]:
            lo = int(w["first_inst"])
            warps[int(w["warp_id"])] = d.inst[lo:lo + int(w["n_insts"])]
        return tb, warps

    # trace_parser.cc:379-385
    def kernel_finalizer(self, kernel):
        kernel.close()

    # post-traces-processing.cpp:85-194 (writer side)
    def regroup(self, trace):
        """kernel-N.trace image (bytes / numpy uint8) -> (kernel-N.traceg image as bytes, RegroupStats)."""
        buf = np.frombuffer(trace, dtype=np.uint8) if not isinstance(trace, np.ndarray) else np.ascontiguousarray(trace, dtype=np.uint8)
        out, n, st = C.c_void_p(), C.c_size_t(), _ffi.RegroupStats()
        self._ck(self.lib.ti_regroup_buffer(self.ctx, buf.ctypes.data if len(buf) else None, len(buf), C.byref(out), C.byref(n), C.byref(st)))
        return C.string_at(out.value, n.value), st

    def regroup_device(self, dev_ptr, nbytes, out_ptr=None, cap=0):
        """Device-resident variant: returns (out_len, RegroupStats); the text goes to out_ptr when given."""
        n, st = C.c_size_t(), _ffi.RegroupStats()
        self._ck(self.lib.ti_regroup_device(self.ctx, C.c_void_p(dev_ptr), nbytes, C.c_void_p(out_ptr) if out_ptr else None, cap, C.byref(n), C.byref(st)))
        return n.value, st

    # post-traces-processing.cpp:44-83
    def regroup_commandlist(self, text):
        n = C.c_size_t()
        st = self.lib.ti_regroup_commandlist(text, len(text), None, 0, C.byref(n))
        if st not in (0, 6):
            self._ck(st)
        out = C.create_string_buffer(max(1, n.value))
        self._ck(self.lib.ti_regroup_commandlist(text, len(text), out, n.value, C.byref(n)))
        return out.raw[:n.value]

    def derive(self, counters):
        """ti_counters_derive: the trace-derivable statistics of SURVEY 8 f3 from a counter block."""
        cnt = np.zeros(1, dtype=R.COUNTERS)
        cnt[0] = counters
        out = _ffi.DerivedStats()
        self._ck(self.lib.ti_counters_derive(cnt.ctypes.data, C.byref(out)))
        return out

    def plan_shards(self, body, n):
        body = np.ascontiguousarray(body, dtype=np.uint8)
        cuts = (C.c_size_t * (n + 1))()
        self._ck(self.lib.ti_plan_shards(body.ctypes.data, len(body), n, cuts))
        return list(cuts)
