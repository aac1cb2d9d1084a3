"""ti_decode_noalloc: the decode entry point for hosts that must keep the CUDA allocator
out of the library's hands -- a simulator that carves all device memory out of one
pool it reserved at start-up (a hidden cudaMalloc could fail or fragment that pool), or
a front end that decodes from a latency-sensitive loop (cudaMalloc / cudaFree
synchronise the device and would stall its other streams).  The caller passes every
input and output buffer; the look-back tables the kernel needs on top of that live in
a fixed __device__ array inside the library (sized for 256 MiB of text).

The test drives it through the C ABI, twice in a row on different inputs, and checks
each result bit for bit against the ordinary entry points on the same input (integer
records and counters: the tolerance of every other parity test here, i.e. none).  The
second input is smaller than the first, so stale table entries left by the first call
would show.  It must pass alone in a fresh process (the first call does the one-time
address look-up) and after the rest of the suite, hence the file name: it is collected
last.
"""
import ctypes as C

import numpy as np
import pytest

import harness
import tiformat as T
from accel_sim_framework_distributed_b200 import _ffi

pytestmark = pytest.mark.gpu


def _existing_path(tp, body, hdr):
    k = tp.open_buffer(body, hdr)
    d = k.run()
    out = dict(inst=d.inst.copy(), rows=d.rows.copy(), txn_off=d.txn_off.copy(), txn=d.txn.copy(), tb=d.tb.copy(),
               warp=d.warp.copy(), counters=np.array(d.counters, dtype=T.COUNTERS))
    tp.kernel_finalizer(k)
    return out


def _no_alloc_path(tp, torch, body, hdr, bufs, stream=None, cap_txn=None, expect_overflow=0, expect_flags=0):
    """bufs: caller-owned device tensors, allocated once and reused by both calls."""
    text = bufs["text"]
    text[:len(body)].copy_(torch.from_numpy(body))
    io = _ffi.NoAllocIO()
    io.d_text, io.len = text.data_ptr(), len(body)
    io.binary_version, io.trace_version = hdr.binary_version, hdr.trace_version
    io.shmem_base_addr, io.local_base_addr = hdr.shmem_base_addr, hdr.local_base_addr
    io.d_inst, io.cap_inst = bufs["inst"].data_ptr(), bufs["inst"].numel() // 32
    io.d_rows, io.cap_mem = bufs["rows"].data_ptr(), bufs["rows"].numel() // 256
    io.d_txn_off = bufs["txn_off"].data_ptr()
    io.d_txn, io.cap_txn = bufs["txn"].data_ptr(), (bufs["txn"].numel() // 32 if cap_txn is None else cap_txn)
    io.d_tb, io.cap_tb = bufs["tb"].data_ptr(), bufs["tb"].numel() // 32
    io.d_warp, io.cap_warp = bufs["warp"].data_ptr(), bufs["warp"].numel() // 16
    io.d_counters, io.d_status = bufs["counters"].data_ptr(), bufs["status"].data_ptr()
    assert bufs["txn_off"].numel() // 4 >= io.cap_mem + 1
    torch.cuda.synchronize()
    st = tp.lib.ti_decode_noalloc(tp.ctx, C.byref(io), stream.cuda_stream if stream is not None else None)
    assert st == 0, tp.lib.ti_last_error()
    # the context stream is a non-blocking stream of the library's: wait for the whole device
    if stream is not None:
        stream.synchronize()
    else:
        torch.cuda.synchronize()
    status = bufs["status"].cpu().numpy().view("<u4")
    assert status[1] == expect_flags and status[3] == expect_overflow, status   # parse flags, overflow bits
    cnt = bufs["counters"].cpu().numpy().view(T.COUNTERS)[0]
    n_inst, n_mem, n_txn, n_tb, n_warp = (int(cnt[f]) for f in ("n_inst", "n_mem", "n_txn", "n_tb", "n_warp"))

    def host(name, dtype, n):
        return bufs[name].cpu().numpy()[:n * np.dtype(dtype).itemsize].view(dtype).copy()

    return dict(inst=host("inst", T.INST, n_inst), rows=host("rows", "<u8", n_mem * 32).reshape(n_mem, 32),
                txn_off=host("txn_off", "<u4", n_mem + 1), txn=host("txn", T.TXN, n_txn), tb=host("tb", T.TB, n_tb),
                warp=host("warp", T.WARP, n_warp), counters=np.array(cnt, dtype=T.COUNTERS))


def test_no_alloc_matches_existing_entry_points_twice(cuda_parser):
    import torch
    tp = cuda_parser
    bufs = _bufs(torch)
    inputs = [
        dict(profile=5, seed=501, n_tb=600, insts_per_warp=12, binary_version=70, block_threads=128),  # ~6 MB, every line shape
        dict(profile=2, seed=502, n_tb=150, insts_per_warp=25),                                        # smaller: 1.5 MB
    ]
    for gen in inputs:
        img = harness.generate_bytes(**gen)
        hdr = tp.parse_header(img[:8192])
        body = np.ascontiguousarray(img[hdr.body_offset:])
        assert len(body) <= bufs["text"].numel()
        want = _existing_path(tp, body, hdr)
        got = _no_alloc_path(tp, torch, body, hdr, bufs)
        for name in ("inst", "rows", "txn_off", "txn", "tb", "warp"):
            assert np.array_equal(got[name], want[name]), name
        for f in T.COUNTERS.names:
            assert np.array_equal(got["counters"][f], want["counters"][f]), f


def test_no_alloc_rejects_oversized_text(cuda_parser):
    tp = cuda_parser
    io = _ffi.NoAllocIO()
    io.d_text, io.len = 16, _ffi.TI_NOALLOC_MAX_TEXT + 1          # never dereferenced: refused on size alone
    io.binary_version = 70
    for f in ("d_inst", "d_rows", "d_txn_off", "d_txn", "d_tb", "d_warp", "d_counters", "d_status"):
        setattr(io, f, 16)
    assert tp.lib.ti_decode_noalloc(tp.ctx, C.byref(io), None) == 7  # TI_ERR_TOO_LARGE
    assert b"exceed" in tp.lib.ti_last_error()


def _bufs(torch):
    u8 = dict(dtype=torch.uint8, device="cuda")
    return dict(text=torch.empty(8 << 20, **u8), inst=torch.empty(32 * 400000, **u8), rows=torch.empty(256 * 200000, **u8),
                txn_off=torch.empty(4 * 200001, **u8), txn=torch.empty(32 * 2000000, **u8), tb=torch.empty(32 * 20000, **u8),
                warp=torch.empty(16 * 100000, **u8), counters=torch.empty(T.COUNTERS.itemsize, **u8),
                status=torch.empty(32, **u8))


def test_no_alloc_against_the_oracle_on_a_caller_stream(cuda_parser, tmp_path):
    """The same entry point checked against the CPU oracle itself (not only against the library's other entry
    points), with the work queued on a stream the caller owns."""
    import os
    import torch
    tp = cuda_parser
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, profile=5, seed=503, n_tb=200, insts_per_warp=12, binary_version=75, block_threads=128)
    img = np.fromfile(path, dtype=np.uint8)
    hdr = tp.parse_header(img[:8192])
    body = np.ascontiguousarray(img[hdr.body_offset:])
    got = _no_alloc_path(tp, torch, body, hdr, _bufs(torch), stream=torch.cuda.Stream())
    o = T.load_dump(harness.oracle_dump(path, str(tmp_path)))
    for f in T.INST.names:
        if f != "opc_off":
            assert np.array_equal(got["inst"][f], o.inst[f]), f
    assert np.array_equal(got["rows"], o.rows) and np.array_equal(got["txn_off"], o.txn_off) and np.array_equal(got["txn"], o.txn)
    for f in ("x", "y", "z", "first_warp", "first_inst"):
        assert np.array_equal(got["tb"][f], o.tb[f]), f
    for f in ("warp_id", "n_insts", "first_inst"):
        assert np.array_equal(got["warp"][f], o.warp[f]), f
    assert not T.compare_counters(got["counters"], o.counters)


def test_no_alloc_undersized_capacity_and_broken_structure(cuda_parser):
    """Documented behaviour of the entry point that cannot retry: with too small a transaction capacity it still
    counts everything, raises the overflow bit in status[3] and writes nothing past the capacity; and a shard
    whose "#END_TB" is missing raises TI_EF_BAD_STRUCTURE in status[1] on the device (the caller never sees
    the host-side counters check of the other entry points)."""
    import torch
    tp = cuda_parser
    img = harness.generate_bytes(profile=2, seed=504, n_tb=100, insts_per_warp=25)
    hdr = tp.parse_header(img[:8192])
    body = np.ascontiguousarray(img[hdr.body_offset:])
    want = _existing_path(tp, body, hdr)
    bufs = _bufs(torch)
    cap = len(want["txn"]) // 2
    bufs["txn"].fill_(0xAB)
    got_cnt = None
    # (the helper asserts status[3] == 4: transaction capacity exceeded)
    text = bufs["text"]; text[:len(body)].copy_(torch.from_numpy(body))
    io = _ffi.NoAllocIO()
    io.d_text, io.len = text.data_ptr(), len(body)
    io.binary_version, io.trace_version = hdr.binary_version, hdr.trace_version
    io.shmem_base_addr, io.local_base_addr = hdr.shmem_base_addr, hdr.local_base_addr
    io.d_inst, io.cap_inst = bufs["inst"].data_ptr(), bufs["inst"].numel() // 32
    io.d_rows, io.cap_mem = bufs["rows"].data_ptr(), bufs["rows"].numel() // 256
    io.d_txn_off = bufs["txn_off"].data_ptr()
    io.d_txn, io.cap_txn = bufs["txn"].data_ptr(), cap
    io.d_tb, io.cap_tb = bufs["tb"].data_ptr(), bufs["tb"].numel() // 32
    io.d_warp, io.cap_warp = bufs["warp"].data_ptr(), bufs["warp"].numel() // 16
    io.d_counters, io.d_status = bufs["counters"].data_ptr(), bufs["status"].data_ptr()
    torch.cuda.synchronize()
    assert tp.lib.ti_decode_noalloc(tp.ctx, C.byref(io), None) == 0
    torch.cuda.synchronize()
    status = bufs["status"].cpu().numpy().view("<u4")
    assert status[3] & 4 and status[1] == 0, status
    got_cnt = bufs["counters"].cpu().numpy().view(T.COUNTERS)[0]
    assert int(got_cnt["n_txn"]) == len(want["txn"]) and int(got_cnt["n_inst"]) == len(want["inst"])   # counted in full
    txn_bytes = bufs["txn"].cpu().numpy()
    assert np.array_equal(txn_bytes[:cap * 32].view(T.TXN), want["txn"][:cap])        # what fits is right
    assert np.all(txn_bytes[cap * 32:(cap + 4096) * 32] == 0xAB)                      # nothing behind the capacity
    # a missing "#END_TB": structure error reported in the device-side status block
    cut = body.tobytes().rfind(b"#END_TB")
    broken = np.frombuffer(body.tobytes()[:cut] + b"\n", dtype=np.uint8).copy()
    text[:len(broken)].copy_(torch.from_numpy(broken))
    io.len, io.cap_txn = len(broken), bufs["txn"].numel() // 32
    torch.cuda.synchronize()
    assert tp.lib.ti_decode_noalloc(tp.ctx, C.byref(io), None) == 0
    torch.cuda.synchronize()
    status = bufs["status"].cpu().numpy().view("<u4")
    assert status[1] & _ffi.EF["BAD_STRUCTURE"], status
