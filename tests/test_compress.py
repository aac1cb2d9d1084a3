"""ti_compress_rows (second half): the tracer's address compressors on the device.

  * against the REFERENCE's own base_stride_compress / base_delta_compress (oracle/_ref/compress_ref: the two
    functions cut out of tracer_tool.cu and compiled as they are), text for text, on rows that hit every branch;
  * round trip: rows -> text (device) -> .traceg lines -> fused decode kernel -> rows, bit-identical -- no
    hand-written generator stands between the two directions.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import edge_cases
import harness

COMPRESS_REF = os.path.join(harness.ROOT, "oracle", "_ref", "compress_ref")


def _compress(tp, rows, masks, enable, to_device=None):
    """Runs ti_compress_rows; returns (text bytes, offsets).  to_device: numpy -> device pointer holder (GPU) or None (emulated)."""
    n = len(masks)
    rows = np.ascontiguousarray(rows, dtype="<u8")
    masks = np.ascontiguousarray(masks, dtype="<u4")
    lib = tp.lib
    nbytes = C.c_size_t()
    if to_device is None:
        off = np.zeros(n + 1, dtype="<u8")
        tp._ck(lib.ti_compress_rows(tp.ctx, rows.ctypes.data, masks.ctypes.data, n, enable, None, 0, off.ctypes.data, C.byref(nbytes)))
        text = np.zeros(max(1, nbytes.value), dtype=np.uint8)
        tp._ck(lib.ti_compress_rows(tp.ctx, rows.ctypes.data, masks.ctypes.data, n, enable, text.ctypes.data, len(text), off.ctypes.data, C.byref(nbytes)))
        return text[:nbytes.value].tobytes(), off
    import torch
    d_rows, d_masks = to_device(rows.view(np.uint8).reshape(-1)), to_device(masks.view(np.uint8).reshape(-1))
    d_off = torch.zeros(8 * (n + 1), dtype=torch.uint8, device="cuda")
    tp._ck(lib.ti_compress_rows(tp.ctx, d_rows.data_ptr(), d_masks.data_ptr(), n, enable, None, 0, d_off.data_ptr(), C.byref(nbytes)))
    d_text = torch.zeros(max(16, nbytes.value), dtype=torch.uint8, device="cuda")
    tp._ck(lib.ti_compress_rows(tp.ctx, d_rows.data_ptr(), d_masks.data_ptr(), n, enable, d_text.data_ptr(), d_text.numel(), d_off.data_ptr(), C.byref(nbytes)))
    return d_text.cpu().numpy()[:nbytes.value].tobytes(), d_off.cpu().numpy().view("<u8")


def _branchy_rows(n, seed):
    """Rows that reach every branch of the two compressors: empty masks, single lanes, a first lane at 31, clean runs,
    runs with one broken difference, lanes behind a gap, negative and int-overflowing strides, wrap-around."""
    rng = np.random.default_rng(seed)
    rows = np.zeros((n, 32), dtype=np.uint64)
    masks = np.zeros(n, dtype=np.uint32)
    for i in range(n):
        kind = i % 10
        base = np.uint64(rng.integers(1 << 20, 1 << 46))
        stride = int(rng.choice([0, 1, 4, 8, 128, -4, -256, 1 << 31, -(1 << 31) - 8, 3]))
        a = (int(base) + stride * np.arange(32, dtype=object)) % (1 << 64)
        a = np.array([int(x) for x in a], dtype=np.uint64)
        if kind == 0:
            m = 0
        elif kind == 1:
            m = 1 << int(rng.integers(0, 32))
        elif kind == 2:
            m = 1 << 31
        elif kind == 3:
            m = 0xFFFFFFFF
        elif kind == 4:
            lo = int(rng.integers(0, 30)); ln = int(rng.integers(2, 33 - lo)); m = ((1 << ln) - 1) << lo
        elif kind == 5:   # a run, a gap, more lanes
            m = 0x0000FF0F
        elif kind == 6:   # one broken difference inside the run
            m = 0xFFFFFFFF; a[int(rng.integers(2, 32))] += np.uint64(4)
        elif kind == 7:   # scattered
            m = int(rng.integers(1, 1 << 32)); a = rng.integers(0, 1 << 48, 32).astype(np.uint64)
        elif kind == 8:   # near 2^64
            m = 0x000000FF; a = (np.uint64(0xFFFFFFFFFFFFFFF0) + np.arange(32, dtype=np.uint64) * np.uint64(8))
        else:
            m = int(rng.integers(1, 1 << 32))
        rows[i] = a
        masks[i] = m
    return rows, masks


def _ref_text(rows, masks, enable, tmp):
    p = os.path.join(tmp, "rows.bin")
    rec = np.zeros(len(masks), dtype=np.dtype([("mask", "<u4"), ("pad", "<u4"), ("addrs", "<u8", (32,))]))
    rec["mask"], rec["addrs"] = masks, rows
    with open(p, "wb") as f:
        f.write(np.uint64(len(masks)).tobytes())
        f.write(rec.tobytes())
    r = subprocess.run([COMPRESS_REF, p, str(enable)], capture_output=True, check=True)
    return r.stdout.split(b"\n")[:-1]


def _check_against_reference(tp, tmp, to_device=None):
    rows, masks = _branchy_rows(400, 5)
    for enable in (1, 0):
        text, off = _compress(tp, rows, masks, enable, to_device)
        want = _ref_text(rows, masks, enable, tmp)
        got = [text[int(off[i]):int(off[i + 1])] for i in range(len(masks))]
        assert got == want, next((i, g, w) for i, (g, w) in enumerate(zip(got, want)) if g != w)


def _round_trip(tp, tmp, to_device=None, **gen):
    """rows decoded from a corpus -> compressed text (device) -> new .traceg -> decoded again: same rows."""
    path = os.path.join(tmp, "kernel-1.traceg")
    harness.generate(path, **gen)
    k = tp.parse_kernel_info(path)
    d = k.run()
    has = d.inst["mem_idx"] != 0xFFFFFFFF
    rows, masks = d.rows.copy(), d.inst["mask"][has].copy()
    assert len(rows) == len(masks) > 0
    inst = d.inst.copy()
    opc = [d.opcode(i) for i in range(len(inst))]
    tp.kernel_finalizer(k)
    # the decode zeroes what lies behind the first run of a base+stride line, so compress what the decode gave back
    # with the same choice the tracer would make for those rows
    text, off = _compress(tp, rows, masks, 1, to_device)
    lines, mi = [], 0
    for i, r in enumerate(inst):
        head = "%04x %08x %d " % (r["pc"], r["mask"], (r["pack_b"] >> 3) & 1)
        if (r["pack_b"] >> 3) & 1:
            head += "R%d " % r["dst"]
        head += opc[i].decode() + " %d " % (r["pack_b"] & 7) + "".join("R%d " % r["src"][j] for j in range(r["pack_b"] & 7))
        if r["mem_idx"] != 0xFFFFFFFF:
            head += "%d " % max(1, int(r["data_size"])) + text[int(off[mi]):int(off[mi + 1])].decode()
            mi += 1
        else:
            head += "0 "
        lines.append(head + "0")
    img = edge_cases.image([(0, 0, 0, [lines, []])])
    p2 = os.path.join(tmp, "kernel-2.traceg")
    open(p2, "wb").write(img)
    k2 = tp.parse_kernel_info(p2)
    d2 = k2.run()
    assert np.array_equal(d2.rows, rows)
    assert np.array_equal(d2.inst["mask"], inst["mask"]) and np.array_equal(d2.inst["mem_idx"], inst["mem_idx"])
    tp.kernel_finalizer(k2)


needs_ref = pytest.mark.skipif(not os.access(COMPRESS_REF, os.X_OK), reason="reference-built compressor oracle not present")


@needs_ref
def test_emu_compress_matches_reference(emu_parser, tmp_path):
    _check_against_reference(emu_parser, str(tmp_path))


def test_emu_compress_round_trip(emu_parser, tmp_path):
    _round_trip(emu_parser, str(tmp_path), profile=2, seed=8, n_tb=3, insts_per_warp=25)
    _round_trip(emu_parser, str(tmp_path), profile=5, seed=9, n_tb=6, insts_per_warp=10, block_threads=64)


def _to_device(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@needs_ref
def test_compress_matches_reference(cuda_parser, tmp_path):
    _check_against_reference(cuda_parser, str(tmp_path), _to_device)


@pytest.mark.gpu
@pytest.mark.parametrize("profile,kw", [(2, dict(n_tb=200, insts_per_warp=25)), (3, dict(n_tb=60, insts_per_warp=12)),
                                        (5, dict(n_tb=100, insts_per_warp=10, block_threads=128))])
def test_compress_round_trip(cuda_parser, tmp_path, profile, kw):
    """BASELINE config 2 / config 3 shapes (and the mixed corpus): rows -> text -> ti_decode_kernel -> rows."""
    _round_trip(cuda_parser, str(tmp_path), _to_device, profile=profile, seed=20 + profile, **kw)
