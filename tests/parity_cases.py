"""Parity checks shared by the CPU-emulated (test_emu_kernel.py) and the GPU
(test_gpu_parity.py) suites.  Bar: bit-exact for records, address rows, transaction
lists and integer counters; 1e-9 relative for the AccelWattch floating-point sums
(BASELINE.json north_star)."""
import glob
import os

import numpy as np

import edge_cases
import harness
import tiformat as T
from accel_sim_framework_distributed_b200 import _ffi, trace_parser

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
POWER_RTOL = 1e-9


def parser_with(tp, over):
    """A parser with the same library but another ti_config (or tp itself)."""
    if not over:
        return tp, False
    lib_path = tp.lib._name
    return trace_parser(lib_path=lib_path, config=T.qv100_config(**over)), True


def check_against_dump(d, o, offsets, skip_counters=("reserved", "n_begin_tb", "n_end_tb", "text_bytes", "n_warp", "n_mem"), tp=None):
    msgs = T.compare(d, o, offsets=offsets, text_a=d.text, names=("cuda", "oracle"))
    msgs += T.compare_counters(d.counters, o.counters, skip=skip_counters)
    msgs += T.compare_power(d.power, o.power, rtol=POWER_RTOL)
    assert not msgs, "\n".join(msgs[:12])
    # when the dump comes from the reference-built oracle it carries what the reference's OWN
    # ptx_file_line_stats_add_smem_bank_conflict / _uncoalesced_gmem calls added up to
    if tp is not None and int(o.counters["reserved"][3]) == 1:
        ds = tp.derive(d.counters)
        assert ds.smem_bank_conflict == int(o.counters["reserved"][1])
        assert ds.uncoalesced_gmem == int(o.counters["reserved"][2])
        assert list(ds.active_lane_distro) == [int(v) for v in np.asarray(o.counters["hist"]).sum(axis=0)]


def golden_names():
    return sorted(os.path.basename(p)[:-7] for p in glob.glob(os.path.join(GOLDEN, "*.traceg")) if os.path.exists(p[:-7] + ".tidump"))


def run_golden(tp, name):
    """CUDA path vs the reference-produced fixture committed under tests/golden."""
    over = edge_cases.VALID.get(name, (None, {}))[1]
    p, own = parser_with(tp, over)
    try:
        k = p.parse_kernel_info(os.path.join(GOLDEN, name + ".traceg"))
        d = k.run()
        o = T.load_dump(os.path.join(GOLDEN, name + ".tidump"))
        check_against_dump(d, o, offsets=False, tp=p)
        p.kernel_finalizer(k)
    finally:
        if own:
            p.close()


def run_generated(tp, tmp_path, over=None, **gen):
    """CUDA path vs the oracle run here on a freshly generated corpus."""
    p, own = parser_with(tp, over or {})
    try:
        path = os.path.join(str(tmp_path), "kernel-1.traceg")
        harness.generate(path, **gen)
        k = p.parse_kernel_info(path)
        d = k.run()
        cfg = T.qv100_config(**over) if over else None
        o = T.load_dump(harness.oracle_dump(path, str(tmp_path), config=cfg))
        check_against_dump(d, o, offsets=harness.oracle_kind() == "port", tp=p)
        p.kernel_finalizer(k)
        return d
    finally:
        if own:
            p.close()


def run_error(tp, tmp_path, name):
    img, flag = edge_cases.ERRORS[name]
    path = os.path.join(str(tmp_path), name + ".traceg")
    open(path, "wb").write(img)
    k = tp.parse_kernel_info(path)
    k.upload()
    k.decode()
    counts = _ffi.Counts()
    st = tp.lib.ti_kernel_sync(k._h, counts)
    tp.kernel_finalizer(k)
    assert st == _ffi.TI_ERR_PARSE, (name, st)
    assert counts.error_flags & _ffi.EF[flag], (name, hex(counts.error_flags))
    # the reported offset lies inside the offending line (the second instruction line of these images)
    body = img[k.hdr.body_offset:]
    first = b"0000 ffffffff 1 R1 FADD 2 R2 R3 0 0\n"
    if flag not in ("BAD_STRUCTURE", "FRAMING", "UNDEF_SPACE") and first in body:
        lo = body.index(first) + len(first)
        hi = body.index(b"\n", lo)
        assert lo <= counts.first_error_line_off < hi, (name, counts.first_error_line_off, lo, hi)
    # the restatement raises the same flag
    assert harness.oracle_flags(path, str(tmp_path)) & _ffi.EF[flag], name


def run_sharded(tp, tmp_path, n_shards, **gen):
    """Decoding byte-range shards cut at #BEGIN_TB and concatenating the results equals
    decoding the file in one piece (the path shards with no exchange)."""
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, **gen)
    k = tp.parse_kernel_info(path)
    whole = k.run()
    want = dict(inst=whole.inst.copy(), rows=whole.rows.copy(), txn=whole.txn.copy(), tb=whole.tb.copy(),
                warp=whole.warp.copy(), counters=whole.counters.copy())
    body = np.fromfile(path, dtype=np.uint8)[k.hdr.body_offset:]
    cuts = tp.plan_shards(body, n_shards)
    assert cuts[0] == 0 and cuts[-1] == len(body) and cuts == sorted(cuts)
    acc = np.zeros(1, dtype=T.COUNTERS)
    parts = []
    for b, e in zip(cuts[:-1], cuts[1:]):
        assert b == 0 or b == e or bytes(body[b:b + 9]) == b"#BEGIN_TB"
        d = k.run(b, e)
        cnt = np.zeros(1, dtype=T.COUNTERS)
        cnt[0] = d.counters
        tp.lib.ti_counters_add(acc.ctypes.data, cnt.ctypes.data)
        inst = d.inst.copy()
        inst["opc_off"] += np.uint32(b)                    # shard-relative -> body-relative
        parts.append((inst, d.rows.copy(), d.txn.copy(), d.tb.copy(), d.warp.copy()))
    inst = np.concatenate([p[0] for p in parts])
    mem_base = np.cumsum([0] + [len(p[1]) for p in parts])
    inst_base = np.cumsum([0] + [len(p[0]) for p in parts])
    warp_base = np.cumsum([0] + [len(p[4]) for p in parts])
    lo = 0
    for i, p in enumerate(parts):
        seg = inst[lo:lo + len(p[0])]
        has = seg["mem_idx"] != 0xFFFFFFFF
        seg["mem_idx"][has] += np.uint32(mem_base[i])
        p[3]["first_inst"] += np.uint32(inst_base[i]); p[3]["first_warp"] += np.uint32(warp_base[i])
        p[3]["text_off"] += np.uint64(cuts[i])
        p[4]["first_inst"] += np.uint32(inst_base[i])
        lo += len(p[0])
    assert np.array_equal(inst, want["inst"])
    assert np.array_equal(np.concatenate([p[1] for p in parts]), want["rows"])
    assert np.array_equal(np.concatenate([p[2] for p in parts]), want["txn"])
    assert np.array_equal(np.concatenate([p[3] for p in parts]), want["tb"])
    assert np.array_equal(np.concatenate([p[4] for p in parts]), want["warp"])
    for f in T.COUNTERS.names:
        if f not in ("reserved",):
            assert np.array_equal(acc[0][f], want["counters"][f]), f
    tp.kernel_finalizer(k)


def run_streamed(tp, tmp_path, shard_bytes, **gen):
    """The pipelined host -> host path returns exactly what the one-piece path returns."""
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, **gen)
    k = tp.parse_kernel_info(path)
    whole = k.run()
    want = {n: getattr(whole, n).copy() for n in ("inst", "rows", "txn_off", "txn", "tb", "warp")}
    wc = np.zeros(1, dtype=T.COUNTERS)
    wc[0] = whole.counters
    d = k.run_streamed(shard_bytes)
    for n, w in want.items():
        assert np.array_equal(getattr(d, n), w), n
    for f in T.COUNTERS.names:
        if f != "reserved":
            assert np.array_equal(d.counters[f], wc[0][f]), f
    # and a second time, reusing the pipeline's buffers
    d = k.run_streamed(shard_bytes)
    for n, w in want.items():
        assert np.array_equal(getattr(d, n), w), n
    tp.kernel_finalizer(k)


def run_rare_paths(tp, tmp_path):
    """The kernel's rare per-chunk paths really run (ti_kernel_path_stats) and give oracle-identical
    results: staging overflow + recompute, headers that cannot be parked (early resolve of the record
    prefixes), chunks with warp-path instructions (early resolve of the row index) and a full
    base+stride descriptor list."""
    tmp = str(tmp_path)

    def decode(img, name):
        path = os.path.join(tmp, name + ".traceg")
        open(path, "wb").write(img)
        k = tp.parse_kernel_info(path)
        d = k.run()
        stats = k.path_stats()
        o = T.load_dump(harness.oracle_dump(path, tmp, name=name + ".tidump"))
        check_against_dump(d, o, offsets=harness.oracle_kind() == "port", tp=tp)
        tp.kernel_finalizer(k)
        return stats

    # [0] the warp path stages more transactions in a chunk than the pool holds (base+delta lines, 32 sectors each),
    # next to base+stride lines whose descriptors share the pool; [2] such chunks resolve the row index early
    lines = ["%04x ffffffff 0 STG.E.64.SYS 2 R4 R6 8 2 0x%x %s 0" % (16 * i, 0x7F1000000000 + 8192 * i, " ".join(["260"] * 31)) for i in range(60)]
    lines += ["%04x ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x%x 128 0" % (0x2000 + 16 * i, 0x7F0000000000 + 4096 * i) for i in range(60)]
    st = decode(edge_cases.image([(b, 0, 0, [lines, lines]) for b in range(2)]), "overflow")
    assert st[0] > 0 and st[2] > 0, list(st)
    # [1] more thread-block / warp headers in a chunk than the parking area holds
    assert decode(edge_cases.VALID["edge_many_headers"][0], "headers")[1] > 0
    # [3] more base+stride instructions in a chunk than descriptors fit
    lines = ["%x 1 0 LDG 0 4 1 0x%x 4 0" % (16 * i, 0x7F3000000000 + 256 * i) for i in range(1500)]   # ~30 bytes: > 256 per chunk
    assert decode(edge_cases.image([(0, 0, 0, [lines, []])]), "notes_full")[3] > 0


def check_compact(k, shard_bytes=0):
    """The compact view (ti_kernel_decode_streamed_compact), expanded on the host by the library's own
    ti_expand_row / ti_expand_txns, is bit-identical to the full view of the same kernel file."""
    whole = k.run()
    want = {n: getattr(whole, n).copy() for n in ("inst", "rows", "txn_off", "txn", "tb", "warp")}
    wc = np.zeros(1, dtype=T.COUNTERS)
    wc[0] = whole.counters
    c = k.run_streamed_compact(shard_bytes)
    for n in ("inst", "tb", "warp"):
        assert np.array_equal(getattr(c, n), want[n]), n
    assert len(c.desc) == len(want["rows"]) and c.counts["n_txn"] == len(want["txn"])
    assert np.array_equal(c.expand_rows(), want["rows"]), "rows"
    txn, off = c.expand_txns()
    assert np.array_equal(off, want["txn_off"]), "txn_off"
    assert np.array_equal(txn, want["txn"]), "txn"
    for f in T.COUNTERS.names:
        if f != "reserved":
            assert np.array_equal(c.counters[f], wc[0][f]), f
    return c


def run_compact(tp, tmp_path, shard_bytes=0, **gen):
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, **gen)
    k = tp.parse_kernel_info(path)
    c = check_compact(k, shard_bytes)
    tp.kernel_finalizer(k)
    return c


def run_compact_golden(tp, name):
    over = edge_cases.VALID.get(name, (None, {}))[1]
    p, own = parser_with(tp, over)
    try:
        k = p.parse_kernel_info(os.path.join(GOLDEN, name + ".traceg"))
        check_compact(k)
        p.kernel_finalizer(k)
    finally:
        if own:
            p.close()


def run_pipeline_stats(tp, tmp_path, **gen):
    """register-file read / write totals and the decode counters against the restated counting rule
    applied to the REFERENCE's own trace_warp_inst_t fields (ref_oracle prints them in its summary line)."""
    import json
    import subprocess
    if harness.oracle_kind() != "reference":
        import pytest
        pytest.skip("needs the reference-built oracle")
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, **gen)
    k = tp.parse_kernel_info(path)
    k.run()
    for group in (4, 8):
        r = subprocess.run([harness.REF_ORACLE, "time", path, "", "convert"], capture_output=True, text=True,
                           env=dict(os.environ, TI_RF_GATING_GROUP=str(group)))
        want = json.loads(r.stdout.strip().splitlines()[-1])["pstats"]
        plain, gated = k.pipeline_stats(0, group), k.pipeline_stats(1, group)
        for got, idx in ((plain, 0), (gated, 1)):
            assert got["regfile_reads"] == want["regfile_reads"][idx] and got["regfile_writes"] == want["regfile_writes"][idx], (group, idx, got, want)
            assert got["non_rf_operands"] == 0 == want["non_rf_operands"]
            for f in ("decoded_insn", "int_decoded_insn", "fp_decoded_insn"):
                assert got[f] == want[f], (f, got, want)
    tp.kernel_finalizer(k)
