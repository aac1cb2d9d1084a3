"""CPU-only: the SAME kernel source (csrc/*.cuh) compiled with g++ against the
lock-step SIMT emulator in tests/simt_emu, checked against the oracle.  This is test
infrastructure for debugging kernel logic without a GPU -- the product never loads
it; the parity tests proper are tests/test_gpu_parity.py (-m gpu)."""
import pytest

import edge_cases
import parity_cases


@pytest.mark.parametrize("name", parity_cases.golden_names())
def test_emu_golden(emu_parser, name):
    parity_cases.run_golden(emu_parser, name)


@pytest.mark.parametrize("profile,kw", [
    (2, dict(n_tb=12, insts_per_warp=25)),                      # several chunks, look-back across CTAs
    (3, dict(n_tb=3, insts_per_warp=12, block_threads=64)),     # 32 sectors per instruction
    (5, dict(n_tb=16, insts_per_warp=10, block_threads=128, binary_version=86)),
])
def test_emu_generated(emu_parser, tmp_path, profile, kw):
    parity_cases.run_generated(emu_parser, tmp_path, profile=profile, seed=40 + profile, **kw)


@pytest.mark.parametrize("name", sorted(edge_cases.ERRORS))
def test_emu_errors(emu_parser, tmp_path, name):
    parity_cases.run_error(emu_parser, tmp_path, name)


def test_emu_sharded(emu_parser, tmp_path):
    parity_cases.run_sharded(emu_parser, tmp_path, 3, profile=5, seed=9, n_tb=12, insts_per_warp=8, block_threads=96)


def test_emu_staging_overflow_and_capacity_retry(emu_parser, tmp_path):
    """More transactions per chunk than the smem staging area (recompute path) and than
    the default output capacity (exact-size re-decode)."""
    import os
    import harness
    import tiformat as T
    lines = ["%04x ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x%x 128 0" % (16 * i, 0x7F0000000000 + 4096 * i) for i in range(150)]
    lines += ["%04x ffffffff 0 STG.E.64.SYS 2 R4 R6 8 2 0x%x %s 0" % (0x2000 + 16 * i, 0x7F1000000000 + 8192 * i, " ".join(["260"] * 31)) for i in range(40)]
    img = edge_cases.image([(b, 0, 0, [lines, lines]) for b in range(2)])
    path = os.path.join(str(tmp_path), "k.traceg")
    open(path, "wb").write(img)
    k = emu_parser.parse_kernel_info(path)
    d = k.run()
    assert len(d.txn) > 32 * 150 * 4
    o = T.load_dump(harness.oracle_dump(path, str(tmp_path)))
    parity_cases.check_against_dump(d, o, offsets=harness.oracle_kind() == "port")
    emu_parser.kernel_finalizer(k)


def test_emu_streamed(emu_parser, tmp_path):
    parity_cases.run_streamed(emu_parser, tmp_path, 40000, profile=5, seed=21, n_tb=24, insts_per_warp=8, block_threads=96)


def test_emu_rare_paths(emu_parser, tmp_path):
    parity_cases.run_rare_paths(emu_parser, tmp_path)


def test_emu_fast_pass_covers_tracer_lines(emu_parser, tmp_path):
    """P2's fast pass (csrc/ti_fast.cuh) takes the tracer-shaped lines once the CTA's opcode memo is
    warm: on a many-chunk corpus most lines must have gone through it (ti_counters.reserved[4]) and the
    result must still equal the oracle's (run_generated)."""
    d = parity_cases.run_generated(emu_parser, tmp_path, profile=2, seed=77, n_tb=40, insts_per_warp=25)
    assert int(d.counters["reserved"][4]) > 0.5 * len(d.inst), (int(d.counters["reserved"][4]), len(d.inst))
    d = parity_cases.run_generated(emu_parser, tmp_path, profile=5, seed=78, n_tb=60, insts_per_warp=12, block_threads=128)
    assert int(d.counters["reserved"][4]) > 0.3 * len(d.inst), (int(d.counters["reserved"][4]), len(d.inst))


def test_emu_streamed_empty_tail_shards_and_error_offset(emu_parser, tmp_path):
    """The host -> host pipeline with shards that come out empty (one thread block longer than a shard): the
    CSR sentinel txn_off[n_mem] must survive, and the result must equal the one-piece decode."""
    import os
    import numpy as np
    import harness
    lines = ["%04x ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x%x 4 0" % (16 * i, 0x7F0000000000 + 256 * i) for i in range(400)]
    img = edge_cases.image([(0, 0, 0, [lines, lines])])
    path = os.path.join(str(tmp_path), "one_big_tb.traceg")
    open(path, "wb").write(img)
    k = emu_parser.parse_kernel_info(path)
    whole = k.run()
    want_off = whole.txn_off.copy()
    d = k.run_streamed(len(img) // 3)          # cuts come out as [0, len, len, len]
    assert np.array_equal(d.txn_off, want_off)
    assert int(d.txn_off[-1]) == len(d.txn)
    emu_parser.kernel_finalizer(k)


@pytest.mark.parametrize("name", parity_cases.golden_names())
def test_emu_compact_golden(emu_parser, name):
    parity_cases.run_compact_golden(emu_parser, name)


def test_emu_compact_generated(emu_parser, tmp_path):
    """Compact results expand to the full ones: a base+stride corpus (nothing but descriptors travels) and
    the mixed one (rows and transactions of list / delta instructions are kept), over several shards."""
    c = parity_cases.run_compact(emu_parser, tmp_path, profile=2, seed=5, n_tb=12, insts_per_warp=25)
    assert c.n_full <= 0.02 * len(c.desc) and c.n_txn_stored <= 0.02 * c.counts["n_txn"]   # (a warp with more than 32 such lines sends the rest down the warp path)
    c = parity_cases.run_compact(emu_parser, tmp_path, shard_bytes=30000, profile=5, seed=6, n_tb=24, insts_per_warp=8, block_threads=96)
    assert 0 < c.n_full < len(c.desc) and 0 < c.n_txn_stored < c.counts["n_txn"]


def test_emu_pipeline_stats(emu_parser, tmp_path):
    parity_cases.run_pipeline_stats(emu_parser, tmp_path, profile=5, seed=71, n_tb=20, insts_per_warp=11, block_threads=96)
