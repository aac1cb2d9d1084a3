"""GPU parity tests: the CUDA library, called through its C ABI, against the CPU
oracle (the reference-built oracle/_ref when it travelled with the repo, else the
port) and against the reference-produced fixtures in tests/golden."""
import itertools
import os

import numpy as np
import pytest

import edge_cases
import harness
import parity_cases
import tiformat as T

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", parity_cases.golden_names())
def test_golden(cuda_parser, name):
    parity_cases.run_golden(cuda_parser, name)


@pytest.mark.parametrize("profile,kw", [
    (1, dict(n_tb=64, insts_per_warp=40, grid=(1, 64, 1), block=(16, 16, 1))),
    (1, dict(n_tb=64, insts_per_warp=40, grid=(1, 64, 1), block=(16, 16, 1), variant=1)),
    (2, dict(n_tb=2000, insts_per_warp=25)),
    (3, dict(n_tb=300, insts_per_warp=20)),
    (4, dict(n_tb=128, insts_per_warp=32, binary_version=80)),
    (6, dict(n_tb=1, insts_per_warp=5, binary_version=80)),
])
def test_generated_profiles(cuda_parser, tmp_path, profile, kw):
    parity_cases.run_generated(cuda_parser, tmp_path, profile=profile, seed=profile, **kw)


CFGS = {
    "qv100": {},
    "fermi": dict(coalesce_arch=20),
    "fermi_skip_l1": dict(coalesce_arch=20, gmem_skip_L1D=1),
    "parts2_bcast": dict(mem_warp_parts=2, shmem_limited_broadcast=1, num_shmem_bank=16, clock_gated_lanes=0, num_dp_units=0),
    "arch13_parts4": dict(coalesce_arch=13, mem_warp_parts=4, clock_gated_lanes=0),
}


@pytest.mark.parametrize("cfg,bv", list(itertools.product(CFGS, [35, 61, 70, 75, 86])))
def test_soup_all_architectures_and_configs(cuda_parser, tmp_path, cfg, bv):
    parity_cases.run_generated(cuda_parser, tmp_path, over=CFGS[cfg], profile=5, seed=7000 + bv, n_tb=200,
                               insts_per_warp=12, binary_version=bv, block_threads=128)


def test_soup_trace_version_2(cuda_parser, tmp_path):
    parity_cases.run_generated(cuda_parser, tmp_path, profile=5, seed=77, n_tb=100, insts_per_warp=12, trace_version=2)


@pytest.mark.parametrize("name", sorted(edge_cases.ERRORS))
def test_malformed_lines_are_flagged(cuda_parser, tmp_path, name):
    parity_cases.run_error(cuda_parser, tmp_path, name)


@pytest.mark.parametrize("n_shards", [2, 5])
def test_byte_range_shards(cuda_parser, tmp_path, n_shards):
    parity_cases.run_sharded(cuda_parser, tmp_path, n_shards, profile=5, seed=13, n_tb=300, insts_per_warp=10, block_threads=128)


def test_staging_overflow_and_capacity_retry(cuda_parser, tmp_path):
    """Stride-128 accesses give 32 sectors per instruction from 45-byte lines: more
    transactions per chunk than the smem staging area and than the default output
    capacity -- exercises the recompute path and the exact-size re-decode."""
    lines = ["%04x ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x%x 128 0" % (16 * i, 0x7F0000000000 + 4096 * i) for i in range(400)]
    img = edge_cases.image([(b, 0, 0, [lines] * 2) for b in range(16)])  # block dim 64: two warps per block
    path = os.path.join(str(tmp_path), "k.traceg")
    open(path, "wb").write(img)
    k = cuda_parser.parse_kernel_info(path)
    d = k.run()
    assert len(d.txn) == 32 * len(d.inst) == 32 * 400 * 2 * 16
    o = T.load_dump(harness.oracle_dump(path, str(tmp_path)))
    parity_cases.check_against_dump(d, o, offsets=harness.oracle_kind() == "port")
    cuda_parser.kernel_finalizer(k)


def test_streamed_empty_tail_shards(cuda_parser, tmp_path):
    """One thread block longer than a shard: ti_plan_shards returns empty tail shards; the pipelined path must
    still return the one-piece result, CSR sentinel included."""
    lines = ["%04x ffffffff 1 R2 LDG.E.SYS 1 R4 4 1 0x%x 4 0" % (16 * i, 0x7F0000000000 + 256 * i) for i in range(400)]
    img = edge_cases.image([(0, 0, 0, [lines, lines])])
    path = os.path.join(str(tmp_path), "one_big_tb.traceg")
    open(path, "wb").write(img)
    k = cuda_parser.parse_kernel_info(path)
    want = k.run().txn_off.copy()
    d = k.run_streamed(len(img) // 3)
    assert np.array_equal(d.txn_off, want) and int(d.txn_off[-1]) == len(d.txn)
    cuda_parser.kernel_finalizer(k)


def test_full_size_properties(cuda_parser):
    """Size-independent properties at BASELINE.json's full config-2 size (1.05 GB of text, 100 000
    thread blocks -- far too large for the CPU oracle to check in seconds): determinism, CSR
    consistency, framing, counter identities, and decode -> re-encode round trip of the address rows."""
    N_TB = 100000
    img = harness.generate_bytes(profile=2, seed=2, n_tb=N_TB, insts_per_warp=25)
    hdr = cuda_parser.parse_header(img[:4096])
    body = img[hdr.body_offset:]
    k = cuda_parser.open_buffer(body, hdr)
    d = k.run()
    n_inst = N_TB * 8 * 25
    assert len(d.inst) == n_inst and len(d.tb) == N_TB and len(d.warp) == 8 * N_TB
    c = d.counters
    assert int(c["hist"].sum()) == n_inst == int(c["op_count"].sum()) == int(c["n_inst"])
    assert int(c["n_begin_tb"]) == int(c["n_end_tb"]) == N_TB
    # CSR offsets are monotone and end at n_txn; every record points at its own row
    off = d.txn_off.astype(np.int64)
    assert off[0] == 0 and off[-1] == len(d.txn) and np.all(np.diff(off) >= 0)
    has = d.inst["mem_idx"] != 0xFFFFFFFF
    assert np.array_equal(d.inst["mem_idx"][has], np.arange(int(has.sum()), dtype=np.uint32))
    # warp headers tile the instruction array
    assert np.array_equal(d.warp["first_inst"], np.arange(8 * N_TB, dtype=np.uint32) * 25)
    # every transaction of this corpus is one 32 B sector covering lanes that hit it
    assert np.all(d.txn["size"] == 32) and np.all(d.txn["addr"] % 32 == 0)
    mask = d.inst["mask"][has]
    # (every 8th row: the full 2 GB array would need several times that in numpy temporaries)
    rows = d.rows[::8]
    act = ((mask[::8, None] >> np.arange(32, dtype=np.uint32)) & 1).astype(bool)
    # round trip: base + 4 * lane for active lanes of the first run, 0 elsewhere (generator's rule)
    base = rows[np.arange(len(rows)), act.argmax(axis=1)]
    lane0 = act.argmax(axis=1)
    want = np.where(act, base[:, None] + 4 * (np.arange(32)[None, :] - lane0[:, None]).astype(np.int64).astype(np.uint64), 0)
    assert np.array_equal(rows[act.any(axis=1)], want[act.any(axis=1)])
    # lanes of a row's transactions partition its active, non-zero lanes
    cnt = np.diff(off)
    lanes = np.bitwise_or.reduceat(d.txn["lane_mask"], off[:-1][cnt > 0]) if len(d.txn) else np.array([])
    assert np.array_equal(lanes, mask[cnt > 0])
    # determinism: a second decode is bit-identical
    inst1, txn1 = d.inst.copy(), d.txn.copy()
    d2 = k.run()
    assert np.array_equal(inst1, d2.inst) and np.array_equal(txn1, d2.txn)
    # the pipelined host-to-host entry point returns the same bytes as the one-piece path
    rows1, off1, tb1, warp1 = d2.rows.copy(), d2.txn_off.copy(), d2.tb.copy(), d2.warp.copy()
    cnt1 = np.array(d2.counters, dtype=T.COUNTERS)
    d3 = k.run_streamed(32 << 20)
    assert np.array_equal(inst1, d3.inst) and np.array_equal(txn1, d3.txn) and np.array_equal(rows1, d3.rows)
    assert np.array_equal(off1, d3.txn_off) and np.array_equal(tb1, d3.tb) and np.array_equal(warp1, d3.warp)
    for f in T.COUNTERS.names:
        if f not in ("reserved", "text_bytes"):
            assert np.array_equal(cnt1[f], d3.counters[f]), f
    # a checksum of checksums: the counter blocks of four byte-range shards add up to the whole file's
    cuts = cuda_parser.plan_shards(body, 4)
    acc = np.zeros(1, dtype=T.COUNTERS)
    n_parts = 0
    for b, e in zip(cuts[:-1], cuts[1:]):
        part = k.run(b, e)
        c1 = np.zeros(1, dtype=T.COUNTERS)
        c1[0] = part.counters
        cuda_parser.lib.ti_counters_add(acc.ctypes.data, c1.ctypes.data)
        n_parts += len(part.inst)
    assert n_parts == n_inst
    for f in T.COUNTERS.names:
        if f not in ("reserved", "text_bytes"):
            assert np.array_equal(acc[0][f], cnt1[f]), f
    cuda_parser.kernel_finalizer(k)


def test_config2_full_size_vs_reference(cuda_parser, tmp_path):
    """BASELINE.json config 2 at its full size (1.05 GB of text, 20 M instructions) against the reference-built
    oracle, array by array: the body is cut into one whole-thread-block slice per host core, `ref_oracle dump`
    runs on the slices in parallel, and every slice's records / address rows / CSR offsets / transactions /
    thread-block and warp tables are compared bit for bit with the matching part of ONE one-piece GPU decode
    (indices that the oracle counts from the start of its slice are shifted by what the earlier slices hold).
    The integer counters of the slices must add up to the GPU's.  Pins what no smaller test reaches: offsets
    beyond 2^24 instructions, the 31-bit prefix fields of the look-back words, look-back chains over more
    chunks than resident CTAs."""
    import subprocess
    import time
    import bench
    t_start = time.time()
    N_TB = 100000
    img = harness.generate_bytes(profile=2, seed=2, n_tb=N_TB, insts_per_warp=25)
    hdr = cuda_parser.parse_header(img[:4096])
    body = img[hdr.body_offset:]
    k = cuda_parser.open_buffer(body, hdr)
    d = k.run()
    kind = harness.oracle_kind()
    exe = harness.REF_ORACLE if kind == "reference" else harness.PORT_ORACLE
    n_slices = max(2, min(os.cpu_count() or 2, 32))
    paths, total = bench.split_sample(img, hdr.body_offset, n_slices, (len(body) + n_slices - 1) // n_slices, str(tmp_path))
    assert total == len(body)
    acc = np.zeros(1, dtype=T.COUNTERS)
    i0 = m0 = x0 = b0 = w0 = 0          # instructions, rows, transactions, thread blocks, warps before the slice
    text0 = 0
    wave = max(1, n_slices // 2)        # two waves: half of the dumps (about 2 GB) on disk at a time
    for lo in range(0, len(paths), wave):
        part = paths[lo:lo + wave]
        procs = [subprocess.Popen([exe, "dump", p, p + ".tidump"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL) for p in part]
        assert all(pr.wait() == 0 for pr in procs)
        for p in part:
            o = T.load_dump(p + ".tidump")
            ni, nm, nx, nb, nw = len(o.inst), len(o.rows), len(o.txn), len(o.tb), len(o.warp)
            g = d.inst[i0:i0 + ni]
            assert len(g) == ni
            for f in T.INST.names:
                if f == "opc_off":
                    continue                      # the oracles keep the opcode strings, not offsets
                want = o.inst[f]
                if f == "mem_idx":
                    want = np.where(want != 0xFFFFFFFF, want + np.uint32(m0), want)
                assert np.array_equal(g[f], want), (p, "inst." + f)
            assert np.array_equal(d.rows[m0:m0 + nm], o.rows), (p, "rows")
            assert np.array_equal(d.txn_off[m0:m0 + nm + 1].astype(np.int64) - x0, o.txn_off.astype(np.int64)), (p, "txn_off")
            assert np.array_equal(d.txn[x0:x0 + nx], o.txn), (p, "txn")
            gt, gw = d.tb[b0:b0 + nb], d.warp[w0:w0 + nw]
            for f in ("x", "y", "z"):
                assert np.array_equal(gt[f], o.tb[f]), (p, "tb." + f)
            assert np.array_equal(gt["first_warp"], o.tb["first_warp"] + np.uint32(w0)) and np.array_equal(gt["first_inst"], o.tb["first_inst"] + np.uint32(i0)), (p, "tb")
            assert np.array_equal(gw["warp_id"], o.warp["warp_id"]) and np.array_equal(gw["n_insts"], o.warp["n_insts"]), (p, "warp")
            assert np.array_equal(gw["first_inst"], o.warp["first_inst"] + np.uint32(i0)), (p, "warp.first_inst")
            # opcode tokens through the GPU's text offsets: the slice's first and last instruction and 200 random ones
            for j in [0, ni - 1] + list(np.random.default_rng(lo).integers(0, ni, 200)):
                r = g[int(j)]
                assert bytes(body[int(r["opc_off"]):int(r["opc_off"]) + int(r["opcode_len"])]) == o.opcodes[int(j)], (p, "opcode", j)
            c1 = np.zeros(1, dtype=T.COUNTERS)
            c1[0] = o.counters
            cuda_parser.lib.ti_counters_add(acc.ctypes.data, c1.ctypes.data)
            i0 += ni; m0 += nm; x0 += nx; b0 += nb; w0 += nw
            os.remove(p + ".tidump")
            os.remove(p)
    assert (i0, m0, x0, b0, w0) == (len(d.inst), len(d.rows), len(d.txn), len(d.tb), len(d.warp))
    for f in T.COUNTERS.names:
        if f not in ("reserved", "text_bytes", "n_begin_tb", "n_end_tb", "n_warp", "n_mem"):
            assert np.array_equal(acc[0][f], d.counters[f]), f
    cuda_parser.kernel_finalizer(k)
    print("full-size differential test: %d slices, oracle=%s, %.1f s" % (len(paths), kind, time.time() - t_start))


@pytest.mark.parametrize("name", [n for n in parity_cases.golden_names() if n.startswith("edge_")])
def test_compact_golden(cuda_parser, name):
    """The stride quirks of trace_parser.cc:86-105 (zero after the first run, u64 wrap, empty masks ...) through the
    compact view: expanded == full."""
    parity_cases.run_compact_golden(cuda_parser, name)


def test_compact_expands_to_full(cuda_parser, tmp_path):
    c = parity_cases.run_compact(cuda_parser, tmp_path, shard_bytes=1 << 20, profile=2, seed=2, n_tb=400, insts_per_warp=25)
    assert c.n_full <= 0.02 * len(c.desc) and c.d2h_bytes < 0.3 * (32 * len(c.inst) + 260 * len(c.desc) + 32 * c.counts["n_txn"])
    c = parity_cases.run_compact(cuda_parser, tmp_path, shard_bytes=300000, profile=5, seed=31, n_tb=300, insts_per_warp=10, block_threads=128)
    assert 0 < c.n_full < len(c.desc)
    parity_cases.run_compact(cuda_parser, tmp_path, profile=3, seed=3, n_tb=100, insts_per_warp=12)


def test_compact_config2_full_size(cuda_parser):
    """BASELINE config 2 at full size through the compact path: vectorised expansion of every descriptor equals the
    full view's rows, the records are identical, and less than a quarter of the bytes travel."""
    img = harness.generate_bytes(profile=2, seed=2, n_tb=100000, insts_per_warp=25)
    hdr = cuda_parser.parse_header(img[:4096])
    body = img[hdr.body_offset:]
    k = cuda_parser.open_buffer(body, hdr)
    full = k.run_streamed(64 << 20)
    inst, rows, txn, off = full.inst.copy(), full.rows.copy(), full.txn.copy(), full.txn_off.copy()
    c = k.run_streamed_compact(64 << 20)
    assert np.array_equal(c.inst, inst)
    d = c.desc
    is_full = (d["meta"] >> 29 & 1).astype(bool)
    first, n = (d["meta"] & 31).astype(np.int64), (d["meta"] >> 5 & 63).astype(np.int64)
    lanes = np.arange(32, dtype=np.int64)[None, :] - first[:, None]
    for lo in range(0, len(d), 1 << 20):      # in slabs: the whole expansion would need several GB of temporaries
        hi = min(len(d), lo + (1 << 20))
        want = rows[lo:hi]
        got = np.where((lanes[lo:hi] >= 0) & (lanes[lo:hi] < n[lo:hi, None]),
                       d["base"][lo:hi, None] + (d["stride"][lo:hi, None].astype(np.int64) * lanes[lo:hi]).astype(np.uint64), np.uint64(0))
        f = is_full[lo:hi]
        if f.any():
            got[f] = c.rows_full[d["base"][lo:hi][f].astype(np.int64)]
        assert np.array_equal(got, want), lo
    # transactions: the library's own expansion on a sample, the stored ones in full
    rng = np.random.default_rng(1)
    for r in rng.integers(0, len(d), 2000):
        t, o = c.expand_txns(int(r), int(r) + 1)
        assert np.array_equal(t, txn[off[r]:off[r + 1]]), r
    assert c.d2h_bytes < 0.25 * (32 * len(inst) + 256 * len(rows) + 4 * len(off) + 32 * len(txn))
    cuda_parser.kernel_finalizer(k)


@pytest.mark.parametrize("shard_bytes", [1 << 20, 300000])
def test_streamed_host_path(cuda_parser, tmp_path, shard_bytes):
    parity_cases.run_streamed(cuda_parser, tmp_path, shard_bytes, profile=5, seed=31, n_tb=800, insts_per_warp=10, block_threads=128)


@pytest.mark.parametrize("profile,bv", [(5, 70), (5, 86), (2, 70), (1, 70)])
def test_pipeline_stats(cuda_parser, tmp_path, profile, bv):
    parity_cases.run_pipeline_stats(cuda_parser, tmp_path, profile=profile, seed=300 + profile + bv, n_tb=150, insts_per_warp=13,
                                    block_threads=128, binary_version=bv)


def test_rare_paths(cuda_parser, tmp_path):
    parity_cases.run_rare_paths(cuda_parser, tmp_path)   # more chunks than resident CTAs (444)
