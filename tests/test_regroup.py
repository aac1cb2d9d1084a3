"""Writer side of the wire format: kernel-N.trace -> kernel-N.traceg.

Oracle: the reference's own post-traces-processing.cpp built as oracle/_ref/ptp_ref, its
Python restatement rawtrace.ptp_port (pinned against it here and against the committed
reference-produced fixtures tests/golden/regroup_*).  The CUDA path (ti_regroup_*) must
return the same bytes; CPU runs use the SIMT-emulated build of the same kernel source.
"""
import glob
import os

import numpy as np
import pytest

import harness
import rawtrace
from accel_sim_framework_distributed_b200 import _ffi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
HAVE_REF = os.access(rawtrace.PTP_REF, os.X_OK)


def cases():
    """name -> raw trace bytes"""
    g5 = bytes(harness.generate_bytes(profile=5, seed=61, n_tb=9, insts_per_warp=7, block_threads=96))
    g2 = bytes(harness.generate_bytes(profile=2, seed=62, n_tb=40, insts_per_warp=25))          # > 2 count chunks
    g3 = bytes(harness.generate_bytes(profile=3, seed=63, n_tb=3, insts_per_warp=6, block_threads=64))
    gb = bytes(harness.generate_bytes(profile=1, seed=64, n_tb=3, insts_per_warp=12, grid=(1, 3, 1), block=(16, 16, 1)))
    c = {
        "soup": rawtrace.ungroup(g5, 1),
        "stream_multi_chunk": rawtrace.ungroup(g2, 2, window=5),
        "divergent_long_lines": rawtrace.ungroup(g3, 3),
        "grid_y": rawtrace.ungroup(gb, 4, window=2),
        "empty_blocks": rawtrace.ungroup(g5, 5, drop_tb=(0, 4, 8)),
        "blank_lines_inside": rawtrace.ungroup(g5, 6, blank_every=5),
        "no_trailing_newline": rawtrace.ungroup(g5, 7)[:-1],
        "sequential_blocks": rawtrace.ungroup(g5, 8, window=1),
    }
    hdr_only = c["soup"][:c["soup"].index(b"#traces format")]
    c["header_only"] = hdr_only
    c["one_line"] = hdr_only + b"0 0 0 1 0008 ffffffff 0 NOP 0 0\n"
    return c


CASES = cases()


def oracle(raw, tmp):
    return rawtrace.ptp_reference(raw, tmp)[0] if HAVE_REF else rawtrace.ptp_port(raw)


# ------------------------------------------------------------------ oracle pinning (CPU)
@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref/ptp_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name", sorted(CASES))
def test_port_matches_reference_tool(tmp_path, name):
    want, _, _ = rawtrace.ptp_reference(CASES[name], tmp_path)
    assert rawtrace.ptp_port(CASES[name]) == want


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "regroup_*.trace"))))
def test_port_matches_golden(path):
    assert rawtrace.ptp_port(open(path, "rb").read()) == open(path + "g", "rb").read()


def test_golden_fixtures_exist():
    assert len(glob.glob(os.path.join(GOLDEN, "regroup_*.trace"))) >= 4


# ------------------------------------------------------------------ shared checks
def check_case(tp, tmp, name):
    raw = CASES[name]
    got, st = tp.regroup(raw)
    want = oracle(raw, tmp)
    assert got == want, "first difference at byte %d" % next((i for i, (a, b) in enumerate(zip(got, want)) if a != b), min(len(got), len(want)))
    assert st.out_bytes == len(want) and st.n_lines == raw.count(b"\n") + 1


def check_golden(tp, path):
    got, _ = tp.regroup(open(path, "rb").read())
    assert got == open(path + "g", "rb").read()


def check_errors(tp):
    hdr = CASES["header_only"]
    bad = {
        "x outside the grid": hdr + b"999999 0 0 0 0008 ffffffff 0 NOP 0 0\n",
        "warp outside the block": hdr + b"0 0 0 77 0008 ffffffff 0 NOP 0 0\n",
        "not four decimals": hdr + b"0 0 zz 0 0008 ffffffff 0 NOP 0 0\n",
        "nothing after the warp id": hdr + b"0 0 0 0\n",
        "instruction before the dims": b"0 0 0 0 0008 ffffffff 0 NOP 0 0\n" + hdr,
        "no dims at all": b"-kernel name = k\n0 0 0 0 0008 ffffffff 0 NOP 0 0\n",
    }
    for why, raw in bad.items():
        with pytest.raises(_ffi.TraceIngestError) as e:
            tp.regroup(raw)
        assert e.value.status == _ffi.TI_ERR_PARSE, why
    # and the context is still usable afterwards
    got, _ = tp.regroup(CASES["one_line"])
    assert got == rawtrace.ptp_port(CASES["one_line"])


def check_round_trip(tp, tmp):
    """generator -> raw trace -> regroup -> decode equals decoding the generator's grouped file."""
    img = harness.generate_bytes(profile=5, seed=91, n_tb=10, insts_per_warp=9, block_threads=128)
    regrouped, _ = tp.regroup(rawtrace.ungroup(bytes(img), 9))
    res = []
    for image in (np.frombuffer(regrouped, dtype=np.uint8), img):
        hdr = tp.parse_header(image[:8192])
        k = tp.open_buffer(np.ascontiguousarray(image[hdr.body_offset:]), hdr)
        d = k.run()
        inst = d.inst.copy()
        inst["opc_off"] = 0                                   # text offsets differ (header lines), nothing else may
        res.append((inst, d.rows.copy(), d.txn.copy(), d.tb[["x", "y", "z", "first_warp", "first_inst"]].copy(), d.warp.copy()))
        tp.kernel_finalizer(k)
    for a, b in zip(*res):
        assert np.array_equal(a, b)


# ------------------------------------------------------------------ CPU: emulated kernels
@pytest.mark.parametrize("name", sorted(CASES))
def test_emu_regroup(emu_parser, tmp_path, name):
    check_case(emu_parser, tmp_path, name)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "regroup_*.trace"))))
def test_emu_regroup_golden(emu_parser, path):
    check_golden(emu_parser, path)


def test_emu_regroup_errors(emu_parser):
    check_errors(emu_parser)


def test_emu_regroup_round_trip(emu_parser, tmp_path):
    check_round_trip(emu_parser, tmp_path)


def test_regroup_commandlist(emu_parser, tmp_path):
    cmds = ("MemcpyHtoD,0x00007f0000000000,4096", "ncclCommInitAll", "kernel-1.trace", "", "ncclGroupStart", "ncclAllReduce",
            "ncclGroupEnd", "kernel-1.trace", "ncclCommDestroy")
    text = ("\n".join(cmds) + "\n").encode()
    got = emu_parser.regroup_commandlist(text)
    if HAVE_REF:
        _, want, _ = rawtrace.ptp_reference(CASES["one_line"], tmp_path, commands=cmds)
        assert got == want
    assert got == b"".join((c + ("g" if c.startswith("kernel") else "")).encode() + b"\n" for c in cmds if c)
    with pytest.raises(_ffi.TraceIngestError):
        emu_parser.regroup_commandlist(b"kernel-1.trace\nfrobnicate\n")


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_regroup(cuda_parser, tmp_path, name):
    check_case(cuda_parser, tmp_path, name)


@pytest.mark.gpu
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "regroup_*.trace"))))
def test_gpu_regroup_golden(cuda_parser, path):
    check_golden(cuda_parser, path)


@pytest.mark.gpu
def test_gpu_regroup_errors(cuda_parser):
    check_errors(cuda_parser)


@pytest.mark.gpu
def test_gpu_regroup_round_trip(cuda_parser, tmp_path):
    check_round_trip(cuda_parser, tmp_path)


@pytest.mark.gpu
def test_gpu_regroup_large_properties(cuda_parser):
    """64 MB raw trace (too slow for the Python port): the regrouped file must decode, hold every
    line exactly once, and be what the reference tool writes when that is available on the box."""
    raw = np.asarray(harness.generate_bytes(profile=2, seed=5, n_tb=6000, insts_per_warp=25, raw=1))
    got, st = cuda_parser.regroup(raw)
    n_inst = 6000 * 8 * 25
    assert st.n_tb_written == 6000 and st.n_tb_empty == 0 and st.n_warp_empty == 0
    image = np.frombuffer(got, dtype=np.uint8)
    hdr = cuda_parser.parse_header(image[:8192])
    k = cuda_parser.open_buffer(np.ascontiguousarray(image[hdr.body_offset:]), hdr)
    d = k.run()
    assert len(d.inst) == n_inst and len(d.tb) == 6000 and len(d.warp) == 48000
    assert np.array_equal(d.tb["x"], np.arange(6000, dtype=np.int32))
    assert np.array_equal(d.warp["n_insts"], np.full(48000, 25, dtype=np.uint32))
    # payload bytes are conserved: every instruction line lost exactly its "x 0 0 w " prefix
    cuda_parser.kernel_finalizer(k)
    got2, _ = cuda_parser.regroup(raw)
    assert got2 == got                                         # deterministic despite the atomics
    if HAVE_REF:
        import tempfile
        with tempfile.TemporaryDirectory() as tmp:
            want, _, _ = rawtrace.ptp_reference(raw.tobytes(), tmp)
        assert got == want


# ------------------------------------------------------------------ the drop-in command-line tool
def _tool_matches_reference(lib_path, tmp):
    """shim/post_traces_processing_b200.cc against the reference binary: same files, same stdout."""
    import shutil
    import subprocess
    libdir, libname = os.path.split(lib_path)
    tool = os.path.join(tmp, "ptp_b200")
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-I" + os.path.join(harness.ROOT, "include"), "-o", tool,
                    os.path.join(harness.ROOT, "shim", "post_traces_processing_b200.cc"),
                    "-L" + libdir, "-l:" + libname, "-Wl,-rpath," + libdir], check=True)
    klist = "MemcpyHtoD,0x00007f2a00000000,65536\nncclCommInitAll\nkernel-1.trace\n\nkernel-2.trace\nkernel-9.trace\nncclCommDestroy\nfrobnicate\nkernel-3.trace\n"
    outs = []
    for who, exe in (("ref", rawtrace.PTP_REF), ("b200", tool)):
        d = os.path.join(tmp, who, "traces")
        os.makedirs(d)
        open(os.path.join(d, "kernel-1.trace"), "wb").write(CASES["empty_blocks"])       # warnings for blocks and warps
        open(os.path.join(d, "kernel-2.trace"), "wb").write(CASES["grid_y"])
        open(os.path.join(d, "kernel-3.trace"), "wb").write(CASES["soup"])              # after the unknown command: never reached
        open(os.path.join(d, "kernelslist"), "w").write(klist)                           # kernel-9.trace does not exist
        r = subprocess.run([exe, "traces/kernelslist"], cwd=os.path.join(tmp, who), capture_output=True)
        assert r.returncode == 0, (who, r.stderr[-400:])
        files = {f: open(os.path.join(d, f), "rb").read() for f in sorted(os.listdir(d))}
        outs.append((r.stdout, files))
    assert outs[0][0] == outs[1][0]                       # stdout: "Processing file", warnings, "Undefined command"
    assert b"Warning: Thread block" in outs[0][0] and b"Warning: Warp" in outs[0][0] and b"Undefined command" in outs[0][0]
    assert sorted(outs[0][1]) == sorted(outs[1][1]) and "kernel-1.traceg" in outs[0][1] and "kernel-3.traceg" not in outs[0][1]
    for f in outs[0][1]:
        assert outs[0][1][f] == outs[1][1][f], f


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref/ptp_ref not built (needs /root/reference)")
def test_tool_matches_reference_emulated(tmp_path):
    _tool_matches_reference(harness.emu_library(), str(tmp_path))


@pytest.mark.gpu
@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref/ptp_ref not built (needs /root/reference)")
def test_tool_matches_reference_cuda(tmp_path):
    _tool_matches_reference(harness.CUDA_LIB, str(tmp_path))
