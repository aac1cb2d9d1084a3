"""The C++ drop-in (shim/trace_parser.h + trace_parser_b200.cc): the SAME driver
source (oracle/tp_driver.cc) is built once against the reference's own
trace_parser.{h,cc} and once against the shim; their complete outputs -- command list,
kernel headers, the stdout echo lines and every inst_trace_t field of every thread
block -- must be byte-identical."""
import os
import shutil
import subprocess

import pytest

import harness
import parity_cases

TP_REF = os.path.join(harness.ROOT, "oracle", "_ref", "tp_ref")
pytestmark = pytest.mark.skipif(not os.access(TP_REF, os.X_OK), reason="reference-built driver not present")

KLIST = ("MemcpyHtoD,0x00007f2a00000000,65536\nncclCommInitAll\nkernel-1.traceg\nncclGroupStart\nncclAllReduce\n"
         "ncclGroupEnd\nkernel-2.traceg\nMemcpyDtoH,0x1,2\n\nkernel-3_0.traceg\nkernel-4.traceg\nncclCommDestroy\n")


def _build_shim(lib_path, out):
    libdir, libname = os.path.split(lib_path)
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-I" + os.path.join(harness.ROOT, "shim"), "-o", out,
                    os.path.join(harness.ROOT, "oracle", "tp_driver.cc"),
                    os.path.join(harness.ROOT, "shim", "trace_parser_b200.cc"),
                    "-L" + libdir, "-l:" + libname, "-Wl,-rpath," + libdir], check=True)
    return out


def _run_both(lib_path, tmp):
    for i, name in enumerate(["edge_v3", "soup_sm70", "backprop_slice", "edge_v2_prefix"]):
        dst = "kernel-%d%s.traceg" % (i + 1, "_0" if i == 2 else "")
        shutil.copy(os.path.join(parity_cases.GOLDEN, name + ".traceg"), os.path.join(tmp, dst))
    open(os.path.join(tmp, "kernelslist.g"), "w").write(KLIST)
    shim = _build_shim(lib_path, os.path.join(tmp, "tp_shim"))
    a = subprocess.run([TP_REF, "./kernelslist.g"], cwd=tmp, capture_output=True)
    b = subprocess.run([shim, "./kernelslist.g"], cwd=tmp, capture_output=True)
    assert a.returncode == 0 and b.returncode == 0, (a.returncode, b.returncode, b.stderr[-500:])
    assert len(a.stdout) > 10000
    assert a.stdout == b.stdout


def test_shim_matches_reference_emulated(tmp_path):
    _run_both(harness.emu_library(), str(tmp_path))


@pytest.mark.gpu
def test_shim_matches_reference_cuda(tmp_path):
    _run_both(harness.CUDA_LIB, str(tmp_path))


def _windowed(lib_path, tmp):
    """A kernel file walked in many small windows (TI_SHIM_WINDOW_BYTES) gives the reference's output byte for byte."""
    harness.generate(os.path.join(tmp, "kernel-1.traceg"), profile=5, seed=61, n_tb=40, insts_per_warp=9, block_threads=96)
    open(os.path.join(tmp, "kernelslist.g"), "w").write("kernel-1.traceg\n")
    shim = _build_shim(lib_path, os.path.join(tmp, "tp_shim"))
    a = subprocess.run([TP_REF, "./kernelslist.g"], cwd=tmp, capture_output=True)
    for window in ("20000", "7000"):
        b = subprocess.run([shim, "./kernelslist.g"], cwd=tmp, capture_output=True, env=dict(os.environ, TI_SHIM_WINDOW_BYTES=window))
        assert a.returncode == 0 and b.returncode == 0, (a.returncode, b.returncode, b.stderr[-500:])
        assert a.stdout == b.stdout, window


def test_shim_windowed_emulated(tmp_path):
    _windowed(harness.emu_library(), str(tmp_path))


@pytest.mark.gpu
def test_shim_windowed_cuda(tmp_path):
    _windowed(harness.CUDA_LIB, str(tmp_path))


def test_shim_unopenable_file_exits_1(tmp_path):
    """trace_parser.cc:224-227,306-309: message on stdout + exit(1)."""
    shim = _build_shim(harness.emu_library(), os.path.join(str(tmp_path), "tp_shim"))
    for exe in (TP_REF, shim):
        r = subprocess.run([exe, "/nonexistent/kernelslist.g"], capture_output=True)
        assert r.returncode == 1 and b"Unable to open file" in r.stdout
