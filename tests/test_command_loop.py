"""The caller of the trace parser: Accel-Sim's command loop.

oracle/_ref/loop_ref is the REFERENCE's own gpu-simulator/main.cc, compiled where it lies over stand-ins for the timing
model (oracle/loop_stubs: a deterministic "GPU" that pulls every thread block of a launched kernel through the trace
parser) and the reference's own trace_parser.cc.  The restated loop (shim/command_loop.h) is built over the same
stand-ins and the C++ drop-in (shim/ + the C ABI).  Their complete stdout -- parser echo lines, launch / finish order,
stream bookkeeping, NCCL commands, statistics lines, the engine's per-kernel instruction counts and checksums -- must
be byte-identical, for a serial and a concurrent window."""
import os
import shutil
import subprocess

import pytest

import harness
import parity_cases

LOOP_REF = os.path.join(harness.ROOT, "oracle", "_ref", "loop_ref")
pytestmark = pytest.mark.skipif(not os.access(LOOP_REF, os.X_OK), reason="reference-built command loop not present")

# serial window: commands of every kind between the kernels
KLIST_SERIAL = ("MemcpyHtoD,0x00007f2a00000000,65536\nncclCommInitAll\nkernel-1.traceg\nkernel-2.traceg\nncclGroupStart\nncclAllReduce\n"
                "ncclGroupEnd\nkernel-3.traceg\nMemcpyHtoD,0x10,32\nkernel-4.traceg\nkernel-5.traceg\nncclCommDestroy\n")
# concurrent window: the reference fills its window with whatever FOLLOWS a kernel command (main.cc:101-107 does not look at
# the command type) and reads commandlist[i] past the end while the window drains (main.cc:84), so a list it can run has its
# kernels in one run and enough commands behind them to drain on
KLIST_WINDOW = ("MemcpyHtoD,0x00007f2a00000000,65536\nncclCommInitAll\nncclGroupStart\nncclAllReduce\nncclGroupEnd\nkernel-1.traceg\nkernel-2.traceg\n"
                "kernel-3.traceg\nkernel-4.traceg\nkernel-5.traceg\n" + "ncclGroupStart\nncclAllReduce\nncclGroupEnd\n" * 3 + "ncclCommDestroy\n")


def _build(lib_path, out):
    libdir, libname = os.path.split(lib_path)
    o = os.path.join(harness.ROOT, "oracle")
    subprocess.run(["g++", "-std=c++17", "-O2", "-w", "-include", "bits/stdc++.h", "-I" + os.path.join(o, "loop_stubs"), "-I" + os.path.join(harness.ROOT, "shim"),
                    '-DTI_LOOP_TRACE_PARSER_H="%s"' % os.path.join(harness.ROOT, "shim", "trace_parser.h"), "-o", out,
                    os.path.join(o, "loop_driver.cc"), os.path.join(harness.ROOT, "shim", "trace_parser_b200.cc"),
                    "-L" + libdir, "-l:" + libname, "-Wl,-rpath," + libdir], check=True)
    return out


def _run(lib_path, tmp):
    # five kernels; the first two share CUDA stream 0: the second one waits for the first
    for i, (stream, n_tb) in enumerate([(0, 5), (0, 3), (7, 4), (9, 6), (11, 2)]):
        p = os.path.join(tmp, "kernel-%d.traceg" % (i + 1))
        harness.generate(p, profile=5, seed=100 + i, n_tb=n_tb, insts_per_warp=7 + i, block_threads=64, kernel_id=i + 1)
        txt = open(p, "rb").read().replace(b"-cuda stream id = 0", b"-cuda stream id = %d" % stream, 1)
        open(p, "wb").write(txt)
    mine = _build(lib_path, os.path.join(tmp, "loop_b200"))
    for concurrent, window, max_running in ((0, 1, 1), (1, 2, 3)):
        open(os.path.join(tmp, "kernelslist.g"), "w").write(KLIST_WINDOW if concurrent else KLIST_SERIAL)
        env = dict(os.environ, TI_LOOP_KERNELSLIST="./kernelslist.g", TI_LOOP_CONCURRENT=str(concurrent), TI_LOOP_WINDOW=str(window),
                   TI_LOOP_MAX_RUNNING=str(max_running))
        a = subprocess.run([LOOP_REF], cwd=tmp, capture_output=True, env=env)
        b = subprocess.run([mine], cwd=tmp, capture_output=True, env=env)
        assert a.returncode == 0 and b.returncode == 0, (a.returncode, b.returncode, a.stderr[-300:], b.stderr[-300:])
        assert a.stdout.count(b"launching kernel name") == 5 and a.stdout.count(b"ENGINE stats") == 5
        assert a.stdout == b.stdout, (concurrent, window)


def test_command_loop_matches_reference_emulated(tmp_path):
    _run(harness.emu_library(), str(tmp_path))


@pytest.mark.gpu
def test_command_loop_matches_reference_cuda(tmp_path):
    _run(harness.CUDA_LIB, str(tmp_path))
