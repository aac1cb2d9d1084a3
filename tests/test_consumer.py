"""The reference's own consumer code on top of the drop-in (towards SURVEY 8 f2).

oracle/_ref/ref_oracle is the reference's trace_parser.cc + trace_driven.cc +
abstract_hardware_model.cc under oracle/ref_driver.cc.  Here the SAME driver and the SAME
unmodified trace_driven.cc / abstract_hardware_model.cc are compiled against shim/trace_parser.h
and linked with shim/trace_parser_b200.cc + the C ABI library instead of the reference's parser:
thread blocks decoded by the kernels, consumed by the reference's parse_from_trace_struct and
generate_mem_accesses.  Its dumps must be byte-identical to the all-reference build's."""
import filecmp
import os
import subprocess

import pytest

import harness
import parity_cases

REF = os.path.join(harness.ROOT, "oracle", "_ref")
OBJS = [os.path.join(REF, f) for f in ("shim_ref_driver.o", "shim_trace_driven.o", "abstract_hardware_model.o", "shim_trace_parser.o")]
SHIM_CUDA = os.path.join(REF, "ref_oracle_shim_cuda")
NAMES = ["edge_v3", "edge_v2_prefix", "edge_ampere", "edge_many_headers", "soup_sm35", "soup_sm61", "soup_sm70", "soup_sm75", "soup_sm86",
         "stream_slice", "divergent_slice", "backprop_slice", "edge_no_bases", "edge_no_trailing_newline"]   # default (QV100) configuration


def _link_emulated(tmp):
    lib = harness.emu_library()
    libdir, libname = os.path.split(lib)
    exe = os.path.join(tmp, "ref_oracle_shim_emu")
    subprocess.run(["g++", "-o", exe] + OBJS + ["-Wl,--gc-sections", "-L" + libdir, "-l:" + libname, "-Wl,-rpath," + libdir], check=True)
    return exe


def _same_dumps(exe, tmp, traces):
    for i, tr in enumerate(traces):
        out = os.path.join(tmp, "shim%d.tidump" % i)
        r = subprocess.run([exe, "dump", tr, out], capture_output=True, text=True)
        assert r.returncode == 0, (tr, r.stdout[-300:], r.stderr[-300:])
        want = tr[:-7] + ".tidump"
        if not os.path.exists(want):                       # generated corpus: the all-reference build, run now
            want = os.path.join(tmp, "ref%d.tidump" % i)
            subprocess.run([harness.REF_ORACLE, "dump", tr, want], check=True, capture_output=True)
        assert filecmp.cmp(out, want, shallow=False), tr


@pytest.mark.skipif(not all(os.path.exists(o) for o in OBJS), reason="oracle/_ref objects not built (needs /root/reference)")
def test_reference_consumer_on_emulated_kernels(tmp_path):
    tmp = str(tmp_path)
    traces = [os.path.join(parity_cases.GOLDEN, n + ".traceg") for n in NAMES]
    gen = os.path.join(tmp, "kernel-7.traceg")
    harness.generate(gen, profile=5, seed=707, n_tb=30, insts_per_warp=11, block_threads=160, binary_version=75)
    _same_dumps(_link_emulated(tmp), tmp, traces + [gen])


@pytest.mark.gpu
@pytest.mark.skipif(not os.access(SHIM_CUDA, os.X_OK), reason="oracle/_ref/ref_oracle_shim_cuda not built (needs /root/reference at build time)")
def test_reference_consumer_on_cuda_kernels(tmp_path):
    tmp = str(tmp_path)
    traces = [os.path.join(parity_cases.GOLDEN, n + ".traceg") for n in NAMES]
    gen = os.path.join(tmp, "kernel-7.traceg")
    harness.generate(gen, profile=2, seed=708, n_tb=400, insts_per_warp=25)
    _same_dumps(SHIM_CUDA, tmp, traces + [gen])
