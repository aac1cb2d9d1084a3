"""C-ABI surface (CPU only): the library loads, exports every symbol the header
declares, its host-only entry points work, and compute fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import edge_cases
import harness
from accel_sim_framework_distributed_b200 import _ffi

pytestmark = pytest.mark.skipif(not os.path.exists(harness.CUDA_LIB), reason="run __graft_entry__.build() first")


@pytest.fixture(scope="module")
def lib():
    return _ffi.load_library(harness.CUDA_LIB)


def test_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(harness.ROOT, "include", "traceingest.h")).read()
    declared = set(re.findall(r"\b(ti_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_ffi.SYMBOLS), declared ^ set(_ffi.SYMBOLS)
    for s in declared:
        assert hasattr(lib, s), s


def test_header_parse(lib):
    img = edge_cases.VALID["edge_v3"][0]
    h = _ffi.KernelHdr()
    assert lib.ti_parse_kernel_header(img, len(img), C.byref(h)) == 0
    assert h.kernel_name == b"_Z4edgePv" and h.kernel_id == 7 and tuple(h.grid_dim) == (2, 1, 1)
    assert tuple(h.tb_dim) == (64, 1, 1) and h.binary_version == 70 and h.trace_version == 3
    assert h.shmem_base_addr == edge_cases.SHMEM and h.local_base_addr == edge_cases.LOCAL
    assert h.nvbit_version == b" 1.5.5"  # the reference keeps the blank after '=' (trace_parser.cc:354)
    assert img[h.body_offset - 1:h.body_offset] == b"\n" and b"#traces format" in img[:h.body_offset]


def test_commandlist_and_memcpy(lib):
    text = (b"MemcpyHtoD,0x00007f0000001000,4096\nncclCommInitAll\nkernel-1.traceg\nncclGroupStart\nncclAllReduce\n"
            b"ncclGroupEnd\nMemcpyDtoH,0x1,2\n\nkernel-2_3.traceg\nncclCommDestroy\n")
    n = C.c_size_t()
    cmds = (_ffi.Command * 16)()
    assert lib.ti_parse_commandlist(text, len(text), cmds, 16, C.byref(n)) == 0
    kinds = [c.kind for c in cmds[:n.value]]
    assert kinds == [2, 4, 1, 6, 8, 7, 1, 5]  # MemcpyDtoH is skipped (trace_parser.cc:279)
    a, cnt = C.c_uint64(), C.c_uint64()
    line = text[:text.index(b"\n")]
    assert lib.ti_parse_memcpy(line, len(line), C.byref(a), C.byref(cnt)) == 0
    assert (a.value, cnt.value) == (0x7F0000001000, 4096)
    assert lib.ti_parse_memcpy(b"MemcpyHtoD,1", 12, C.byref(a), C.byref(cnt)) == _ffi.TI_ERR_PARSE


def test_plan_shards(lib):
    img = np.frombuffer(harness.generate_bytes(profile=2, seed=5, n_tb=40, insts_per_warp=6).tobytes(), dtype=np.uint8)
    cuts = (C.c_size_t * 5)()
    assert lib.ti_plan_shards(img.ctypes.data, len(img), 4, cuts) == 0
    cuts = list(cuts)
    assert cuts[0] == 0 and cuts[-1] == len(img) and cuts == sorted(cuts)
    for c in cuts[1:-1]:
        assert bytes(img[c:c + 9]) == b"#BEGIN_TB" and img[c - 1] == 10


def test_no_cpu_fallback(lib):
    """Without a CUDA device every compute entry point refuses to run."""
    if lib.ti_device_count() > 0:
        pytest.skip("a GPU is present")
    from accel_sim_framework_distributed_b200 import records
    cfg = np.array([records.qv100_config()])
    ctx = C.c_void_p()
    assert lib.ti_ctx_create(cfg.ctypes.data, C.byref(ctx)) == 8  # TI_ERR_NO_DEVICE
    assert not ctx.value


def test_counters_finalize_matches_reference_style_sums(lib):
    """Host finalisation of the integer block vs the doubles the oracle accumulates
    instruction by instruction like the reference (1e-9 relative)."""
    import tiformat as T
    import parity_cases
    for name in ("soup_sm70", "edge_subwarps_broadcast", "backprop_slice"):
        o = T.load_dump(os.path.join(parity_cases.GOLDEN, name + ".tidump"))
        over = edge_cases.VALID.get(name, (None, {}))[1]
        cfg = np.array([T.qv100_config(**over)])
        pw = np.zeros(1, dtype=T.POWER)
        cnt = np.array([o.counters])
        # ti_counters_finalize needs a context (GPU); the same arithmetic is reachable through the emulated build
        emu = _ffi.load_library(harness.emu_library())
        ctx = C.c_void_p()
        assert emu.ti_ctx_create(cfg.ctypes.data, C.byref(ctx)) == 0
        assert emu.ti_counters_finalize(ctx, cnt.ctypes.data, pw.ctypes.data) == 0
        emu.ti_ctx_destroy(ctx)
        assert not T.compare_power(pw[0], o.power, rtol=1e-9)
