"""Error behaviour of the C ABI (include/traceingest.h): status codes instead of the reference's
exit() / assert(), and a context that stays usable after an error."""
import ctypes as C
import os

import numpy as np
import pytest

import edge_cases
import harness
from accel_sim_framework_distributed_b200 import _ffi
from accel_sim_framework_distributed_b200 import records as R
from accel_sim_framework_distributed_b200 import trace_parser

ST = {v: k for k, v in _ffi.STATUS_NAMES.items()}


def check_contract(tp, tmp):
    lib = tp.lib
    # unopenable trace file: the reference prints "Unable to open file" and exits (trace_parser.cc:306-309)
    with pytest.raises(_ffi.TraceIngestError) as e:
        tp.parse_kernel_info(os.path.join(tmp, "nope.traceg"))
    assert e.value.status == ST["TI_ERR_IO"] and "Unable to open file" in str(e.value)
    # unsupported binary version: trace_driven.cc:113-118 prints and exits
    img = edge_cases.image([(0, 0, 0, [edge_cases.W0[:3]])], bv=100)
    p = os.path.join(tmp, "sm100.traceg")
    open(p, "wb").write(img)
    with pytest.raises(_ffi.TraceIngestError) as e:
        tp.parse_kernel_info(p)
    assert e.value.status == ST["TI_ERR_BINARY_VERSION"]
    # call order
    good = os.path.join(tmp, "good.traceg")
    open(good, "wb").write(edge_cases.VALID["edge_v3"][0])
    k = tp.parse_kernel_info(good)
    assert lib.ti_kernel_decode(k._h) == ST["TI_ERR_STATE"]                       # decode before upload
    assert lib.ti_kernel_sync(k._h, None) == ST["TI_ERR_STATE"]                   # sync before decode
    view = _ffi.View()
    assert lib.ti_kernel_fetch(k._h, C.byref(view)) == ST["TI_ERR_STATE"]         # fetch before sync
    assert lib.ti_kernel_upload(k._h, 10, 5) == ST["TI_ERR_INVALID_ARG"]          # begin > end
    assert lib.ti_kernel_upload(k._h, 0, k.body_len + 1) == ST["TI_ERR_INVALID_ARG"]
    lm = _ffi.LocalMap()
    off = np.zeros(2, dtype=np.uint32)
    n = C.c_size_t()
    issue = np.zeros(1, dtype=R.LOCAL_ISSUE)
    assert lib.ti_kernel_coalesce_local(k._h, C.byref(lm), issue.ctypes.data, 1, None, 0, off.ctypes.data, C.byref(n)) == ST["TI_ERR_STATE"]
    d = k.run()                                                                   # and the object still works
    assert len(d.inst) == 66
    assert lib.ti_kernel_coalesce_local(k._h, C.byref(lm), issue.ctypes.data, 1, None, 0, off.ctypes.data, C.byref(n)) == ST["TI_ERR_INVALID_ARG"]  # zeroed map
    tp.kernel_finalizer(k)
    # null / misaligned arguments
    h = C.c_void_p()
    hdr = _ffi.KernelHdr()
    assert lib.ti_kernel_open_buffer(tp.ctx, None, 10, C.byref(hdr), C.byref(h)) == ST["TI_ERR_INVALID_ARG"]
    assert lib.ti_kernel_open_device(tp.ctx, C.c_void_p(8), 100, C.byref(hdr), C.byref(h)) in (ST["TI_ERR_INVALID_ARG"], ST["TI_ERR_BINARY_VERSION"])
    out_len = C.c_size_t()
    assert lib.ti_regroup_device(tp.ctx, C.c_void_p(8), 100, None, 0, C.byref(out_len), None) == ST["TI_ERR_INVALID_ARG"]   # not 16-byte aligned
    assert lib.ti_regroup_buckets(None, None, None, None, None) == ST["TI_ERR_INVALID_ARG"]
    assert lib.ti_strerror(ST["TI_ERR_PARSE"]) and lib.ti_strerror(12345)
    # a configuration the reference itself aborts on (abstract_hardware_model.cc:436)
    cfg = np.array([R.qv100_config()], dtype=R.CONFIG)
    cfg["coalesce_arch"] = 12
    ctx = C.c_void_p()
    assert lib.ti_ctx_create(cfg.ctypes.data, C.byref(ctx)) == ST["TI_ERR_INVALID_ARG"]
    cfg = np.array([R.qv100_config()], dtype=R.CONFIG)
    cfg["mem_warp_parts"] = 3
    assert lib.ti_ctx_create(cfg.ctypes.data, C.byref(ctx)) == ST["TI_ERR_INVALID_ARG"]
    # parse error: status, flags and the offset of the first bad line; then a good file again
    bad_img, flag = edge_cases.ERRORS["unknown_opcode"]
    pb = os.path.join(tmp, "bad.traceg")
    open(pb, "wb").write(bad_img)
    kb = tp.parse_kernel_info(pb)
    kb.upload(); kb.decode()
    counts = kb.sync(check=False)
    assert counts.error_flags & _ffi.EF[flag]
    body = bad_img[kb.hdr.body_offset:]
    assert body[counts.first_error_line_off:].startswith(b"0010 ffffffff 1 R1 FROBNICATE")
    assert lib.ti_kernel_fetch(kb._h, C.byref(view)) == ST["TI_ERR_STATE"]        # nothing to fetch after a failed sync
    tp.kernel_finalizer(kb)
    k2 = tp.parse_kernel_info(good)
    assert len(k2.run().inst) == 66
    tp.kernel_finalizer(k2)


def test_contract_emulated(emu_parser, tmp_path):
    check_contract(emu_parser, str(tmp_path))


@pytest.mark.gpu
def test_contract_cuda(cuda_parser, tmp_path):
    check_contract(cuda_parser, str(tmp_path))


@pytest.mark.gpu
def test_bad_device_ordinal_is_no_device():
    cfg = R.qv100_config()
    with pytest.raises(_ffi.TraceIngestError) as e:
        trace_parser(config=cfg, lib_path=harness.CUDA_LIB, device=99)
    assert e.value.status == ST["TI_ERR_NO_DEVICE"]
