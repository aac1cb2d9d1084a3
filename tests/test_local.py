"""Local-memory translation + coalescing: ti_kernel_coalesce_local against the
reference's own generate_mem_accesses fed by translate_local_memaddr (oracle/_ref/ref_oracle
`local` mode; the translation is restated there because shader.cc cannot be linked) and
against the committed reference-produced fixtures tests/golden/*.tilocal."""
import os
import subprocess

import numpy as np
import pytest

import harness
import parity_cases
import tiformat as T
from accel_sim_framework_distributed_b200 import _ffi
from accel_sim_framework_distributed_b200 import records as R

GOLDEN = parity_cases.GOLDEN
# (num_shader, n_thread_per_shader, local_mem_map, padded_threads_per_cta, max_cta_per_shader)
MAPS = {"qv100_cta_map": (80, 2048, 1, 128, 8), "legacy_map": (80, 2048, 0, 128, 8), "tiny": (3, 256, 1, 96, 2)}
FIXTURES = [("soup_sm70", "qv100_cta_map"), ("soup_sm35", "legacy_map"), ("soup_sm86", "tiny"), ("edge_v3", "qv100_cta_map")]


def load_tilocal(path):
    b = open(path, "rb").read()
    assert b[:8] == b"TILOCAL1"
    n, nt = (int(v) for v in np.frombuffer(b[8:24], dtype="<u8"))
    iss = np.frombuffer(b[24:24 + 16 * n], dtype="<u4").reshape(-1, 4)
    txn = np.frombuffer(b[24 + 16 * n:24 + 16 * n + 32 * nt], dtype=T.TXN)
    return iss, txn


def local_map(m):
    lm = _ffi.LocalMap()
    lm.num_shader, lm.n_thread_per_shader, lm.local_mem_map, lm.padded_threads_per_cta, lm.max_cta_per_shader = m
    lm.local_generic_start = _ffi.TI_LOCAL_GENERIC_START
    return lm


def run_oracle(trace, m, tmp):
    out = os.path.join(str(tmp), "o.tilocal")
    r = subprocess.run([harness.REF_ORACLE, "local", trace, out] + [str(v) for v in m], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def check(tp, trace, tilocal, m):
    iss, want = load_tilocal(tilocal)
    assert len(iss) > 0
    k = tp.parse_kernel_info(trace)
    d = k.run()
    # the oracle issued every local load / store of the file: the decode must agree on which those are
    space = d.inst["pack_c"] & 0xf
    is_local = (space == 2) & ((d.inst["op"] == 8) | (d.inst["op"] == 11))   # local_space; LOAD_OP / STORE_OP (abstract_hardware_model.h enums)
    assert np.array_equal(np.nonzero(is_local)[0], iss[:, 0])
    ok = iss[:, 3] != 0xFFFFFFFF                                              # issues where no reference assert trips
    issues = np.zeros(int(ok.sum()), dtype=R.LOCAL_ISSUE)
    issues["inst"], issues["sid"], issues["tid0"] = iss[ok, 0], iss[ok, 1], iss[ok, 2]
    txn, off = k.coalesce_local(local_map(m), issues)
    assert np.array_equal(np.diff(off.astype(np.int64)), iss[ok, 3])
    assert len(txn) == len(want)
    for f in T.TXN.names:
        assert np.array_equal(txn[f], want[f]), f
    # an issue that trips an assert in the reference is refused
    for row in iss[~ok]:
        bad = np.zeros(1, dtype=R.LOCAL_ISSUE)
        bad["inst"], bad["sid"], bad["tid0"] = row[0], row[1], row[2]
        with pytest.raises(_ffi.TraceIngestError):
            k.coalesce_local(local_map(m), bad)
    # so is one that does not name a local load / store
    not_local = np.zeros(1, dtype=R.LOCAL_ISSUE)
    not_local["inst"] = int(np.nonzero(~is_local)[0][0])
    with pytest.raises(_ffi.TraceIngestError):
        k.coalesce_local(local_map(m), not_local)
    tp.kernel_finalizer(k)


def fixture_paths(name, mapname):
    return os.path.join(GOLDEN, name + ".traceg"), os.path.join(GOLDEN, "%s.%s.tilocal" % (name, mapname))


@pytest.mark.parametrize("name,mapname", FIXTURES)
def test_emu_local_golden(emu_parser, name, mapname):
    trace, fx = fixture_paths(name, mapname)
    check(emu_parser, trace, fx, MAPS[mapname])


@pytest.mark.skipif(harness.oracle_kind() != "reference", reason="needs oracle/_ref/ref_oracle")
def test_emu_local_generated(emu_parser, tmp_path):
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, profile=5, seed=123, n_tb=20, insts_per_warp=12, block_threads=128)
    check(emu_parser, path, run_oracle(path, MAPS["qv100_cta_map"], tmp_path), MAPS["qv100_cta_map"])


@pytest.mark.gpu
@pytest.mark.parametrize("name,mapname", FIXTURES)
def test_gpu_local_golden(cuda_parser, name, mapname):
    trace, fx = fixture_paths(name, mapname)
    check(cuda_parser, trace, fx, MAPS[mapname])


@pytest.mark.gpu
@pytest.mark.skipif(harness.oracle_kind() != "reference", reason="needs oracle/_ref/ref_oracle")
@pytest.mark.parametrize("mapname", sorted(MAPS))
def test_gpu_local_generated(cuda_parser, tmp_path, mapname):
    path = os.path.join(str(tmp_path), "kernel-1.traceg")
    harness.generate(path, profile=5, seed=321, n_tb=300, insts_per_warp=14, block_threads=96 if mapname == "tiny" else 128, binary_version=75)
    check(cuda_parser, path, run_oracle(path, MAPS[mapname], tmp_path), MAPS[mapname])
