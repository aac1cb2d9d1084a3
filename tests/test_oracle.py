"""The oracle itself (CPU only): the restatement (oracle/port_oracle) against
(a) the reference-produced golden fixtures and (b) the reference-built oracle run here
when /root/reference is present.  Parity of the port is PINNED by these tests."""
import itertools
import os
import subprocess

import pytest

import edge_cases
import harness
import parity_cases
import tiformat as T


@pytest.fixture(scope="module", autouse=True)
def _built():
    subprocess.run(["make", "-s", "-C", os.path.join(harness.ROOT, "oracle"), "port"], check=True)


@pytest.mark.parametrize("name", parity_cases.golden_names())
def test_port_matches_reference_golden(name, tmp_path):
    over = edge_cases.VALID.get(name, (None, {}))[1]
    cfg = T.qv100_config(**over) if over else None
    out = harness.oracle_dump(os.path.join(parity_cases.GOLDEN, name + ".traceg"), str(tmp_path), config=cfg, kind="port")
    a, b = T.load_dump(os.path.join(parity_cases.GOLDEN, name + ".tidump")), T.load_dump(out)
    msgs = T.compare(a, b, offsets=False, names=("ref", "port"))
    msgs += T.compare_counters(a.counters, b.counters) + T.compare_power(a.power, b.power, rtol=0)
    assert not msgs, msgs[:8]


CFGS = {
    "qv100": {},
    "fermi": dict(coalesce_arch=20),
    "fermi_skip_l1": dict(coalesce_arch=20, gmem_skip_L1D=1),
    "parts2_bcast": dict(mem_warp_parts=2, shmem_limited_broadcast=1, num_shmem_bank=16, clock_gated_lanes=0, num_dp_units=0),
    "arch13_parts4": dict(coalesce_arch=13, mem_warp_parts=4, clock_gated_lanes=0),
}


@pytest.mark.skipif(harness.oracle_kind() != "reference", reason="reference-built oracle not present")
@pytest.mark.parametrize("cfg,bv", list(itertools.product(CFGS, [35, 60, 70, 75, 86])))
def test_port_matches_reference_fuzz(cfg, bv, tmp_path):
    path = os.path.join(str(tmp_path), "k.traceg")
    harness.generate(path, profile=5, seed=1000 + bv, n_tb=24, insts_per_warp=10, binary_version=bv, block_threads=128)
    c = T.qv100_config(**CFGS[cfg])
    a = T.load_dump(harness.oracle_dump(path, str(tmp_path), config=c, kind="reference", name="ref.tidump"))
    b = T.load_dump(harness.oracle_dump(path, str(tmp_path), config=c, kind="port", name="port.tidump"))
    msgs = T.compare(a, b, offsets=False, names=("ref", "port"))
    msgs += T.compare_counters(a.counters, b.counters) + T.compare_power(a.power, b.power, rtol=0)
    assert not msgs, msgs[:8]


@pytest.mark.parametrize("name", sorted(edge_cases.ERRORS))
def test_port_flags_malformed_lines(name, tmp_path):
    from accel_sim_framework_distributed_b200 import _ffi
    img, flag = edge_cases.ERRORS[name]
    path = os.path.join(str(tmp_path), name + ".traceg")
    open(path, "wb").write(img)
    assert harness.oracle_flags(path, str(tmp_path)) & _ffi.EF[flag]


def test_readme_example_line(tmp_path):
    """The one instruction line the reference documents (util/tracer_nvbit/README.md:89)."""
    img = edge_cases.image([(0, 0, 0, [["0000 ffffffff 1 R1 IMAD.MOV.U32 2 R255 R255 0"]])])
    path = os.path.join(str(tmp_path), "r.traceg")
    open(path, "wb").write(img)
    d = T.load_dump(harness.oracle_dump(path, str(tmp_path), kind="port"))
    r = d.inst[0]
    assert (r["pc"], r["mask"], r["dst"], list(r["src"][:2]), r["pack_b"] & 7, (r["pack_b"] >> 3) & 1) == (0, 0xFFFFFFFF, 1, [255, 255], 2, 1)
    assert d.opcodes[0] == b"IMAD.MOV.U32" and r["mem_idx"] == 0xFFFFFFFF
