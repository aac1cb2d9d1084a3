#!/usr/bin/env python3
"""Regenerates tests/golden/: small .traceg inputs + the outputs of the REFERENCE's own
code for them (oracle/_ref/ref_oracle, i.e. the reference's trace_parser.cc,
trace_driven.cc and abstract_hardware_model.cc compiled from /root/reference).

The reference ships no golden vectors for this path; these files are the
pin: they travel with the repo, so the oracle port and the CUDA kernels are checked
against reference-produced bytes even where /root/reference does not exist.

    python tests/golden/make_golden.py        (needs oracle/_ref/ref_oracle)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import edge_cases  # noqa: E402
import harness  # noqa: E402
import tiformat  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")


def main():
    assert os.access(REF, os.X_OK), "build oracle/_ref first (make -C oracle ref)"
    cases = {}
    for name, (img, over) in edge_cases.VALID.items():
        cases[name] = (img, over)
    # one small soup per architecture table + one config-2 / config-3 shaped slice
    for bv in (35, 61, 70, 75, 86):
        cases["soup_sm%d" % bv] = (bytes(harness.generate_bytes(profile=5, seed=100 + bv, n_tb=6, insts_per_warp=8,
                                                                binary_version=bv, block_threads=96)), {})
    cases["stream_slice"] = (bytes(harness.generate_bytes(profile=2, seed=2, n_tb=4, insts_per_warp=25)), {})
    cases["divergent_slice"] = (bytes(harness.generate_bytes(profile=3, seed=3, n_tb=2, insts_per_warp=10, block_threads=64)), {})
    cases["backprop_slice"] = (bytes(harness.generate_bytes(profile=1, seed=1, n_tb=3, insts_per_warp=34, grid=(1, 3, 1), block=(16, 16, 1))), {})
    for name, (img, over) in sorted(cases.items()):
        tr = os.path.join(HERE, name + ".traceg")
        open(tr, "wb").write(img)
        cmd = [REF, "dump", tr, os.path.join(HERE, name + ".tidump")]
        if over:
            cfg = os.path.join(HERE, name + ".cfg")
            import numpy as np
            np.array([tiformat.qv100_config(**over)]).tofile(cfg)
            cmd.append(cfg)
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, (name, r.stdout, r.stderr)
        print(name, len(img), r.stdout.strip())


def regroup_fixtures():
    """Raw kernel-N.trace inputs + what the REFERENCE's post-traces-processing.cpp (oracle/_ref/ptp_ref)
    writes for them: the pin of ti_regroup_* and of rawtrace.ptp_port."""
    import tempfile
    import rawtrace
    import test_regroup
    for name in ("soup", "grid_y", "empty_blocks", "blank_lines_inside", "no_trailing_newline", "header_only", "one_line"):
        raw = test_regroup.CASES[name]
        with tempfile.TemporaryDirectory() as tmp:
            g, _, out = rawtrace.ptp_reference(raw, tmp)
        open(os.path.join(HERE, "regroup_%s.trace" % name), "wb").write(raw)
        open(os.path.join(HERE, "regroup_%s.traceg" % name), "wb").write(g)
        print("regroup_" + name, len(raw), len(g), out.count("Warning"), "warnings")


def local_fixtures():
    """tests/golden/<trace>.<map>.tilocal: the reference's generate_mem_accesses on every local load / store
    after translate_local_memaddr with the placement rule of `ref_oracle local`."""
    import test_local
    for name, mapname in test_local.FIXTURES:
        trace, fx = test_local.fixture_paths(name, mapname)
        r = subprocess.run([REF, "local", trace, fx] + [str(v) for v in test_local.MAPS[mapname]], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        iss, txn = test_local.load_tilocal(fx)
        print(os.path.basename(fx), len(iss), "issues", len(txn), "transactions")


if __name__ == "__main__":
    regroup_fixtures()
    local_fixtures()
    main()
