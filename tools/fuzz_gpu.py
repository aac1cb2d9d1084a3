#!/usr/bin/env python3
"""Extended differential run on the GPU box: many seeds x profiles x architectures x coalescer configurations against
the CPU oracle (tests/parity_cases.run_generated), beyond what the test suite runs every time.
usage: python tools/fuzz_gpu.py [n_seeds] [thread-block multiplier]"""
import os
import sys
import tempfile
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import harness
import parity_cases
from accel_sim_framework_distributed_b200 import trace_parser

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30
mul = int(sys.argv[2]) if len(sys.argv) > 2 else 1
tp = trace_parser(lib_path=harness.CUDA_LIB)
overs = [None, dict(coalesce_arch=20), dict(mem_warp_parts=4, shmem_limited_broadcast=1, num_shmem_bank=16, clock_gated_lanes=0),
         dict(coalesce_arch=35, gmem_skip_L1D=1)]
done = 0
with tempfile.TemporaryDirectory() as tmp:
    for seed in range(n):
        for profile, kw in ((5, dict(n_tb=mul * (8 + seed % 7), insts_per_warp=6 + seed % 9, block_threads=32 * (1 + seed % 5), binary_version=(35, 61, 70, 75, 86)[seed % 5])),
                            (3, dict(n_tb=mul * (2 + seed % 3), insts_per_warp=8, block_threads=64)),
                            (2, dict(n_tb=mul * (10 + seed % 13), insts_per_warp=25)),
                            (1, dict(n_tb=3, insts_per_warp=20 + seed % 15, grid=(1, 3, 1), block=(16, 16, 1))),
                            (4, dict(n_tb=mul * 4, insts_per_warp=12, block_threads=128, binary_version=86))):
            parity_cases.run_generated(tp, tmp, over=overs[(seed + profile) % len(overs)], profile=profile, seed=1000 + 17 * seed + profile, **kw)
            done += 1
print("fuzz ok: %d corpora, oracle=%s" % (done, harness.oracle_kind()))
