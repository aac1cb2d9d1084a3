#!/usr/bin/env python3
"""Device-resident timing of ti_regroup_device (writer side) beside the reference tool.
usage (GPU box): python tools/regroup_bench.py [n_tb] [reps]"""
import os
import subprocess
import sys
import tempfile
import time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import json
import numpy as np
import torch
import harness
import rawtrace
from accel_sim_framework_distributed_b200 import trace_parser

n_tb = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
raw = np.asarray(harness.generate_bytes(profile=2, seed=2, n_tb=n_tb, insts_per_warp=25, raw=1))
tp = trace_parser(lib_path=harness.CUDA_LIB)
dev = torch.empty(len(raw) + 16, dtype=torch.uint8, device="cuda")
dev[:len(raw)].copy_(torch.from_numpy(raw))
out = torch.empty(len(raw) + (64 << 20), dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
ms = []
for i in range(reps + 2):
    n, st = tp.regroup_device(dev.data_ptr(), len(raw), out.data_ptr(), out.numel())
    if i >= 2:
        ms.append(st.device_ms)
best, med = min(ms), sorted(ms)[len(ms) // 2]
res = {"workload": "config2 as a raw .trace: %d thread blocks x 8 warps x 25 lines, warps interleaved" % n_tb, "in_bytes": len(raw), "out_bytes": n,
       "lines": int(st.n_lines), "ms_median": med, "ms_best": best, "in_GB_per_s": len(raw) / med / 1e6,
       "alg_bytes_GB_per_s": (len(raw) + n) / med / 1e6}
# the reference tool on one host core (it is single-threaded), on a bounded sample
sample_tb = min(n_tb, 6000)
sample = harness.generate_bytes(profile=2, seed=2, n_tb=sample_tb, insts_per_warp=25, raw=1).tobytes()
if os.access(rawtrace.PTP_REF, os.X_OK):
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.perf_counter()
        g, _, _ = rawtrace.ptp_reference(sample, tmp)
        dt = time.perf_counter() - t0
    res["cpu_reference"] = {"GB_per_s": len(sample) / dt / 1e9, "cores": 1, "sample": "%d thread blocks (%.0f MB), file in and out" % (sample_tb, len(sample) / 1e6)}
    got = bytes(out[:n].cpu().numpy()) if n_tb == sample_tb else None
    if got is not None:
        res["matches_reference"] = got == g
print(json.dumps(res))
