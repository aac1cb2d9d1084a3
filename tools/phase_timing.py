#!/usr/bin/env python3
"""Tuning aid: per-phase cycle sums of thread 0 of every CTA (build with -DTI_PHASE_TIMING).
usage (GPU box): TI_LIB=build/libti_phase.so python tools/phase_timing.py [workload]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import torch
import bench
import harness
from accel_sim_framework_distributed_b200 import trace_parser

wl = sys.argv[1] if len(sys.argv) > 1 else "config2"
img = bench.make_corpus(wl, 2)
tp = trace_parser(lib_path=harness.CUDA_LIB)
hdr = tp.parse_header(img[:8192])
body = img[hdr.body_offset:]
dev = torch.empty(len(body) + 16, dtype=torch.uint8, device="cuda")
dev[:len(body)].copy_(torch.from_numpy(body))
torch.cuda.synchronize()
k = tp.open_device(dev.data_ptr(), len(body), hdr, keepalive=dev)
k.upload()
for _ in range(2):
    k.decode(); k.sync()
out = (C.c_ulonglong * 16)()
tp.lib.ti_debug_phase_cycles(out)
k.decode(); k.sync()
ms = k.decode_ms()
tp.lib.ti_debug_phase_cycles(out)
names = ["top of loop", "P0 wait text", "P1 classify", "W line starts", "W tokenize + stride + counters", "F tail (recompute, barrier)",
         "H structural lines", "M warp path", "T resolve chains 0/1", "headers + records out", "loop tail", "(in S + F') S + look-back of F", "(in S + F') F rows", "(in S + F') F offsets, patch, transactions", "-"]
tot = sum(out[:15])
print("decode %.3f ms; thread-0 cycles per phase (share of the CTA's time):" % ms)
for n, v in zip(names, out):
    print("  %-40s %6.2f%%" % (n, 100.0 * v / tot))
