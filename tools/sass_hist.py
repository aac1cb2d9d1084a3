#!/usr/bin/env python3
"""Opcode histogram of the shipped cubin's ti_decode_kernel (cuobjdump -sass): proves the Blackwell bulk-copy / mbarrier
instructions are there and counts the spill sites.  usage: tools/sass_hist.py [libtraceingest.so]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "accel_sim_framework_distributed_b200", "csrc", "libtraceingest.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur, per = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_]*(?:\.[A-Z0-9_]+)*)", line)
    if cur and m:
        per[cur][m.group(1)] += 1
for fn, cnt in per.items():
    if "ti_decode_kernel" not in fn:
        continue
    n = sum(cnt.values())
    print("%s: %d SASS instructions (%d KB)" % (fn, n, n * 16 // 1024))
    base = collections.Counter()
    for k, v in cnt.items():
        base[k.split(".")[0]] += v
    print("  by mnemonic: " + ", ".join("%s %d" % kv for kv in base.most_common(24)))
    for key in ("UBLKCP", "SYNCS", "LDL", "STL", "MATCH", "REDUX", "VOTE", "SHFL", "BAR", "ATOMS", "RED", "LDS", "STS", "LDG", "STG", "LD.", "ST."):
        hits = {k: v for k, v in cnt.items() if k.startswith(key.rstrip("."))} if not key.endswith(".") else {k: v for k, v in cnt.items() if k.split(".")[0] == key[:-1]}
        if hits:
            print("  %-7s %s" % (key.rstrip("."), ", ".join("%s x%d" % kv for kv in sorted(hits.items()))))
cold = {fn: sum(c.values()) for fn, c in per.items() if "ti_decode_kernel" not in fn and ("decode_addresses" in fn or "coalesce" in fn or "shared_cycles" in fn)}
for fn, n in cold.items():
    print("out-of-line (cold) %s: %d instructions" % (fn, n))
