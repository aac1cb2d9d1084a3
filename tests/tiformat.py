"""Oracle dump container (oracle/dump_format.h) loader and field-by-field comparison
helpers.  Shared by the tests, bench.py and smoke()."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from accel_sim_framework_distributed_b200.records import (  # noqa: E402,F401
    CONFIG, COUNTERS, INST, POWER, POWER_FIELDS, TB, TXN, WARP, qv100_config)


class Dump:
    pass


def load_dump(path):
    raw = np.fromfile(path, dtype=np.uint8)
    assert bytes(raw[:7]) == b"TIDUMP2", path
    n_inst, n_mem, n_txn, n_tb, n_warp, n_blob = (int(v) for v in raw[8:56].view("<u8"))
    off = 56
    d = Dump()

    def take(dtype, n):
        nonlocal off
        nbytes = np.dtype(dtype).itemsize * n
        a = raw[off:off + nbytes].view(dtype)
        off += nbytes
        return a

    d.inst = take(INST, n_inst)
    d.rows = take("<u8", n_mem * 32).reshape(n_mem, 32)
    d.txn_off = take("<u4", n_mem + 1)
    if (n_mem + 1) & 1:
        off += 4
    d.txn = take(TXN, n_txn)
    d.tb = take(TB, n_tb)
    d.warp = take(WARP, n_warp)
    d.counters = take(COUNTERS, 1)[0]
    d.power = take(POWER, 1)[0]
    blob = bytes(raw[off:off + n_blob])
    d.opcodes = blob.split(b"\0")[:-1] if n_blob else []
    return d


def opcodes_from_text(inst, text):
    """Opcode strings of every record, cut out of the shard text."""
    return [bytes(text[int(o):int(o) + int(n)]) for o, n in zip(inst["opc_off"], inst["opcode_len"])]


def compare(a, b, text_a=None, text_b=None, offsets=True, names=("a", "b")):
    """Field-by-field comparison of two result sets; returns a list of messages
    (empty == bit-exact).  `offsets=False` ignores the text offsets a
    reference-built dump cannot know and compares opcode strings instead."""
    msgs = []

    def chk(label, x, y):
        if x.shape != y.shape:
            msgs.append("%s: shape %s vs %s" % (label, x.shape, y.shape))
            return
        if not np.array_equal(x, y):
            bad = np.flatnonzero((x != y).reshape(len(x), -1).any(axis=1)) if x.ndim else [0]
            msgs.append("%s: %d mismatches, first at %d: %s=%s %s=%s" % (
                label, len(bad), bad[0], names[0], x[bad[0]], names[1], y[bad[0]]))

    for f in INST.names:
        if f == "opc_off" and not offsets:
            continue
        if len(a.inst) != len(b.inst):
            msgs.append("inst count %d vs %d" % (len(a.inst), len(b.inst)))
            break
        chk("inst." + f, a.inst[f], b.inst[f])
    chk("rows", a.rows, b.rows)
    chk("txn_off", a.txn_off, b.txn_off)
    for f in TXN.names:
        if len(a.txn) != len(b.txn):
            msgs.append("txn count %d vs %d" % (len(a.txn), len(b.txn)))
            break
        chk("txn." + f, a.txn[f], b.txn[f])
    for f in TB.names:
        if f in ("text_off", "line_len") and not offsets:
            continue
        if len(a.tb) != len(b.tb):
            msgs.append("tb count %d vs %d" % (len(a.tb), len(b.tb)))
            break
        chk("tb." + f, a.tb[f], b.tb[f])
    for f in ("warp_id", "n_insts", "first_inst"):
        if len(a.warp) != len(b.warp):
            msgs.append("warp count %d vs %d" % (len(a.warp), len(b.warp)))
            break
        chk("warp." + f, a.warp[f], b.warp[f])
    if not offsets:
        oa = a.opcodes if text_a is None else opcodes_from_text(a.inst, text_a)
        ob = b.opcodes if text_b is None else opcodes_from_text(b.inst, text_b)
        if oa != ob:
            k = next((i for i, (x, y) in enumerate(zip(oa, ob)) if x != y), min(len(oa), len(ob)))
            msgs.append("opcode strings differ first at %d" % k)
    return msgs


def compare_counters(ca, cb, skip=("n_begin_tb", "n_end_tb", "text_bytes", "n_warp", "n_mem", "reserved")):
    msgs = []
    for f in COUNTERS.names:
        if f in skip:
            continue
        if not np.array_equal(ca[f], cb[f]):
            msgs.append("counters.%s differ: %s vs %s" % (f, ca[f], cb[f]))
    return msgs


def compare_power(pa, pb, rtol=1e-9):
    """AccelWattch floating-point sums: north_star tolerance 1e-9 relative."""
    msgs = []
    for f in POWER_FIELDS:
        x, y = float(pa[f]), float(pb[f])
        if abs(x - y) > rtol * max(abs(x), abs(y)):
            msgs.append("power.%s: %r vs %r" % (f, x, y))
    return msgs


