"""TEST INFRASTRUCTURE for the writer side: raw kernel-N.trace images and the
oracles of the .trace -> .traceg regrouping.

  ungroup()        builds a raw trace (what the NVBit tool writes, tracer_tool.cu:685-756: every
                   instruction line prefixed "tbx tby tbz warp ") from a grouped .traceg image, the
                   warps interleaved the way concurrently running warps would be
  ptp_reference()  runs the reference's own post-traces-processing.cpp (oracle/_ref/ptp_ref)
  ptp_port()       restatement of group_per_block in Python, each step citing the reference
                   (post-traces-processing.cpp, "PTP"); pinned against ptp_reference by
                   tests/test_regroup.py and against tests/golden/regroup_*.traceg
"""
import math
import os
import random
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
PTP_REF = os.path.join(ROOT, "oracle", "_ref", "ptp_ref")


def ungroup(traceg, seed, window=3, drop_tb=(), blank_every=0):
    """traceg: bytes of a trace_version >= 3 .traceg.  Thread blocks are issued in waves of `window`;
    inside a wave the next line comes from a random warp that still has lines (per-warp order kept)."""
    rng = random.Random(seed)
    text = traceg.decode()
    head, _, body = text.partition("\n#BEGIN_TB")
    header = [ln for ln in head.split("\n")]
    while header and header[-1] == "":
        header.pop()
    blocks = []
    for blk in ("\n#BEGIN_TB" + body).split("\n#BEGIN_TB")[1:]:
        m = re.search(r"thread block = (\d+),(\d+),(\d+)", blk)
        xyz = tuple(int(v) for v in m.groups())
        warps = []
        for wm in re.finditer(r"warp = (\d+)\ninsts = (\d+)\n((?:[^\n#w][^\n]*\n)*)", blk):
            lines = [ln for ln in wm.group(3).split("\n") if ln]
            assert len(lines) == int(wm.group(2))
            warps.append((int(wm.group(1)), lines))
        blocks.append((xyz, warps))
    out = list(header) + [""]          # the header already ends with its "#traces format = ..." line
    n = 0
    for w0 in range(0, len(blocks), window):
        live = []
        for bi, (xyz, warps) in enumerate(blocks[w0:w0 + window]):
            if w0 + bi in drop_tb:
                continue
            for wid, lines in warps:
                if lines:
                    live.append([xyz, wid, lines, 0])
        while live:
            k = rng.randrange(len(live))
            xyz, wid, lines, at = live[k]
            out.append("%d %d %d %d %s" % (xyz[0], xyz[1], xyz[2], wid, lines[at]))
            n += 1
            if blank_every and n % blank_every == 0:
                out.append("")
            live[k][3] += 1
            if live[k][3] == len(lines):
                live.pop(k)
    return ("\n".join(out) + "\n").encode()


def ptp_reference(raw, tmpdir, commands=("kernel-1.trace",)):
    """Runs the reference tool on a kernelslist holding `commands`; kernel-1.trace = raw.
    Returns (kernel-1.traceg bytes, kernelslist.g bytes, stdout)."""
    assert os.access(PTP_REF, os.X_OK), "build oracle/_ref first (make -C oracle ref)"
    d = os.path.join(str(tmpdir), "ptp")
    os.makedirs(d, exist_ok=True)
    open(os.path.join(d, "kernel-1.trace"), "wb").write(raw)
    open(os.path.join(d, "kernelslist"), "w").write("\n".join(commands) + "\n")
    for f in ("kernel-1.traceg", "kernelslist.g"):
        if os.path.exists(os.path.join(d, f)):
            os.remove(os.path.join(d, f))
    r = subprocess.run([PTP_REF, os.path.join(d, "kernelslist")], capture_output=True, text=True)
    g = open(os.path.join(d, "kernel-1.traceg"), "rb").read() if os.path.exists(os.path.join(d, "kernel-1.traceg")) else None
    lg = open(os.path.join(d, "kernelslist.g"), "rb").read()
    return g, lg, r.stdout


def ptp_port(raw):
    """group_per_block (PTP:85-194) restated.  Returns the .traceg bytes."""
    out = []
    lines = raw.decode("latin-1").split("\n")          # while (!eof) getline: newlines + 1 pieces (PTP:103-105)
    grid = block = None
    insts = None
    for line in lines:
        if len(line) == 0 or line[0] == "#":           # PTP:107-110
            out.append(line + "\n")
            continue
        if line[0] == "-":                             # PTP:112-137
            m = re.match(r"-grid dim = \((-?\d+),(-?\d+),(-?\d+)\)", line)
            if m and line.split()[:2] == ["-grid", "dim"]:
                grid = [int(v) for v in m.groups()]
            m = re.match(r"-block dim = \((-?\d+),(-?\d+),(-?\d+)\)", line)
            if m and line.split()[:2] == ["-block", "dim"]:
                block = [int(v) for v in m.groups()]
            if grid and block and insts is None:
                w = int(math.ceil(float(block[0] * block[1] * block[2]) / 32))
                insts = [dict(init=False, xyz=(0, 0, 0), warps=[[] for _ in range(w)]) for _ in range(grid[0] * grid[1] * grid[2])]
            out.append(line + "\n")
            continue
        m = re.match(r"\s*(\d+)\s+(\d+)\s+(\d+)\s+(\d+)", line)   # ss >> x >> y >> z >> warp (PTP:139-141)
        x, y, z, w = (int(v) for v in m.groups())
        tb = z * grid[1] * grid[0] + y * grid[0] + x              # PTP:142-143
        t = insts[tb]
        if not t["init"]:                                         # PTP:144-149
            t["init"], t["xyz"] = True, (x, y, z)
        t["warps"][w].append(line[m.end() + 1:])                  # substr(tellg() + 1), PTP:154-156
    for t in insts or []:                                         # PTP:159-192
        if not (t["init"] and t["warps"]):
            continue
        out.append("\n#BEGIN_TB\n")
        out.append("\nthread block = %d,%d,%d\n" % t["xyz"])
        for j, wl in enumerate(t["warps"]):
            out.append("\nwarp = %d\ninsts = %d\n" % (j, len(wl)))
            out.extend(ln + "\n" for ln in wl)
        out.append("\n#END_TB\n")
    return "".join(out).encode("latin-1")
