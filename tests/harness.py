"""Test/bench harness: synthetic corpus generation (tools/libtigen.so), running the
CPU oracles, and building the SIMT-emulated library used by the CPU-only tests."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "tools", "libtigen.so")
REF_ORACLE = os.path.join(ROOT, "oracle", "_ref", "ref_oracle")
PORT_ORACLE = os.path.join(ROOT, "oracle", "port_oracle")
# TI_LIB: tuning builds of the same CUDA library (bench sweeps only)
CUDA_LIB = os.environ.get("TI_LIB") or os.path.join(ROOT, "accel_sim_framework_distributed_b200", "csrc", "libtraceingest.so")
EMU_LIB = os.path.join(ROOT, "tests", "simt_emu", "libtraceingest_emu.so")


class GenParams(C.Structure):
    _fields_ = [
        ("seed", C.c_uint64), ("profile", C.c_uint32), ("grid", C.c_uint32 * 3), ("block", C.c_uint32 * 3),
        ("insts_per_warp", C.c_uint32), ("binary_version", C.c_uint32), ("kernel_id", C.c_uint32),
        ("trace_version", C.c_uint32), ("n_threads", C.c_uint32), ("variant", C.c_uint32),
        ("shmem_base", C.c_uint64), ("local_base", C.c_uint64), ("raw", C.c_uint32), ("pad_", C.c_uint32),
    ]


def _gen_lib():
    src = os.path.join(ROOT, "tools", "tigen.cc")
    if not os.path.exists(GEN) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(GEN)):
        subprocess.run(["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-o", GEN,
                        os.path.join(ROOT, "tools", "tigen.cc"), "-lpthread"], check=True)
    lib = C.CDLL(GEN)
    lib.tigen_generate.restype = C.c_size_t
    lib.tigen_generate.argtypes = [C.POINTER(GenParams), C.c_void_p, C.c_size_t]
    return lib


def generate_bytes(profile, seed, n_tb, insts_per_warp, binary_version=70, block_threads=256, trace_version=3,
                   variant=0, kernel_id=1, grid=None, block=None, size_hint=None, raw=0,
                   shmem_base=0x00007f8c00000000, local_base=0x00007f8c10000000):
    """Returns the whole .traceg image as a numpy uint8 array."""
    lib = _gen_lib()
    p = GenParams()
    p.seed, p.profile, p.insts_per_warp = seed, profile, insts_per_warp
    g = grid or (n_tb, 1, 1)
    b = block or (block_threads, 1, 1)
    for i in range(3):
        p.grid[i], p.block[i] = g[i], b[i]
    p.binary_version, p.kernel_id, p.trace_version, p.variant = binary_version, kernel_id, trace_version, variant
    p.shmem_base, p.local_base = shmem_base, local_base
    p.raw = raw
    cap = size_hint or 0
    buf = np.empty(max(cap, 1), dtype=np.uint8)
    n = lib.tigen_generate(C.byref(p), buf.ctypes.data if cap else None, cap)
    if n > cap:
        buf = np.empty(n, dtype=np.uint8)
        n = lib.tigen_generate(C.byref(p), buf.ctypes.data, n)
    return buf[:n]


def generate(path, **kw):
    img = generate_bytes(**kw)
    img.tofile(path)
    return len(img)


def oracle_kind():
    """'reference' when the reference-built oracle travelled with the repo, else 'port'."""
    return "reference" if os.access(REF_ORACLE, os.X_OK) else "port"


def oracle_dump(trace_path, tmpdir, config=None, kind=None, name="oracle.tidump"):
    """Runs a CPU oracle on a .traceg file; returns the dump path.  Raises on failure."""
    kind = kind or oracle_kind()
    exe = REF_ORACLE if kind == "reference" else PORT_ORACLE
    if kind == "port" and not os.path.exists(exe):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "port"], check=True)
    out = os.path.join(tmpdir, name)
    cmd = [exe, "dump", trace_path, out]
    if config is not None:
        cfg = os.path.join(tmpdir, "cfg.bin")
        np.array([config]).tofile(cfg)
        cmd.append(cfg)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode not in (0, 5):
        raise RuntimeError("oracle failed (%d): %s %s" % (r.returncode, r.stdout[-400:], r.stderr[-400:]))
    return out


def oracle_flags(trace_path, tmpdir, config=None):
    """error_flags the port oracle raises for a file (the reference itself would assert)."""
    import json
    cmd = [PORT_ORACLE, "dump", trace_path, os.path.join(tmpdir, "flags.tidump")]
    if config is not None:
        cfg = os.path.join(tmpdir, "cfg.bin")
        np.array([config]).tofile(cfg)
        cmd.append(cfg)
    r = subprocess.run(cmd, capture_output=True, text=True)
    return json.loads(r.stdout.strip().splitlines()[-1])["error_flags"]


def emu_library():
    """Builds (if stale) the SIMT-emulated build of the SAME kernel source with g++.
    Test infrastructure only: lets the CPU-only suite exercise the kernel logic."""
    csrc = os.path.join(ROOT, "accel_sim_framework_distributed_b200", "csrc")
    emu = os.path.join(ROOT, "tests", "simt_emu")
    srcs = [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cu", ".cuh", ".cpp", ".h", ".inc"))]
    srcs += [os.path.join(emu, f) for f in ("simt_emu.h", "simt_emu.cc", "cuda_runtime_emu.h")]
    srcs.append(os.path.join(ROOT, "include", "traceingest.h"))
    if not os.path.exists(EMU_LIB) or any(os.path.getmtime(s) > os.path.getmtime(EMU_LIB) for s in srcs):
        subprocess.run(["g++", "-std=c++17", "-O2", "-g", "-DTI_EMU", "-I" + emu, "-shared", "-fPIC", "-o", EMU_LIB,
                        "-x", "c++", os.path.join(csrc, "ti_api.cu"), os.path.join(csrc, "ti_host.cpp"),
                        os.path.join(emu, "simt_emu.cc")], check=True)
    return EMU_LIB
