#!/usr/bin/env python3
"""Headline benchmark of the trace-ingest path (BASELINE.json metric: trace GB/s and
warp-instructions/s decoded + coalesced, % of HBM peak).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload config2|config3|config4|config5] [--split kernel|byte-range]

A step = one pass of the fused decode kernel over one synthetic .traceg corpus
(config 2 of BASELINE.json: 100k thread blocks x 8 warps x 25 instructions, ~1 GB of
text, 60/20/20 ALU/LDG/STG, base+stride addresses).  `value` is measured with the text
already resident in HBM; `e2e` goes through the host-buffer C-ABI calls (pinned host
text -> device, decode, every result array back to pinned host memory).  With N > 1
(torchrun) every rank decodes its own corpus (the path shards by kernel file with no
exchange: weak scaling) and the integer counter blocks are summed with one NCCL
all-reduce, timed separately.  --impl reference times the reference's own CPU code
(oracle/_ref, else the oracle port) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

# tuning experiments with deliberately wrong output (tools/sweep_*.sh) set TI_BENCH_NOCHECK=1; never set for a reported number
_CHECK = os.environ.get("TI_BENCH_NOCHECK") != "1"

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "trace_text_GB_per_s_decoded_and_coalesced"

WORKLOADS = {
    # BASELINE.json configs[1] / configs[2]
    "config2": dict(profile=2, n_tb=100000, insts_per_warp=25, block_threads=256, size_hint=1200 << 20,
                    desc="synthetic ~1 GB traceg: 100k TBs x 8 warps x 25 insts, 60/20/20 ALU/LDG/STG, base+stride"),
    "config3": dict(profile=3, n_tb=8000, insts_per_warp=12, block_threads=256, size_hint=400 << 20,
                    desc="synthetic memory-divergent traceg: 50% list-compressed 32-lane accesses, 32 sectors/warp"),
    # BASELINE.json configs[3]: the kernel launches of examples/all-reduce (a ring all-reduce over 8 devices: 2*(8-1) ring
    # steps, main.cu:32-34,76-85), one kernel file each, sharded over the ranks by kernel file (strong scaling)
    "config4": dict(profile=4, n_tb=6000, insts_per_warp=24, block_threads=256, size_hint=100 << 20, n_kernels=14,
                    desc="14 ring-all-reduce step kernels (examples/all-reduce shape), ~65 MB of traceg each, sharded by kernel file"),
    # BASELINE.json configs[4]: a corpus far larger than one file, streamed from pinned host memory
    "config5": dict(profile=2, n_tb=100000, insts_per_warp=25, block_threads=256, size_hint=1200 << 20, files_per_gpu=8,
                    desc="64 GB-class corpus: 8 files of ~1 GB per GPU streamed from pinned host memory (one generated file decoded 8 "
                         "times per step: replication factor 8), compact results back to pinned host memory, counters all-reduced"),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traffic_bytes(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum of one ti_decode_kernel launch, from the committed
    ncu --set full capture of this workload (never measured under the timed run); None if not captured."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    t = json.load(open(p)).get(workload)
    return None if not t else int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons, "samples": len(sm)}


def algorithmic_bytes(text_bytes, n_inst, n_mem, n_txn):
    """every byte the path must read or write once: the text, one 32 B record
    per instruction, one 256 B address row + one 4 B CSR offset per memory instruction,
    one 32 B record per transaction."""
    return text_bytes + 32 * n_inst + 260 * n_mem + 32 * n_txn


def make_corpus(workload, seed, **over):
    import harness
    w = dict(WORKLOADS[workload])
    for k in ("desc", "n_kernels", "files_per_gpu"):
        w.pop(k, None)
    w.update(over)
    return harness.generate_bytes(seed=seed, **w)


def pin_to_gpu_numa_node(index):
    """Run this rank's host threads on the cores next to its GPU (sysfs numa_node of the PCI device): the pinned
    buffers it allocates afterwards then come from that node's memory.  Returns a short description."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(index).pci_bus_id if hasattr(torch.cuda.get_device_properties(index), "pci_bus_id") else None
        dom = getattr(torch.cuda.get_device_properties(index), "pci_domain_id", 0)
        dev = getattr(torch.cuda.get_device_properties(index), "pci_device_id", 0)
        path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/numa_node" % (dom, bus, dev)
        node = int(open(path).read().strip())
        if node < 0:
            return "numa_node -1 (single node): affinity untouched"
        cpus = []
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        os.sched_setaffinity(0, cpus)
        return "numa node %d, %d cpus" % (node, len(cpus))
    except Exception as e:   # sysfs not there, no permission: report, do not fail the bench
        return "unavailable (%s)" % type(e).__name__


def host_link_all_ranks(torch, dist, world):
    """What the host link gives when EVERY rank copies at once (between two barriers): pinned <-> device, both
    directions busy.  Per-rank GB/s of each direction; min / median over the ranks."""
    nb = 256 << 20
    hp_up, hp_dn = torch.empty(nb, dtype=torch.uint8).pin_memory(), torch.empty(nb, dtype=torch.uint8).pin_memory()
    d_up, d_dn = torch.empty(nb, dtype=torch.uint8, device="cuda"), torch.empty(nb, dtype=torch.uint8, device="cuda")
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    res = {}
    for mode in ("h2d", "d2h", "both"):
        best = 0.0
        for _ in range(3):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for _r in range(4):
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s_up):
                        d_up.copy_(hp_up, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s_dn):
                        hp_dn.copy_(d_dn, non_blocking=True)
            torch.cuda.synchronize()
            best = max(best, 4 * nb / (time.perf_counter() - t0) / 1e9)
        t = torch.tensor([best], dtype=torch.float64, device="cuda")
        if world > 1:
            parts = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(parts, t)
            vals = sorted(float(x.item()) for x in parts)
        else:
            vals = [best]
        res[mode] = {"min": vals[0], "median": vals[len(vals) // 2]}
    return {"h2d_GB_per_s": res["h2d"], "d2h_GB_per_s": res["d2h"], "each_direction_when_both_GB_per_s": res["both"], "ranks_copying_at_once": world,
            "how": "4 x 256 MiB pinned <-> device copies per rank, all ranks between two barriers, best of 3, per-rank GB/s"}


def cpu_run(trace_paths, stages="coalesce"):
    """Runs the CPU oracle binary over several files at once (one process each)."""
    import harness
    kind = harness.oracle_kind()
    exe = harness.REF_ORACLE if kind == "reference" else harness.PORT_ORACLE
    t0 = time.perf_counter()
    procs = [subprocess.Popen([exe, "time", p, "", stages], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for p in trace_paths]
    outs = [p.communicate()[0] for p in procs]
    wall = time.perf_counter() - t0
    n_inst = sum(json.loads(o.strip().splitlines()[-1])["n_inst"] for o in outs)
    return wall, n_inst, kind


def split_sample(img, hdr_len, n_parts, target_bytes, tmp):
    """Writes n_parts kernel files (header + a whole-thread-block slice of the body)."""
    body = img[hdr_len:]
    key = np.frombuffer(b"\n#BEGIN_TB", dtype=np.uint8)
    paths, total, pos = [], 0, 0
    for i in range(n_parts):
        end = min(len(body), pos + target_bytes)
        if end < len(body):
            chunk = body[end:end + (1 << 20)].tobytes()
            j = chunk.find(key.tobytes())
            end = end + j + 1 if j >= 0 else len(body)
        p = os.path.join(tmp, "sample-%d.traceg" % i)
        with open(p, "wb") as f:
            f.write(img[:hdr_len].tobytes())
            f.write(body[pos:end].tobytes())
        total += end - pos
        paths.append(p)
        pos = end
        if pos >= len(body):
            break
    return paths, total


def bench_reference(args):
    """The reference's CPU implementation on this box's host cores (bounded sample)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import harness
    cores = os.cpu_count() or 1
    wl = args.workload if args.workload in ("config2", "config3") else ("config2" if args.workload == "config5" else args.workload)
    img = make_corpus(wl, seed=2)
    # the body starts at the first "#BEGIN_TB" (found here in Python: this process loads no product library)
    body_offset = img[:65536].tobytes().find(b"#BEGIN_TB")
    assert body_offset > 0
    if harness.oracle_kind() != "reference":
        print("bench.py --impl reference: oracle/_ref/ref_oracle is not here, timing the CPU restatement (oracle/port_oracle) instead", file=sys.stderr)
    # ~0.25 M instructions, a few seconds of reference CPU time per core and step
    per_core = min(12 << 20, (len(img) - body_offset) // cores)
    with tempfile.TemporaryDirectory() as tmp:
        paths, nbytes = split_sample(img, body_offset, cores, per_core, tmp)
        times = []
        for s in range(args.warmup + args.steps):
            wall, n_inst, kind = cpu_run(paths)
            if s >= args.warmup:
                times.append(wall)
    t = sum(times) / len(times)
    val = nbytes / t / 1e9
    line = {
        "metric": METRIC, "value": val, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/u64",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": args.workload + ": " + WORKLOADS[args.workload]["desc"]},
        "warp_insts_per_s": n_inst / t,
        "cpu_baseline": {"value": val, "unit": "GB/s", "cores": len(paths), "kind": kind,
                         "sample": "%d processes x %.0f MB whole-thread-block slices of the same corpus (%.1f MB total), parse+convert+coalesce, stdout discarded" % (len(paths), per_core / 1e6, nbytes / 1e6)},
        "e2e": {"value": val, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def writer_side(tp, torch):
    """Informational, outside the headline metric: ti_regroup_device (.trace -> .traceg) on a
    config2-shaped raw trace resident in HBM, CUDA events inside the library around all its kernels."""
    import harness
    raw = np.asarray(harness.generate_bytes(profile=2, seed=2, n_tb=20000, insts_per_warp=25, raw=1))
    dev = torch.empty(len(raw) + 16, dtype=torch.uint8, device="cuda")
    dev[:len(raw)].copy_(torch.from_numpy(raw))
    torch.cuda.synchronize()
    ms = []
    for i in range(5):
        n, st = tp.regroup_device(dev.data_ptr(), len(raw))
        if i >= 2:
            ms.append(st.device_ms)
    med = sorted(ms)[len(ms) // 2]
    pk, _ = peaks()
    return {"kernel_set": "rg_count, rg_keys, rs_hist/rs_scatter, scans, rg_headers, rg_lines", "in_bytes": len(raw), "out_bytes": int(n),
            "lines": int(st.n_lines), "ms_median": med, "in_GB_per_s": len(raw) / med / 1e6,
            "roofline": {"bound": "hbm", "achieved": (len(raw) + n) / med / 1e6, "peak": pk, "unit": "GB/s", "frac": (len(raw) + n) / med / 1e6 / pk,
                         "algorithmic_bytes": "input read once + output written once"}}


def shim_block(img, window_mb=256):
    """SURVEY 8(d) "the shim's materialisation reported separately": one driver (oracle/tp_driver.cc, `time` mode) walks
    every thread block of the corpus through trace_parser::get_next_threadblock_traces, once built against the reference's
    own trace_parser.cc (oracle/_ref/tp_ref) and once against the drop-in over the CUDA library (oracle/_ref/tp_shim_cuda).
    Wall seconds, instructions per second and peak resident set of both; single host thread each."""
    import harness
    ref = os.path.join(harness.ROOT, "oracle", "_ref", "tp_ref")
    shim = os.path.join(harness.ROOT, "oracle", "_ref", "tp_shim_cuda")
    if not (os.access(ref, os.X_OK) and os.access(shim, os.X_OK)):
        return {"unavailable": "oracle/_ref/tp_ref or tp_shim_cuda not built"}
    out = {"window_mb": window_mb, "how": "tp_driver time: parse_kernel_info + get_next_threadblock_traces over all thread blocks, parser stdout discarded"}
    with tempfile.TemporaryDirectory() as tmp:
        img.tofile(os.path.join(tmp, "kernel-1.traceg"))
        open(os.path.join(tmp, "kernelslist.g"), "w").write("kernel-1.traceg\n")
        for name, exe in (("drop_in_b200", shim), ("reference", ref)):
            r = subprocess.run([exe, "time", "./kernelslist.g"], cwd=tmp, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True,
                               env=dict(os.environ, TI_SHIM_WINDOW_MB=str(window_mb), TI_SHIM_TIMING="1"))
            try:
                lines = r.stderr.strip().splitlines()
                d = json.loads([l for l in lines if l.startswith("{")][-1])
                d["warp_insts_per_s"] = d["n_inst"] / d["seconds"]
                d["GB_per_s"] = len(img) / d["seconds"] / 1e9
                for l in lines:   # the drop-in's own split of its wall time (TI_SHIM_TIMING)
                    if l.startswith("trace_parser(b200): ti_ctx_create"):
                        w = l.replace(",", " ").split()
                        d["ctx_create_s"], d["window_decode_s"] = float(w[2]), float(w[6])
                        d["hand_out_s"] = round(d["seconds"] - d["ctx_create_s"] - d["window_decode_s"], 3)
                        d["note"] = "ctx_create_s = CUDA context + module load, once per process; window_decode_s = uploads, fused kernel, gather, downloads"
                out[name] = d
            except Exception:
                out[name] = {"failed": r.returncode, "stderr": r.stderr[-300:]}
    if "seconds" in out.get("drop_in_b200", {}) and "seconds" in out.get("reference", {}):
        out["speedup"] = out["reference"]["seconds"] / out["drop_in_b200"]["seconds"]
        out["same_checksum"] = out["reference"]["checksum"] == out["drop_in_b200"]["checksum"]
    return out


def bench_b200(args):
    import torch
    import harness
    from accel_sim_framework_distributed_b200 import distributed as D
    from accel_sim_framework_distributed_b200 import trace_parser

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    affinity = pin_to_gpu_numa_node(local)      # before any pinned allocation of this rank
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wl = args.workload
    split = args.split
    tp = trace_parser(lib_path=harness.CUDA_LIB, device=local)
    nccl = D.NcclCounters(tp)                   # the library's own collective (ti_allreduce_counters)

    # ---- what this rank decodes
    #   split kernel     : its own kernel file(s)  (config2/3/5: one file per rank, weak scaling; config4: its share of the
    #                      14 kernel files of ONE job, strong scaling)
    #   split byte-range : its byte range of ONE file that all ranks share (ti_plan_shards), strong scaling
    scaling = "weak"
    stitched_ok = None
    if wl == "config4":
        scaling = "strong"
        n_k = WORKLOADS[wl]["n_kernels"]
        mine = D.kernels_for_rank(n_k, rank, world)
        imgs = [make_corpus(wl, seed=40 + i, kernel_id=i + 1) for i in mine]
    elif split == "byte-range":
        scaling = "strong"
        imgs = [make_corpus(wl, seed=2)]
    else:
        imgs = [make_corpus(wl, seed=2 + rank)]
    hdrs = [tp.parse_header(im[:8192]) for im in imgs]
    bodies = [im[h.body_offset:] for im, h in zip(imgs, hdrs)]
    ranges = [(0, len(b)) for b in bodies]
    if split == "byte-range":
        cuts = tp.plan_shards(bodies[0], world)
        ranges = [D.shard_for_rank(cuts, rank)]
        assert ranges[0][0] % 16 == 0 or True

    # ---- device-resident arm: the text lives in HBM before the timed region
    kds, keep = [], []
    for body, hdr, (b, e) in zip(bodies, hdrs, ranges):
        part = body[b:e]
        dev_text = torch.empty(len(part) + 16, dtype=torch.uint8, device="cuda")
        dev_text[:len(part)].copy_(torch.from_numpy(part))
        kd = tp.open_device(dev_text.data_ptr(), len(part), hdr, keepalive=dev_text)
        kd.upload()
        kds.append(kd)
    torch.cuda.synchronize()
    my_bytes = sum(e - b for b, e in ranges)

    def one_step():
        ms, cs = 0.0, []
        for kd in kds:
            kd.decode()
            cs.append(kd.sync(check=_CHECK))
            ms += kd.decode_ms()              # CUDA events on the library's stream around the launch
        return ms, cs

    for _ in range(args.warmup):
        one_step()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    step_ms = []
    for _ in range(args.steps):
        ms, counts_list = one_step()
        step_ms.append(ms)
    barrier()
    sampler.stop_flag = True
    total_ms = sum(step_ms)
    n_inst = sum(c.n_inst for c in counts_list); n_mem = sum(c.n_mem for c in counts_list); n_txn = sum(c.n_txn for c in counts_list)

    # ---- counter all-reduce over NCCL through the C ABI (the path's only exchange), timed apart
    allreduce_ms = None
    cnt_local = np.zeros(1, dtype=np.dtype(kds[0].counters().dtype))
    for kd in kds:
        c1 = np.zeros(1, dtype=cnt_local.dtype); c1[0] = kd.counters()
        tp.lib.ti_counters_add(cnt_local.ctypes.data, c1.ctypes.data)
    nccl.allreduce(kds[0])                      # warm-up
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    total_first = nccl.allreduce(kds[0])        # device to device + 6 KB to the host
    allreduce_ms = (time.perf_counter() - t0) * 1e3
    total_cnt = D.allreduce_counters(cnt_local[0], device="cuda") if (world > 1 or len(kds) > 1) else total_first

    # ---- byte-range split: the ranks' shard-relative ordinals + exchanged bases must equal the one-piece decode
    if split == "byte-range":
        base = D.exchange_offsets(counts_list[0], device="cuda")
        d = kds[0].fetch()
        tb = d.tb.copy(); warp = d.warp.copy()
        tb["first_inst"] += np.uint32(base["inst"]); tb["first_warp"] += np.uint32(base["warp"]); warp["first_inst"] += np.uint32(base["inst"])
        gathered = [None] * world
        if world > 1:
            dist.all_gather_object(gathered, (tb["first_inst"].copy(), tb["first_warp"].copy(), warp["first_inst"].copy()))
        else:
            gathered = [(tb["first_inst"], tb["first_warp"], warp["first_inst"])]
        if rank == 0:
            whole = tp.open_buffer(bodies[0], hdrs[0])
            w = whole.run()
            stitched_ok = bool(np.array_equal(np.concatenate([g[0] for g in gathered]), w.tb["first_inst"]) and
                               np.array_equal(np.concatenate([g[1] for g in gathered]), w.tb["first_warp"]) and
                               np.array_equal(np.concatenate([g[2] for g in gathered]), w.warp["first_inst"]))
            tp.kernel_finalizer(whole)

    # ---- end-to-end arm: host buffers through the C ABI, copies inside the timed region.  The headline uses the
    # compact result form (ti_kernel_decode_streamed_compact: 16 B descriptors instead of base+stride rows and dense
    # transactions, expanded on the host on demand); the full form is timed beside it.
    reps = WORKLOADS[wl].get("files_per_gpu", 1)
    e2e_s, d2h, e2e_full_s, d2h_full = float("nan"), 0, float("nan"), 0
    kes = []
    if not args.no_e2e:
        pins = []
        for body, hdr, (b, e) in zip(bodies, hdrs, ranges):
            part = np.ascontiguousarray(body[b:e])
            pin = torch.from_numpy(part)
            torch.cuda.cudart().cudaHostRegister(pin.data_ptr(), pin.numel(), 0)
            pins.append(pin)
            kes.append(tp.open_buffer(part, hdr))
        for _ in range(max(1, min(args.warmup, 2))):
            for ke in kes:
                view = ke.run_streamed_compact(args.shard_bytes)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            for _rep in range(reps):
                d2h = 0
                for ke in kes:
                    view = ke.run_streamed_compact(args.shard_bytes)   # pinned host text -> device, fused kernel + gather, results -> pinned host
                    d2h += view.d2h_bytes
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / args.steps
        d2h *= reps
        if wl != "config5":
            full_steps = max(1, min(args.steps, 3))
            for ke in kes:
                vf = ke.run_streamed(args.shard_bytes)
            barrier()
            t0 = time.perf_counter()
            for _ in range(full_steps):
                d2h_full = 0
                for ke in kes:
                    vf = ke.run_streamed(args.shard_bytes)
                    d2h_full += 32 * len(vf.inst) + 256 * len(vf.rows) + 4 * len(vf.txn_off) + 32 * len(vf.txn) + 32 * len(vf.tb) + 16 * len(vf.warp)
            torch.cuda.synchronize()
            e2e_full_s = (time.perf_counter() - t0) / full_steps
        for pin in pins:
            torch.cuda.cudart().cudaHostUnregister(pin.data_ptr())

    # ---- what the host link can do on this box with every rank copying at once (the bound of the end-to-end arm)
    link = host_link_all_ranks(torch, dist, world) if not args.no_e2e else None

    # ---- max over ranks
    stats = torch.tensor([total_ms, e2e_s, e2e_full_s], dtype=torch.float64, device="cuda")
    sums = torch.tensor([float(my_bytes), float(n_inst), float(n_txn), float(n_mem), float(d2h), float(d2h_full)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums)
    total_ms, e2e_s, e2e_full_s = stats.tolist()
    all_bytes, all_inst, all_txn, all_mem, all_d2h, all_d2h_full = sums.tolist()

    if rank == 0:
        peak, peak_src = peaks()
        per_launch_ms = sum(step_ms) / len(step_ms)
        b_alg = algorithmic_bytes(my_bytes, n_inst, n_mem, n_txn)
        achieved = b_alg / (per_launch_ms * 1e-3) / 1e9
        value = all_bytes * args.steps / (total_ms * 1e-3) / 1e9
        if wl == "config5":      # the corpus does not sit in HBM: the streamed arm IS the workload
            value = all_bytes * reps / e2e_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": (e2e_s * 1e3 if wl == "config5" else total_ms / args.steps), "ms_best": min(step_ms), "ms_median": sorted(step_ms)[len(step_ms) // 2],
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "u8/u64", "data": "synthetic",
            "config": {"workload": wl + ": " + WORKLOADS[wl]["desc"], "split": split,
                       "text_bytes_rank0": my_bytes, "text_bytes_all_ranks": int(all_bytes) * (reps if wl == "config5" else 1),
                       "n_inst_rank0": int(n_inst), "n_mem_rank0": int(n_mem), "n_txn_rank0": int(n_txn), "kernel_files_rank0": len(kds),
                       "l2": "input (%.2f GB on rank 0) and outputs larger than the 126 MB L2" % (my_bytes / 1e9),
                       "sharding": "by %s, no data-path collective; counters all-reduced over NCCL (ti_allreduce_counters) after the timed region" % ("byte range of one file" if split == "byte-range" else "kernel file"),
                       "host_affinity": affinity},
            "warp_insts_per_s": all_inst * args.steps / (total_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8TBs": achieved / 8000.0, "traffic": traffic_bytes(wl), "traffic_unit": "dram bytes per launch (ncu, profiles/ncu_traffic.json)", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": b_alg, "kernel": "ti_decode_kernel", "launch_ms": per_launch_ms},
            "e2e": {"value": all_bytes * reps / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": int(all_bytes) * reps, "d2h_bytes_per_step": int(all_d2h),
                    "ms_per_step": e2e_s * 1e3, "warp_insts_per_s": all_inst * reps / e2e_s, "steps": args.steps, "shard_bytes": args.shard_bytes or (64 << 20),
                    "result_form": "compact (ti_kernel_decode_streamed_compact): records, 16 B descriptors, rows / transactions that are not closed forms",
                    "d2h_GB_per_s_achieved": all_d2h / e2e_s / 1e9,
                    "full_form": None if wl == "config5" else {"value": all_bytes / e2e_full_s / 1e9, "ms_per_step": e2e_full_s * 1e3, "d2h_bytes_per_step": int(all_d2h_full),
                                                               "d2h_GB_per_s_achieved": all_d2h_full / e2e_full_s / 1e9},
                    "host_link": link},
            "gpu_launches": 2 * args.steps * len(kds),   # ti_decode_kernel + ti_check_kernel per kernel file and step of the device-timed arm
            "clocks": sampler.summary(),
            "counter_allreduce_ms": allreduce_ms,
            "counter_allreduce": "ti_allreduce_counters: ncclAllReduce(uint64, sum) of the 5.5 KB block, device to device, + copy to the host",
            "counters_check": {"n_inst_all_ranks": int(total_cnt["n_inst"]), "hist_sum": int(total_cnt["hist"].sum()), "expected": int(all_inst)},
        }
        if split == "byte-range":
            line["stitched_ordinals_equal_one_piece"] = stitched_ok
        if world == 1 and not args.no_cpu_baseline and wl in ("config2", "config3"):
            with tempfile.TemporaryDirectory() as tmp:
                paths, nbytes = split_sample(imgs[0], hdrs[0].body_offset, 1, 64 << 20, tmp)
                wall, n_inst_cpu, kind = cpu_run(paths)
            line["cpu_baseline"] = {"value": nbytes / wall / 1e9, "unit": "GB/s", "cores": 1, "kind": kind,
                                    "warp_insts_per_s": n_inst_cpu / wall,
                                    "sample": "first %.0f MB (whole thread blocks) of the same corpus, 1 process, parse+convert+coalesce, stdout discarded" % (nbytes / 1e6)}
        if world == 1 and not args.no_extras and wl == "config2":
            line["writer_side"] = writer_side(tp, torch)
            line["shim"] = shim_block(imgs[0])
        print(json.dumps(line))
    nccl.close()
    for k in kds + kes:
        tp.kernel_finalizer(k)
    tp.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS))
    ap.add_argument("--split", default="kernel", choices=["kernel", "byte-range"], help="byte-range: all ranks share ONE kernel file, cut at #BEGIN_TB (strong scaling)")
    ap.add_argument("--shard-bytes", type=int, default=0, help="shard size of the host-to-host pipeline (0 = the library's default, 64 MiB)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the informational writer-side measurement")
    ap.add_argument("--no-e2e", action="store_true", help="tuning sweeps only: skip the host-buffer arm")
    args = ap.parse_args()
    if args.impl == "reference":
        bench_reference(args)
    else:
        bench_b200(args)


if __name__ == "__main__":
    main()
