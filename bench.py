#!/usr/bin/env python3
"""Headline benchmark of the trace-ingest path (BASELINE.json metric: trace GB/s and
warp-instructions/s decoded + coalesced, % of HBM peak).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload config2|config3]

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


def make_corpus(workload, seed):
    import harness
    w = dict(WORKLOADS[workload])
    w.pop("desc")
    return harness.generate_bytes(seed=seed, **w)


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
    from accel_sim_framework_distributed_b200 import _ffi
    cores = os.cpu_count() or 1
    img = make_corpus(args.workload, seed=2)
    lib = _ffi.load_library(harness.CUDA_LIB)
    hdr = _ffi.KernelHdr()
    lib.ti_parse_kernel_header(img.ctypes.data, min(len(img), 8192), C.byref(hdr))
    # ~0.25 M instructions, a few seconds of reference CPU time per core and step
    per_core = min(12 << 20, (len(img) - int(hdr.body_offset)) // cores)
    with tempfile.TemporaryDirectory() as tmp:
        paths, nbytes = split_sample(img, hdr.body_offset, cores, per_core, tmp)
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


def bench_b200(args):
    import torch
    import harness
    from accel_sim_framework_distributed_b200 import trace_parser

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # every rank decodes its own kernel file (seed = 2 + rank): sharding by kernel file, SURVEY 8e
    img = make_corpus(args.workload, seed=2 + rank)
    tp = trace_parser(lib_path=harness.CUDA_LIB, device=local)
    hdr = tp.parse_header(img[:8192])
    body = img[hdr.body_offset:]

    # ---- device-resident arm: the text lives in HBM before the timed region
    dev_text = torch.empty(len(body) + 16, dtype=torch.uint8, device="cuda")
    dev_text[:len(body)].copy_(torch.from_numpy(body))
    torch.cuda.synchronize()
    kd = tp.open_device(dev_text.data_ptr(), len(body), hdr, keepalive=dev_text)
    kd.upload()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        kd.decode(); counts = kd.sync(check=_CHECK)
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    step_ms = []
    for _ in range(args.steps):
        kd.decode()
        counts = kd.sync(check=_CHECK)
        step_ms.append(kd.decode_ms())   # CUDA events on the library's stream around the launch
    barrier()
    sampler.stop_flag = True
    total_ms = sum(step_ms)
    cnt = kd.counters()

    # ---- NCCL all-reduce of the integer counter block (the path's only exchange), timed apart
    allreduce_ms = None
    if world > 1:
        from accel_sim_framework_distributed_b200 import distributed as D
        D.allreduce_counters(cnt, device="cuda")  # warm-up
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); total_cnt = D.allreduce_counters(cnt, device="cuda"); e1.record()
        torch.cuda.synchronize()
        allreduce_ms = e0.elapsed_time(e1)
    else:
        total_cnt = cnt

    # ---- end-to-end arm: host buffers through the C ABI, copies inside the timed region.  The headline uses the
    # compact result form (ti_kernel_decode_streamed_compact: 16 B descriptors instead of base+stride rows and dense
    # transactions, expanded on the host on demand); the full form is timed beside it.
    ke = tp.open_buffer(body, hdr)
    e2e_s, d2h, e2e_full_s, d2h_full = float("nan"), 0, float("nan"), 0
    if not args.no_e2e:
        pin = torch.from_numpy(body)
        torch.cuda.cudart().cudaHostRegister(pin.data_ptr(), pin.numel(), 0)
        for _ in range(max(1, min(args.warmup, 2))):
            view = ke.run_streamed_compact()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            view = ke.run_streamed_compact()   # pinned host text -> device, fused kernel + gather, every result array -> pinned host
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / args.steps
        d2h = view.d2h_bytes
        assert view.counts["n_inst"] == counts.n_inst and view.counts["n_txn"] == counts.n_txn
        full_steps = max(1, min(args.steps, 3))
        vf = ke.run_streamed()
        barrier()
        t0 = time.perf_counter()
        for _ in range(full_steps):
            vf = ke.run_streamed()
        torch.cuda.synchronize()
        e2e_full_s = (time.perf_counter() - t0) / full_steps
        d2h_full = 32 * len(vf.inst) + 256 * len(vf.rows) + 4 * len(vf.txn_off) + 32 * len(vf.txn) + 32 * len(vf.tb) + 16 * len(vf.warp)
        torch.cuda.cudart().cudaHostUnregister(pin.data_ptr())

    # ---- what the host link can do on this box (the bound of the end-to-end arm): pinned copies, both directions at once
    link = None
    if not args.no_e2e and rank == 0:
        nb = 1 << 30
        hp_up, hp_dn = torch.empty(nb, dtype=torch.uint8).pin_memory(), torch.empty(nb, dtype=torch.uint8).pin_memory()
        d_up, d_dn = torch.empty(nb, dtype=torch.uint8, device="cuda"), torch.empty(nb, dtype=torch.uint8, device="cuda")
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
        best = {}
        for mode in ("h2d", "d2h", "both"):
            ts = []
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                if mode in ("h2d", "both"):
                    with torch.cuda.stream(s_up):
                        d_up.copy_(hp_up, non_blocking=True)
                if mode in ("d2h", "both"):
                    with torch.cuda.stream(s_dn):
                        hp_dn.copy_(d_dn, non_blocking=True)
                torch.cuda.synchronize()
                ts.append(time.perf_counter() - t0)
            best[mode] = nb / min(ts) / 1e9
        link = {"h2d_GB_per_s": best["h2d"], "d2h_GB_per_s": best["d2h"], "each_direction_when_both_GB_per_s": best["both"],
                "how": "1 GiB pinned <-> device copies, best of 3, wall clock around a synchronize"}
        del hp_up, hp_dn, d_up, d_dn

    # ---- max over ranks
    stats = torch.tensor([total_ms, e2e_s, e2e_full_s], dtype=torch.float64, device="cuda")
    sums = torch.tensor([float(len(body)), float(counts.n_inst), float(counts.n_txn), float(counts.n_mem)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums)
    total_ms, e2e_s, e2e_full_s = stats.tolist()
    all_bytes, all_inst, all_txn, all_mem = sums.tolist()

    if rank == 0:
        peak, peak_src = peaks()
        per_launch_ms = sum(step_ms) / len(step_ms)
        b_alg = algorithmic_bytes(len(body), counts.n_inst, counts.n_mem, counts.n_txn)
        achieved = b_alg / (per_launch_ms * 1e-3) / 1e9
        value = all_bytes * args.steps / (total_ms * 1e-3) / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "ms_best": min(step_ms), "ms_median": sorted(step_ms)[len(step_ms) // 2],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8/u64", "data": "synthetic",
            "config": {"workload": args.workload + ": " + WORKLOADS[args.workload]["desc"],
                       "text_bytes_per_gpu": len(body), "n_inst_per_gpu": int(counts.n_inst), "n_mem_per_gpu": int(counts.n_mem),
                       "n_txn_per_gpu": int(counts.n_txn), "l2": "input (%.2f GB) and outputs larger than the 126 MB L2" % (len(body) / 1e9),
                       "sharding": "one kernel file per GPU, no data-path collective; counters all-reduced (NCCL) after the timed region"},
            "warp_insts_per_s": all_inst * args.steps / (total_ms * 1e-3),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_nominal_8TBs": achieved / 8000.0, "traffic": traffic_bytes(args.workload), "traffic_unit": "dram bytes per launch (ncu, profiles/ncu_traffic.json)", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": b_alg, "kernel": "ti_decode_kernel", "launch_ms": per_launch_ms},
            "e2e": {"value": all_bytes / e2e_s / 1e9, "unit": "GB/s", "h2d_bytes_per_step": len(body), "d2h_bytes_per_step": int(d2h),
                    "ms_per_step": e2e_s * 1e3, "warp_insts_per_s": all_inst / e2e_s, "steps": args.steps,
                    "result_form": "compact (ti_kernel_decode_streamed_compact): records, 16 B descriptors, rows / transactions that are not closed forms",
                    "d2h_GB_per_s_achieved": d2h / e2e_s / 1e9,
                    "full_form": {"value": all_bytes / e2e_full_s / 1e9, "ms_per_step": e2e_full_s * 1e3, "d2h_bytes_per_step": int(d2h_full),
                                  "d2h_GB_per_s_achieved": d2h_full / e2e_full_s / 1e9},
                    "host_link": link},
            "gpu_launches": 2 * args.steps,   # ti_decode_kernel + ti_check_kernel per step of the device-timed arm
            "clocks": sampler.summary(),
            "counter_allreduce_ms": allreduce_ms,
            "counters_check": {"n_inst_all_ranks": int(total_cnt["n_inst"]), "hist_sum": int(total_cnt["hist"].sum())},
        }
        if world == 1 and not args.no_cpu_baseline:
            with tempfile.TemporaryDirectory() as tmp:
                paths, nbytes = split_sample(img, hdr.body_offset, 1, 64 << 20, tmp)
                wall, n_inst, kind = cpu_run(paths)
            line["cpu_baseline"] = {"value": nbytes / wall / 1e9, "unit": "GB/s", "cores": 1, "kind": kind,
                                    "warp_insts_per_s": n_inst / wall,
                                    "sample": "first %.0f MB (whole thread blocks) of the same corpus, 1 process, parse+convert+coalesce, stdout discarded" % (nbytes / 1e6)}
        if world == 1 and not args.no_extras:
            line["writer_side"] = writer_side(tp, torch)
        print(json.dumps(line))
    tp.kernel_finalizer(kd)
    tp.kernel_finalizer(ke)
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
