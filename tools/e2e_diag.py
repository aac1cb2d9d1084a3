"""Where the host-to-host arm's time goes: the streamed compact decode at several shard sizes against a bare
upload of the same pinned text (run on the GPU box: python tools/e2e_diag.py)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests'))

def main():
    torch.cuda.init()
    print("affinity:", bench.pin_to_gpu_numa_node(0))
    import harness
    from accel_sim_framework_distributed_b200 import trace_parser
    tp = trace_parser(lib_path=harness.CUDA_LIB, device=0)
    img = bench.make_corpus("config2", seed=2)
    hdr = tp.parse_header(img[:8192])
    body = img[hdr.body_offset:]
    part = np.ascontiguousarray(body)
    pin = torch.from_numpy(part)
    torch.cuda.cudart().cudaHostRegister(pin.data_ptr(), pin.numel(), 0)
    dev = torch.empty(pin.numel(), dtype=torch.uint8, device="cuda")
    for piece in (64 << 20, 256 << 20):
        for _ in range(2):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for o in range(0, pin.numel(), piece):
                dev[o:o + piece].copy_(pin[o:o + piece], non_blocking=True)
            torch.cuda.synchronize(); dt = time.perf_counter() - t0
        print("bare upload, %d MiB pieces: %.2f ms  %.1f GB/s" % (piece >> 20, dt * 1e3, pin.numel() / dt / 1e9))
    # both directions at once, with the copy sizes of the compact results of a 64 MiB shard (records, descriptors, CSR, headers)
    down_sizes = [40 << 20, 8 << 20, 2 << 20, 800 << 10, 200 << 10]
    hdst = [torch.empty(sz * 16, dtype=torch.uint8).pin_memory() for sz in down_sizes]
    dsrc = [torch.empty(sz, dtype=torch.uint8, device="cuda") for sz in down_sizes]
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        with torch.cuda.stream(s_up):
            for o in range(0, pin.numel(), 64 << 20):
                dev[o:o + (64 << 20)].copy_(pin[o:o + (64 << 20)], non_blocking=True)
        with torch.cuda.stream(s_dn):
            for i in range(16):
                for h, d, sz in zip(hdst, dsrc, down_sizes):
                    h[i * sz:(i + 1) * sz].copy_(d, non_blocking=True)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print("bare duplex: %.0f MB up + %.0f MB down in %.2f ms" % (pin.numel() / 1e6, 16 * sum(down_sizes) / 1e6, dt * 1e3))
    ke = tp.open_buffer(part, hdr)
    trace = bool(os.environ.get("TI_STREAM_TRACE"))   # with a -DTI_STREAM_TRACE build of the library: one timeline
    if trace:
        os.environ.pop("TI_STREAM_TRACE")
        for _ in range(2): ke.run_streamed_compact(64 << 20)
        os.environ["TI_STREAM_TRACE"] = "1"
        ke.run_streamed_compact(64 << 20)
        return
    for mb in (16, 32, 64, 128, 256):
        for _ in range(2): ke.run_streamed_compact(mb << 20)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(5): v = ke.run_streamed_compact(mb << 20)
        torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
        print("streamed compact, %3d MiB shards: %.2f ms  %.1f GB/s  (device decode + gather, summed: %.2f ms; down %.0f MB)" % (mb, dt * 1e3, pin.numel() / dt / 1e9, ke.decode_ms(), v.d2h_bytes / 1e6))
main()
