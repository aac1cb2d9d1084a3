"""ti_allreduce_counters: the C-ABI collective (NCCL, device to device) against ti_counters_add of the ranks'
blocks.  One process per GPU over every GPU the box has (two or more: the real exchange; one: a one-rank
communicator, which still goes through ncclCommInitRank / ncclAllReduce)."""
import os
import socket

import numpy as np
import pytest

import harness
import tiformat as T

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    from accel_sim_framework_distributed_b200 import distributed as D
    from accel_sim_framework_distributed_b200 import trace_parser

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    tp = trace_parser(lib_path=harness.CUDA_LIB, device=rank)
    img = harness.generate_bytes(profile=5, seed=900 + rank, n_tb=200 + 50 * rank, insts_per_warp=12, block_threads=128)
    hdr = tp.parse_header(img[:8192])
    k = tp.open_buffer(np.ascontiguousarray(img[hdr.body_offset:]), hdr)
    d = k.run()
    mine = np.zeros(1, dtype=T.COUNTERS)
    mine[0] = d.counters
    nc = D.NcclCounters(tp)
    total = nc.allreduce(k)
    again = nc.allreduce(k)                     # the kernel's own block is untouched: a second call gives the same sum
    assert all(np.array_equal(total[f], again[f]) for f in T.COUNTERS.names)
    # after the pipelined path the file's block lives on the host: the collective puts it on the device itself
    k.run_streamed(1 << 20)
    streamed = nc.allreduce(k)
    np.save(os.path.join(out_dir, "mine%d.npy" % rank), mine)
    np.save(os.path.join(out_dir, "total%d.npy" % rank), np.array([total, streamed], dtype=T.COUNTERS))
    nc.close()
    tp.kernel_finalizer(k)
    tp.close()
    if world > 1:
        dist.destroy_process_group()


def test_allreduce_counters_over_nccl(tmp_path):
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 8)
    assert world >= 1
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from accel_sim_framework_distributed_b200 import _ffi
    lib = _ffi.load_library(harness.CUDA_LIB)
    want = np.zeros(1, dtype=T.COUNTERS)
    for r in range(world):
        m = np.load(os.path.join(str(tmp_path), "mine%d.npy" % r))
        lib.ti_counters_add(want.ctypes.data, m.ctypes.data)
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), "total%d.npy" % r))
        for which in (0, 1):
            for f in T.COUNTERS.names:
                if f not in ("reserved", "text_bytes") or which == 0:
                    assert np.array_equal(got[which][f], want[0][f]), (r, which, f)
