"""N > 1 host logic on CPU: two processes, torch.distributed with the gloo backend.

The path shards with no data-path collective; the only exchange is the sum of
the integer counter block.  Each rank decodes its share with the SIMT-emulated build of the
kernel source (test infrastructure; on the GPU box bench.py does the same over NCCL with the
CUDA library), the blocks are all-reduced, and every rank must end with the counters and the
AccelWattch doubles the CPU oracle computes for the whole input.

  level 1  BASELINE config 4: all-reduce-shaped kernel traces, one kernel file per rank
  level 2  one kernel file cut into per-rank byte ranges by ti_plan_shards
"""
import os
import socket

import numpy as np
import pytest

import harness
import tiformat as T

WORLD = 2
POWER_RTOL = 1e-9          # north_star: fp64 AccelWattch coefficients within 1e-9 relative
SKIP = ("reserved", "text_bytes")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, mode, paths, out_dir):
    import torch.distributed as dist
    from accel_sim_framework_distributed_b200 import distributed as D
    from accel_sim_framework_distributed_b200 import trace_parser

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        tp = trace_parser(lib_path=harness.emu_library())
        acc = np.zeros(1, dtype=T.COUNTERS)
        if mode == "by_kernel":
            for i in D.kernels_for_rank(len(paths), rank, world):
                k = tp.parse_kernel_info(paths[i])
                d = k.run()
                cnt = np.zeros(1, dtype=T.COUNTERS)
                cnt[0] = d.counters
                tp.lib.ti_counters_add(acc.ctypes.data, cnt.ctypes.data)
                tp.kernel_finalizer(k)
        else:
            k = tp.parse_kernel_info(paths[0])
            body = np.fromfile(paths[0], dtype=np.uint8)[k.hdr.body_offset:]
            b, e = D.shard_for_rank(tp.plan_shards(body, world), rank)
            assert b == 0 or b == e or bytes(body[b:b + 9]) == b"#BEGIN_TB"
            d = k.run(b, e)
            acc[0] = d.counters
            # global ordinals: shard-relative first_inst / first_warp + this rank's bases
            base = D.exchange_offsets(d.counts)
            tb = d.tb.copy()
            tb["first_inst"] += np.uint32(base["inst"]); tb["first_warp"] += np.uint32(base["warp"])
            np.save(os.path.join(out_dir, "tb%d.npy" % rank), tb)
            np.save(os.path.join(out_dir, "base%d.npy" % rank), np.array([base["inst"], base["mem"], base["txn"], base["tb"], base["warp"],
                                                                         base["totals"]["inst"], base["totals"]["tb"]], dtype=np.int64))
            tp.kernel_finalizer(k)
        total = D.allreduce_counters(acc[0])
        power = D.finalize(tp, total)
        np.save(os.path.join(out_dir, "cnt%d.npy" % rank), np.array([total], dtype=T.COUNTERS))
        np.save(os.path.join(out_dir, "pow%d.npy" % rank), np.array([power], dtype=T.POWER))
        np.save(os.path.join(out_dir, "own%d.npy" % rank), acc)
        tp.close()
    finally:
        dist.destroy_process_group()


def _spawn(mode, paths, out_dir):
    import torch.multiprocessing as mp
    harness.emu_library()   # build once in the parent, not concurrently in the ranks
    mp.spawn(_worker, args=(WORLD, _free_port(), mode, paths, out_dir), nprocs=WORLD, join=True)
    cnt = [np.load(os.path.join(out_dir, "cnt%d.npy" % r))[0] for r in range(WORLD)]
    pw = [np.load(os.path.join(out_dir, "pow%d.npy" % r))[0] for r in range(WORLD)]
    own = [np.load(os.path.join(out_dir, "own%d.npy" % r))[0] for r in range(WORLD)]
    return cnt, pw, own


def _oracle_sum(paths, tmp):
    """Counter block of the whole input from the CPU oracle: sum over kernel files."""
    tot, dumps = None, []
    for i, p in enumerate(paths):
        o = T.load_dump(harness.oracle_dump(p, tmp, name="o%d.tidump" % i))
        dumps.append(o)
        c = np.array([o.counters], dtype=T.COUNTERS)
        if tot is None:
            tot = c
        else:
            for f in T.COUNTERS.names:
                tot[0][f] = tot[0][f] + c[0][f]
    return tot[0], dumps


def _check(cnt, pw, want_cnt, want_pw=None):
    for r in range(WORLD):
        msgs = T.compare_counters(cnt[r], want_cnt, skip=SKIP + ("n_begin_tb", "n_end_tb", "n_warp", "n_mem"))
        assert not msgs, "rank %d: %s" % (r, msgs[:6])
    # every rank derives the same doubles, bit for bit
    assert pw[0].tobytes() == pw[1].tobytes()
    if want_pw is not None:
        msgs = T.compare_power(pw[0], want_pw, rtol=POWER_RTOL)
        assert not msgs, msgs[:6]


@pytest.mark.timeout(600)
def test_gloo_sharded_by_kernel_file(tmp_path):
    """config 4 shape: three all-reduce-shaped kernels over two ranks (2 + 1)."""
    tmp = str(tmp_path)
    paths = []
    for i in range(3):
        p = os.path.join(tmp, "kernel-%d.traceg" % (i + 1))
        harness.generate(p, profile=4, seed=400 + i, n_tb=6, insts_per_warp=12, block_threads=128, kernel_id=i + 1)
        paths.append(p)
    cnt, pw, own = _spawn("by_kernel", paths, tmp)
    want, _ = _oracle_sum(paths, tmp)
    _check(cnt, pw, want)
    assert int(own[0]["n_inst"]) > 0 and int(own[1]["n_inst"]) > 0          # both ranks really worked
    assert int(cnt[0]["n_inst"]) == int(own[0]["n_inst"]) + int(own[1]["n_inst"])


@pytest.mark.timeout(600)
def test_gloo_sharded_by_byte_range(tmp_path):
    """One kernel file, cut at #BEGIN_TB into one range per rank; the oracle sees the whole file."""
    tmp = str(tmp_path)
    p = os.path.join(tmp, "kernel-1.traceg")
    harness.generate(p, profile=5, seed=77, n_tb=14, insts_per_warp=9, block_threads=96)
    cnt, pw, own = _spawn("by_range", [p], tmp)
    want, dumps = _oracle_sum([p], tmp)
    _check(cnt, pw, want, dumps[0].power)
    assert int(own[0]["n_inst"]) > 0 and int(own[1]["n_inst"]) > 0
    # exchange_offsets: rank 1's bases are rank 0's totals, and the shifted thread-block table is the whole file's
    b0, b1 = (np.load(os.path.join(tmp, "base%d.npy" % r)) for r in range(WORLD))
    assert not b0[:5].any() and b1[0] == int(own[0]["n_inst"]) and b1[3] == int(own[0]["n_tb"]) and b1[5] == int(cnt[0]["n_inst"])
    tb = np.concatenate([np.load(os.path.join(tmp, "tb%d.npy" % r)) for r in range(WORLD)])
    whole = dumps[0].tb
    assert len(tb) == len(whole) == b1[6]
    for f in ("x", "y", "z", "first_warp", "first_inst"):
        assert np.array_equal(tb[f], whole[f]), f


def test_rank_assignment_helpers():
    from accel_sim_framework_distributed_b200 import distributed as D
    for n in (0, 1, 5, 8, 17):
        for world in (1, 2, 3, 8):
            got = [D.kernels_for_rank(n, r, world) for r in range(world)]
            assert sum(got, []) == list(range(n))
            assert max(len(g) for g in got) - min(len(g) for g in got) <= 1
    assert D.shard_for_rank([0, 10, 25], 1) == (10, 25)
    with pytest.raises(ValueError):
        D.kernels_for_rank(4, 2, 2)
    with pytest.raises(ValueError):
        D.shard_for_rank([0, 10], 1)
