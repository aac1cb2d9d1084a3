"""Multi-GPU host logic of the trace-ingest path: one process per GPU, no data-path collective.

The reference has no distributed mode for this path (it is single-threaded C++,
gpu-simulator/main.cc:60-190 walks the command list one kernel at a time).  The path
shards at two levels:

  level 1  kernel files of a kernelslist.g -> ranks   (kernels_for_rank)
  level 2  byte ranges of one kernel file, cut at "\\n#BEGIN_TB" by ti_plan_shards
           (shard_for_rank)

Records, rows and transactions never cross ranks.  The only exchange is the integer
AccelWattch counter block: summed as uint64 over all ranks (allreduce_counters), after
which every rank derives bit-identical doubles with ti_counters_finalize.  torch.distributed
is the plumbing: NCCL on GPUs (bench.py), gloo in the CPU tests.
"""
import numpy as np

from . import records as R


def kernels_for_rank(n_kernels, rank, world):
    """Contiguous blocks of the command list's kernel launches, sizes differing by at most one
    (contiguous rather than round-robin: a rank then owns one run of the list, and the k-th
    thread block of the k-th kernel keeps its file order inside the rank)."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("rank %r of world %r" % (rank, world))
    base, extra = divmod(n_kernels, world)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def shard_for_rank(cuts, rank):
    """(begin, end) byte range of `rank` from the world+1 cut points ti_plan_shards returned."""
    if not 0 <= rank < len(cuts) - 1:
        raise ValueError("rank %r with %d cut points" % (rank, len(cuts)))
    return int(cuts[rank]), int(cuts[rank + 1])


def allreduce_counters(counters, device=None, group=None):
    """Sum one rank's ti_counters block (numpy record of dtype records.COUNTERS) over all
    ranks; returns the summed record.  All fields are uint64 event counts, so the block is
    reduced as one int64 vector (two's-complement addition is the same operation)."""
    import torch
    import torch.distributed as dist

    block = np.zeros(1, dtype=R.COUNTERS)
    block[0] = counters
    vec = np.frombuffer(block.tobytes(), dtype=np.int64).copy()
    t = torch.from_numpy(vec)
    if device is not None:
        t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return np.frombuffer(t.cpu().numpy().tobytes(), dtype=R.COUNTERS)[0]


def exchange_offsets(counts, device=None, group=None):
    """Level-2 sharding: every rank decoded one byte range of the same kernel file and holds
    shard-relative record indices.  All-gathers the five per-rank totals (n_inst, n_mem, n_txn,
    n_tb, n_warp -- 40 bytes per rank, the only other exchange of the path) and returns this
    rank's bases: adding them to first_inst / first_warp / mem_idx / txn_off gives the file-global
    ordinals, i.e. the reference's file-order thread-block hand-out across GPUs."""
    import torch
    import torch.distributed as dist

    names = ("n_inst", "n_mem", "n_txn", "n_tb", "n_warp")
    mine = torch.tensor([int(getattr(counts, n) if hasattr(counts, n) else counts[n]) for n in names], dtype=torch.int64)
    if device is not None:
        mine = mine.to(device)
    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    if world > 1:
        parts = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        allc = torch.stack(parts).cpu().numpy()
    else:
        allc = mine.cpu().numpy()[None, :]
    base = allc[:rank].sum(axis=0) if rank else np.zeros(len(names), dtype=np.int64)
    out = {n[2:]: int(b) for n, b in zip(names, base)}
    out["totals"] = {n[2:]: int(t) for n, t in zip(names, allc.sum(axis=0))}
    return out


def finalize(parser, counters):
    """ti_counters_finalize through a trace_parser: integer block -> AccelWattch doubles."""
    cnt = np.zeros(1, dtype=R.COUNTERS)
    cnt[0] = counters
    pw = np.zeros(1, dtype=R.POWER)
    parser._ck(parser.lib.ti_counters_finalize(parser.ctx, cnt.ctypes.data, pw.ctypes.data))
    return pw[0]


class NcclCounters:
    """The same reduction through the library's own C-ABI collective (ti_allreduce_counters): an NCCL
    communicator created from a unique id that rank 0 draws and torch.distributed carries to the other ranks
    (any transport would do: it is 128 bytes), then one device-to-device ncclAllReduce per call."""

    def __init__(self, parser, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        self._p, self._C = parser, C
        world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        rank = dist.get_rank(group) if world > 1 else 0
        uid = np.zeros(128, dtype=np.uint8)
        if rank == 0:
            parser._ck(parser.lib.ti_nccl_unique_id(uid.ctypes.data))
        if world > 1:
            t = torch.from_numpy(uid).cuda()
            dist.broadcast(t, src=0, group=group)
            uid = t.cpu().numpy()
        self.comm = C.c_void_p()
        parser._ck(parser.lib.ti_nccl_comm_create(parser.ctx, uid.ctypes.data, world, rank, C.byref(self.comm)))
        self.world, self.rank = world, rank

    def allreduce(self, kernel):
        """Sum of the last decode's counter block of `kernel` (a KernelTrace) over the ranks -> numpy record."""
        out = np.zeros(1, dtype=R.COUNTERS)
        self._p._ck(self._p.lib.ti_allreduce_counters(kernel._h, self.comm, None, out.ctypes.data))
        return out[0]

    def close(self):
        if self.comm:
            self._p.lib.ti_nccl_comm_destroy(self.comm)
            self.comm = None
