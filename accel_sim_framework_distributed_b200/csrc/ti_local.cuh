// Local-memory address translation + coalescing.
//
// The one memory space the decode kernel leaves alone: the reference rewrites the addresses of
// local-space loads / stores per SM and hardware thread when the instruction issues
// (trace_driven.cc:593-601 -> shader_core_ctx::translate_local_memaddr, shader.cc:1714-1783) and
// coalesces only then (abstract_hardware_model.cc:509-584, up to eight 4-byte accesses per lane).
// That is a function of (trace, placement), so it runs once the scheduler has placed the CTA:
// the caller passes, per issued instruction, the SM and the hardware thread id of lane 0.
// One warp per issued instruction.
#ifndef TI_LOCAL_CUH
#define TI_LOCAL_CUH

#include "ti_kernels.cuh"

namespace ti {

struct LocalParams {
  const ti_inst_rec *inst;
  uint64_t n_inst;
  const uint64_t *rows;
  uint64_t n_rows;
  const ti_local_issue *issues;
  uint64_t n_issues;
  ti_local_map map;
  DecodeParams model;       // coalesce_arch, gmem_skip_L1D (segment sizing)
  uint32_t *txn_cnt;        // [n_issues]     pass 1: transactions per issue
  const uint32_t *txn_off;  // [n_issues + 1] pass 2: exclusive scan of txn_cnt
  ti_txn_rec *txn;          // pass 2
  uint64_t cap_txn;
  uint32_t *error;          // bit 0: issue does not name a local load/store; bit 1: a translate_local_memaddr assert
};

constexpr int kLocalMaxAcc = 8;   // MAX_ACCESSES_PER_INSN_PER_THREAD (abstract_hardware_model.h:1050)

// shader.cc:1714-1783.  Returns the number of 4-byte accesses (0 on a tripped assert).
TI_D unsigned translate_local(const ti_local_map &m, uint64_t localaddr, uint32_t sid, uint32_t tid, unsigned datasize, uint64_t (&out)[kLocalMaxAcc], bool &bad) {
  uint32_t thread_base, max_threads;   // `unsigned` arithmetic in the reference: wraps at 2^32
  if (m.local_mem_map) {
    thread_base = 4u * (m.padded_threads_per_cta * (sid + m.num_shader * (tid / m.padded_threads_per_cta)) + tid % m.padded_threads_per_cta);
    max_threads = m.padded_threads_per_cta * m.max_cta_per_shader * m.num_shader;
  } else {
    thread_base = 4u * (m.n_thread_per_shader * sid + tid);
    max_threads = m.num_shader * m.n_thread_per_shader;
  }
  if (!((uint64_t)thread_base < 4ull * max_threads)) { bad = true; return 0; }
  if (datasize >= 4) {
    const unsigned n = datasize / 4;
    if (datasize % 4 || n > (unsigned)kLocalMaxAcc || localaddr % 4) { bad = true; return 0; }
    for (unsigned i = 0; i < n; ++i) out[i] = (localaddr / 4 + i) * max_threads * 4 + thread_base + m.local_generic_start;
    return n;
  }
  if (datasize == 0 || (localaddr + datasize - 1) / 4 != localaddr / 4) { bad = true; return 0; }
  out[0] = (localaddr / 4) * max_threads * 4 + localaddr % 4 + thread_base + m.local_generic_start;
  return 1;
}

// pass 0 counts (txn == nullptr), pass 1 writes.
__global__ void __launch_bounds__(256) ti_local_kernel(const LocalParams L) {
  const unsigned lane = threadIdx.x & 31;
  const uint64_t warp0 = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
  for (uint64_t q = warp0; q < L.n_issues; q += n_warps) {
    const ti_local_issue is = L.issues[q];
    bool bad_issue = is.inst >= L.n_inst;
    ti_inst_rec r;
    memset(&r, 0, sizeof r);
    if (!bad_issue) r = L.inst[is.inst];
    const unsigned op = r.op, space = r.pack_c & 0xf;
    bad_issue = bad_issue || space != SPACE_LOCAL || !(op == OP_LOAD || op == OP_STORE) || r.mem_idx == 0xFFFFFFFFu || r.mem_idx >= L.n_rows;
    if (bad_issue) {
      if (lane == 0) { atomicOr(L.error, 1u); if (!L.txn) L.txn_cnt[q] = 0; }
      continue;
    }
    const uint32_t mask = r.mask;
    const unsigned ds = r.data_size, cache_op = r.pack_c >> 4 & 7;
    const unsigned is_write = (op == OP_STORE || (r.pack_b >> 4 & 3) == 2) ? 1u : 0u;
    const unsigned seg = segment_size(L.model, ds, cache_op);
    const unsigned dsc = ds >= 4 ? 4u : ds;                        // data_size_coales (AHM.cc:519-526)
    // the lane's translated addresses; the coalescer stops at the first zero one (AHM.cc:532-535)
    uint64_t a[kLocalMaxAcc];
    unsigned n_acc = 0;
    bool bad = false;
    if (mask >> lane & 1) {
      n_acc = translate_local(L.map, L.rows[(uint64_t)r.mem_idx * 32 + lane], is.sid, is.tid0 + lane, ds, a, bad);
      for (unsigned i = 0; i < n_acc; ++i) if (a[i] == 0) { n_acc = i; break; }
    }
    bad = __any_sync(kFull, bad) || (mask && !seg);
    if (bad) {
      if (lane == 0) { atomicOr(L.error, 2u); if (!L.txn) L.txn_cnt[q] = 0; }
      continue;
    }
    // ascending block order over every touch of the warp (std::map in the reference): repeatedly
    // take the smallest block not yet emitted and merge all touches that fall in it
    uint32_t n_txn = 0;
    uint64_t last = 0;
    bool first = true;
    const uint64_t base = L.txn ? L.txn_off[q] : 0;
    for (;;) {
      uint64_t mine = ~0ull;
      for (unsigned i = 0; i < n_acc; ++i) {
        const uint64_t b = a[i] & ~(uint64_t)(seg - 1);
        if ((first || b > last) && b < mine) mine = b;
      }
      const uint64_t blk = warp_min_u64(mine);
      if (blk == ~0ull) break;
      Txn t;
      t.block = blk; t.lanes = 0; t.chunks = 0; t.bytes[0] = t.bytes[1] = t.bytes[2] = t.bytes[3] = 0;
      bool hit = false;
      for (unsigned i = 0; i < n_acc; ++i) {
        if ((a[i] & ~(uint64_t)(seg - 1)) != blk) continue;
        hit = true;
        const unsigned idx = (unsigned)(a[i] & 127);
        t.chunks |= 1u << (idx / 32);
        const unsigned nb = idx + dsc > 128 ? 128 - idx : dsc;    // clipped at MAX_MEMORY_ACCESS_SIZE
        uint32_t w[4];
        byte_words(idx, nb, w);
        t.bytes[0] |= w[0]; t.bytes[1] |= w[1]; t.bytes[2] |= w[2]; t.bytes[3] |= w[3];
      }
      t.lanes = __ballot_sync(kFull, hit);
      t.chunks = __reduce_or_sync(kFull, t.chunks);
#pragma unroll
      for (int k = 0; k < 4; ++k) t.bytes[k] = __reduce_or_sync(kFull, t.bytes[k]);
      if (L.txn && lane == 0 && base + n_txn < L.cap_txn) txn_store(L.txn + base + n_txn, reduce_and_send(t, seg, is_write ? 5u : 1u, is_write));   // LOCAL_ACC_W / LOCAL_ACC_R
      ++n_txn;
      last = blk;
      first = false;
    }
    if (!L.txn && lane == 0) L.txn_cnt[q] = n_txn;
  }
}

}  // namespace ti
#endif
