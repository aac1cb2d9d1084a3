// Body of the fused kernel; included from ti_kernels.cuh.
#ifndef TI_KERNEL_MAIN_CUH
#define TI_KERNEL_MAIN_CUH

namespace ti {

// Optional per-phase cycle accounting (tuning builds only, -DTI_PHASE_TIMING): thread 0 of
// every CTA stamps clock64() at the phase boundaries and adds the deltas to a device array.
#if defined(TI_PHASE_TIMING) && !defined(TI_EMU)
__device__ unsigned long long g_phase_cycles[16];
#define TI_STAMP(i) do { if (threadIdx.x == 0) { const long long t_ = clock64(); atomicAdd(&g_phase_cycles[i], (unsigned long long)(t_ - t_prev)); t_prev = t_; } } while (0)
#else
#define TI_STAMP(i) do { } while (0)
#endif

TI_D void flush_counters(uint32_t *cnt, unsigned long long *g) {
  for (int i = threadIdx.x; i < kSmCnt; i += kThreads) {
    const uint32_t v = cnt[i];
    if (v) {
      // smem slots [0, kCntConst] mirror the same u64 words; then n_begin, n_end, smem_cycles and the
      // number of "insts = n" lines (ti_counters.reserved[0], must equal the warp headers)
      const int dst = i <= kCntConst ? i : (i == kCntConst + 1 ? kCntNInst + 5 : i == kCntConst + 2 ? kCntNInst + 6 : i == kCntConst + 3 ? kCntNInst + 8 : kCntNInst + 9);
      atomicAdd(g + dst, (unsigned long long)v);
      cnt[i] = 0;
    }
  }
}

__global__ void __launch_bounds__(kThreads, TI_MIN_BLOCKS) ti_decode_kernel(const DecodeParams P) {
  TI_DYN_SMEM(smem);
  ti_inst_rec *recs = reinterpret_cast<ti_inst_rec *>(smem + oRecs);
  ti_txn_rec *stage = reinterpret_cast<ti_txn_rec *>(smem + oPool);      // pool slot s = stage[s]
  uint16_t *slowq = reinterpret_cast<uint16_t *>(smem + oSlowQ);
  uint32_t *cnt = reinterpret_cast<uint32_t *>(smem + oCnt);
  uint16_t *nlmap = reinterpret_cast<uint16_t *>(smem + oNlMap);
  const uint32_t *nlmap32 = reinterpret_cast<const uint32_t *>(smem + oNlMap);
  uint16_t *dlmap = reinterpret_cast<uint16_t *>(smem + oDlMap);
  const uint32_t *dlmap32 = reinterpret_cast<const uint32_t *>(smem + oDlMap);
  uint32_t *wscratch = reinterpret_cast<uint32_t *>(smem + oTok + (threadIdx.x >> 5) * kWarpScratch);
  uint16_t *tokpos = reinterpret_cast<uint16_t *>(wscratch);
  Misc *ms = reinterpret_cast<Misc *>(smem + oMisc);
  MemoEnt *memo = reinterpret_cast<MemoEnt *>(smem + oMemo);

  const unsigned tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint32_t lt = (1u << lane) - 1;
  uint16_t *wl_inst = reinterpret_cast<uint16_t *>(smem + oWlInst) + wid * kWarpInst;   // this warp's line lists
  uint16_t *wl_sl = reinterpret_cast<uint16_t *>(smem + oWlSl) + wid * kWarpSl;
  for (int i = tid; i < kSmCnt; i += kThreads) cnt[i] = 0;
  for (int i = tid; i < kMemoSlots * 4; i += kThreads) reinterpret_cast<unsigned long long *>(memo)[i] = 0;
  unsigned done_chunks = 0;
  unsigned long long fast_lines = 0;   // instruction lines the fast tokenizer decoded, this thread (ti_counters.reserved[4], a path statistic)
#if defined(TI_PHASE_TIMING) && !defined(TI_EMU)
  long long t_prev = clock64();
#endif

  // ---- stage-in helpers.  Buffer offset j holds text byte (chunk * kChunk - 16 + j).  The
  // part of [0, 16 + kChunk + kSlack) that lies inside the shard on 16 B boundaries comes by
  // one TMA bulk copy (issued by one thread); the rest -- the virtual '\n' before the shard,
  // the tail past its end, a last partial vector, the sentinel -- is filled by all threads.
  auto bulk_range = [&](uint32_t c, uint32_t &v_lo, uint32_t &v_hi) {  // vector range of the bulk copy
    const uint64_t cb = (uint64_t)c * kChunk;
    v_lo = c == 0 ? 1u : 0u;
    const uint64_t full = P.len / 16;                                   // whole vectors of the shard
    const uint64_t first = cb / 16 - (c == 0 ? 0 : 1);                  // text vector held by buffer vector v_lo
    const uint64_t avail = full > first ? full - first : 0;
    const uint64_t want = (uint64_t)(16 + kChunk + kSlack) / 16 - v_lo;
    v_hi = v_lo + (uint32_t)(avail < want ? avail : want);
  };
  auto issue_bulk = [&](uint32_t c, unsigned buf) {                     // one thread
    uint32_t v_lo, v_hi;
    bulk_range(c, v_lo, v_hi);
    uint8_t *dst = smem + oText + buf * kTextBytes;
    const unsigned bytes = (v_hi - v_lo) * 16;
    fence_proxy_async();                       // the buffer was last touched through the generic proxy
    mbar_expect_tx(&ms->mbar[buf], bytes);
    if (bytes) bulk_g2s(dst + v_lo * 16, P.text + ((uint64_t)c * kChunk - 16 + (uint64_t)v_lo * 16), bytes, &ms->mbar[buf]);
  };
  auto fill_edges = [&](uint32_t c, unsigned buf) {                     // all threads
    uint32_t v_lo, v_hi;
    bulk_range(c, v_lo, v_hi);
    uint4 *tv = reinterpret_cast<uint4 *>(smem + oText + buf * kTextBytes);
    const uint64_t cb = (uint64_t)c * kChunk;
    for (uint32_t v = tid; v < (uint32_t)(kTextBytes / 16); v += kThreads) {
      if (v >= v_lo && v < v_hi) continue;
      const int64_t pos = (int64_t)cb - 16 + (int64_t)v * 16;
      uint32_t w[4] = {0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au, 0x0a0a0a0au};
      if (v < (uint32_t)((16 + kChunk + kSlack) / 16) && pos >= 0)
        for (int b = 0; b < 16 && (uint64_t)pos + b < P.len; ++b)   // the shard's last, partial vector
          w[b >> 2] = (w[b >> 2] & ~(0xffu << (8 * (b & 3)))) | (uint32_t)P.text[pos + b] << (8 * (b & 3));
      tv[v] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  };


  // ---- F: a chunk's address rows, CSR offsets and transactions leave, in file order.  This is where
  // chains 2 (memory instructions) and 3 (transactions) resolve.  It normally runs one iteration after
  // the chunk was decoded (`pend`), so that every predecessor has had a whole iteration to publish;
  // everything it needs was left in shared memory under the chunk's parity `par`: the base+stride
  // descriptors, txn_cnt / txn_loc / txn_slot / m_io, the per-warp offsets (Misc::pw) and, in the pool,
  // what the warp path staged.  The records left earlier with a scratch value in mem_idx; the row index
  // is patched here.  After the look-back every warp handles the items of its own range.
  // `recompute` (CTA-uniform): the staging area overflowed, so the spilled warp-path instructions are
  // coalesced again from their address rows and their records, both read back from global memory (the
  // warp that reads a record is the one whose bulk store wrote it).
  auto finish = [&](uint32_t c, unsigned par, uint32_t n_mem, uint32_t n_txn_chunk, uint64_t base_inst,
                    bool have_mem, uint64_t base_mem, bool recompute, const LbPre *pre) {
    if (lane == 0) bulk_store_wait_all();              // this warp's record block is in global memory: the patch below lands on top of it
    __syncwarp();
    uint64_t ex[4];
    uint64_t base_txn;
    if (have_mem) { cta_resolve(P.chain[3], nullptr, c, chain_pack(n_txn_chunk, 0), 0, ex, ms->lb, pre); base_txn = ex[0]; }
    else { cta_resolve(P.chain[2], P.chain[3], c, chain_pack(n_mem, 0), chain_pack(n_txn_chunk, 0), ex, ms->lb, pre); base_mem = ex[0]; base_txn = ex[2]; }
    if (tid == 0) {
      uint32_t ov = 0;
      if (base_mem + n_mem > P.cap_mem) ov |= 2;
      if (base_txn + n_txn_chunk > P.cap_txn) ov |= 4;
      if (base_txn + n_txn_chunk > 0xfffffffeull || base_mem + n_mem > 0x7ffffffeull) ov |= 32;
      if (ov) atomicOr(&P.g->overflow, ov);
      if (c == P.n_chunks - 1) {                       // last chunk: CSR sentinel + totals
        if (base_mem + n_mem <= P.cap_mem) P.txn_off[base_mem + n_mem] = (uint32_t)(base_txn + n_txn_chunk + P.off_txn);
        P.counters[kCntNInst + 1] = base_mem + n_mem;
        P.counters[kCntNInst + 2] = base_txn + n_txn_chunk;
      }
    }
    TI_STAMP(11);  // (S +) F: look-back
    const WarpPre pw = ms->pw[par][wid];
    const StrideNote *wnotes = reinterpret_cast<const StrideNote *>(smem + oNotes) + (par * kWarps + wid) * kNotesPerWarp;
    const StrideNote *cnotes = reinterpret_cast<const StrideNote *>(smem + oNotes) + par * kWarps * kNotesPerWarp;
    const uint16_t *txn_slot = reinterpret_cast<const uint16_t *>(smem + oTxnSlot) + par * (align16(kItems * 2) / 2) + wid * kWarpInst;
    const uint16_t *txn_loc = reinterpret_cast<const uint16_t *>(smem + oTxnLoc) + par * (align16(kItems * 2) / 2) + wid * kWarpInst;
    const uint8_t *txn_cnt = smem + oTxnCnt + par * align16(kItems) + wid * kWarpInst;
    const uint8_t *m_io = smem + oMIo + par * align16(kItems) + wid * kWarpInst;
    const uint64_t wmem = base_mem + pw.mem0, wtxn = base_txn + pw.txn0, winst = base_inst + pw.inst0;
    // compact mode: the 16 B descriptor stands for the row (and, for a dense progression, for the transactions)
    if (P.desc) {
      for (uint32_t fo = lane; fo < pw.n_notes; fo += 32) {
        const StrideNote nt = wnotes[fo];
        const uint64_t row = wmem + ((nt.meta >> 18 & 0x3ff) - wid * kWarpInst);
        if (!(nt.meta >> 28 & 1) && row < P.cap_mem) {
          ti_mem_desc dd;
          dd.base = nt.base; dd.stride = nt.stride; dd.meta = nt.meta & 0x3ffffu;
          P.desc[row] = dd;
        }
      }
    }
    // address rows of the base+stride instructions: half a warp per row, two addresses (16 B) per lane,
    // so every store instruction covers two whole 256 B rows
    for (uint32_t f0 = 0; f0 < (P.desc ? 0u : (uint32_t)pw.n_notes); f0 += 2) {
      const uint32_t fo = f0 + (lane >> 4);
      if (fo < pw.n_notes) {
        const StrideNote nt = wnotes[fo];
        const uint32_t meta = nt.meta;
        const uint64_t row = wmem + ((meta >> 18 & 0x3ff) - wid * kWarpInst);
        if (!(meta >> 28 & 1) && row < P.cap_mem) {
          const unsigned first = meta & 31, n = meta >> 5 & 63, l = (lane & 15) * 2;
          const uint64_t a0 = l - first < n ? nt.base + (uint64_t)(int64_t)nt.stride * (uint64_t)(l - first) : 0;
          const uint64_t a1 = l + 1 - first < n ? nt.base + (uint64_t)(int64_t)nt.stride * (uint64_t)(l + 1 - first) : 0;
          reinterpret_cast<ulonglong2 *>(P.rows + row * 32)[lane & 15] = make_ulonglong2(a0, a1);
        }
      }
    }
    TI_STAMP(12);  // F: rows
    // per memory instruction: CSR offset and the row index in its record
    for (uint32_t j = lane; j < pw.n_mem; j += 32) {
      const uint64_t row = wmem + j;
      if (row < P.cap_mem) P.txn_off[row] = (uint32_t)(wtxn + txn_loc[j] + P.off_txn);
      const uint64_t ri = winst + m_io[j];
      if (ri < P.cap_inst) P.inst[ri].mem_idx = (uint32_t)(row + P.off_mem);
    }
    // its transactions: four lanes per instruction.  A dense base+stride progression has every transaction in
    // closed form (lane q writes transactions q, q+4, ...); the copy of what the warp path staged is split the
    // same way; the streaming merge of other progressions is one lane's work.
    for (uint32_t j = lane >> 2; j < pw.n_mem; j += 8) {
      const uint32_t sub = lane & 3;
      const uint32_t n = txn_cnt[j], slot = txn_slot[j];
      const uint64_t at = wtxn + txn_loc[j];
      if (!n || slot == 0xFFFFu || at >= P.cap_txn) continue;
      const unsigned room = (unsigned)(P.cap_txn - at < n ? P.cap_txn - at : n);
      if (slot & 0x8000u) {                            // base+stride: straight from the descriptor
        const StrideNote nt = cnotes[slot & 0x7fffu];
        const uint32_t meta = nt.meta;
        StrideDesc d;
        d.base = nt.base; d.stride = nt.stride; d.first = meta & 31; d.n = meta >> 5 & 63;
        const unsigned ds = 1u << (meta >> 11 & 7), seg = 32u << (meta >> 14 & 3), is_write = meta >> 16 & 1;
        if (meta >> 17 & 1) {
          if (!P.desc)
            for (unsigned k = sub; k < room; k += 4) txn_store(P.txn + at + k, dense_txn(d, ds, seg, is_write ? 4u : 0u, is_write, k));
        } else if (sub == 0) stride_emit(d, ds, seg, is_write ? 4u : 0u, is_write, P.txn + at, room);
      } else {                                         // warp path: staged in the pool, in 16 B units
        const uint4 *src = reinterpret_cast<const uint4 *>(stage) + (slot & 0x3fffu);
        if (slot & 0x4000u) {                            // compact form: one unit per transaction
          for (uint32_t k = sub; k < room; k += 4) txn_store(P.txn + at + k, txn_unpack_run(src[k]));
        } else {
          uint4 *dst = reinterpret_cast<uint4 *>(P.txn + at);
          for (uint32_t k = sub; k < room; k += 4) { dst[2 * k] = src[2 * k]; dst[2 * k + 1] = src[2 * k + 1]; }
        }
      }
    }
    TI_STAMP(13);  // F: offsets, record patch, transactions
    if (recompute) {                                   // CTA-uniform
      for (uint32_t j = 0; j < pw.n_mem; ++j) {
        if (txn_slot[j] != 0xFFFFu || !txn_cnt[j]) continue;
        const uint64_t at = wtxn + txn_loc[j];
        const uint64_t ri = winst + m_io[j];
        if (ri >= P.cap_inst) continue;                // the record table overflowed: the host decodes again
        ti_inst_rec rr;
        {
          const uint4 *g = reinterpret_cast<const uint4 *>(P.inst + ri);
          uint4 *d = reinterpret_cast<uint4 *>(&rr);
          d[0] = __ldcg(g); d[1] = __ldcg(g + 1);
        }
        const ti_inst_rec *rp = &rr;
        const uint64_t row = wmem + j;
        const uint64_t addr = row < P.cap_mem ? __ldcg(P.rows + row * 32 + lane) : 0;   // written by the warp path
        const unsigned ds = rp->data_size, pack_b = rp->pack_b, op = rp->op;
        const unsigned seg = segment_size(P, ds, rp->pack_c >> 4 & 7);
        const unsigned is_write = (op == OP_STORE || op == OP_TSTORE || (pack_b >> 4 & 3) == 2) ? 1u : 0u;
        Txn t2[2];
        unsigned pos[2] = {~0u, ~0u};
        uint32_t fl = 0;
        unsigned m;
        if (pack_b >> 6 & 1) m = coalesce_atomic(P, rp->mask, addr, ds, seg, t2[0], pos[0], fl);
        else m = coalesce_global(P, rp->mask, addr, ds, seg, t2, pos, wscratch);
        const unsigned type = is_write ? 4u : 0u;
        if (pos[0] < m && at + pos[0] < P.cap_txn) txn_store(P.txn + at + pos[0], reduce_and_send(t2[0], seg, type, is_write));
        if (pos[1] < m && at + pos[1] < P.cap_txn) txn_store(P.txn + at + pos[1], reduce_and_send(t2[1], seg, type, is_write));
      }
    }
    __syncthreads();   // the pool may be reused
  };
  bool pend = false, pend_have_mem = false, pend_recompute = false;   // a chunk whose rows and transactions are still parked in shared memory (CTA-uniform)
  uint32_t pend_chunk = 0, pend_n_mem = 0, pend_n_txn = 0;
  unsigned pend_par = 0;
  uint64_t pend_base_inst = 0, pend_base_mem = 0;

  // prologue: first ticket, first bulk copy
  if (tid == 0) {
    mbar_init(&ms->mbar[0], 1);
    mbar_init(&ms->mbar[1], 1);
    ms->chunk_next = atomicAdd(&P.g->ticket, 1u);
    if (ms->chunk_next < P.n_chunks) issue_bulk(ms->chunk_next, 0);
    ms->next_mem = 0; ms->stage_used = 0; ms->err = 0; ms->first_err = 0xffffffffu; ms->n_slowq = 0; ms->n_general = 0;
  }
  __syncthreads();
  if (ms->chunk_next < P.n_chunks) fill_edges(ms->chunk_next, 0);

  for (unsigned iter = 0;; ++iter) {
    __syncthreads();  // previous chunk fully drained; chunk_next and the edge fills of this chunk's buffer visible
    const unsigned buf = TI_TEXT_BUFS == 2 ? (iter & 1) : 0u;
    const unsigned par = iter & 1;
    const uint32_t chunk = ms->chunk_next;
    if (chunk >= P.n_chunks) break;
    uint8_t *text = smem + oText + buf * kTextBytes;
    const uint32_t *text32 = reinterpret_cast<const uint32_t *>(text);
    uint4 *text_v = reinterpret_cast<uint4 *>(text);
    StrideNote *notes = reinterpret_cast<StrideNote *>(smem + oNotes) + (par * kWarps + wid) * kNotesPerWarp;   // this warp's descriptors
    uint16_t *txn_slot = reinterpret_cast<uint16_t *>(smem + oTxnSlot) + par * (align16(kItems * 2) / 2);      // chunk-wide, indexed by item
    uint16_t *txn_loc = reinterpret_cast<uint16_t *>(smem + oTxnLoc) + par * (align16(kItems * 2) / 2);
    uint8_t *txn_cnt = smem + oTxnCnt + par * align16(kItems);
    uint8_t *m_io = smem + oMIo + par * align16(kItems);
    const uint64_t cb = (uint64_t)chunk * kChunk;            // text offset of the chunk
    const uint32_t tbase = (uint32_t)(cb - 16);              // text offset = tbase + smem offset (mod 2^32)
    const uint32_t lim = (uint32_t)(P.len - cb < (uint64_t)kChunk ? P.len - cb : (uint64_t)kChunk);
    TI_STAMP(0);   // top-of-loop
    mbar_wait(&ms->mbar[buf], TI_TEXT_BUFS == 2 ? ((iter >> 1) & 1) : (iter & 1));  // P0: this chunk's bulk copy has landed
    TI_STAMP(1);

    // ---------------------------------------------------------------- P1 classify every staged byte
    // newline and delimiter bitmaps of the whole staged region: any thread can then find the end of a
    // line or the tokens of a line with find-first-set
    for (int v = tid; v < kTextBytes / 16; v += kThreads) {
      uint32_t nl16, dl16;
      cls_masks16(text_v[v], nl16, dl16);
      nlmap[v] = (uint16_t)nl16; dlmap[v] = (uint16_t)dl16;
    }
    __syncthreads();
    TI_STAMP(2);   // P1

    // ---------------------------------------------------------------- W the warp's own range
    // A line start is a byte that follows '\n' and is not '\n' itself (blank lines never become lines,
    // trace_parser.cc:408).  Lane l scans kLaneBits bytes of the range; a shuffle scan orders the starts.
    uint32_t n_inst_w, n_sl_w, n_tb_w, n_warp_w;
    {
      const unsigned p = 16 + wid * kRange + lane * kLaneBits;
      unsigned long long nl = win32(nlmap32, p);
      if (kLaneBits > 32) nl |= (unsigned long long)win32(nlmap32, p + 32) << 32;
      const unsigned long long lanebits = kLaneBits == 64 ? ~0ull : (1ull << (kLaneBits & 63)) - 1;
      nl &= lanebits;
      const unsigned prev = (nlmap32[(p - 1) >> 5] >> ((p - 1) & 31)) & 1;
      unsigned long long st = ((nl << 1) | prev) & ~nl & lanebits;
      const unsigned end = 16 + lim;                        // first smem offset behind the shard's text
      if (p >= end) st = 0;
      else if (p + kLaneBits > end) st &= (1ull << (end - p)) - 1;
      unsigned long long isinst = 0, packed = 0;            // which starts are instruction lines; counts inst | sl << 16 | tb << 32 | warp << 48
      for (unsigned long long b = st; b; b &= b - 1) {
        const unsigned k = __ffsll((long long)b) - 1;
        const unsigned c = text[p + k];
        if (hexval(c) < 16) { packed += 1ull; isinst |= 1ull << k; }
        else packed += (1ull << 16) + (c == 't' ? 1ull << 32 : 0ull) + (c == 'w' ? 1ull << 48 : 0ull);
      }
      unsigned long long inc = packed;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const unsigned long long u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
      const unsigned long long tot = __shfl_sync(kFull, inc, 31);
      n_inst_w = (uint32_t)tot & 0xffff; n_sl_w = (uint32_t)(tot >> 16) & 0xffff; n_tb_w = (uint32_t)(tot >> 32) & 0xffff; n_warp_w = (uint32_t)(tot >> 48);
      if (n_inst_w > (uint32_t)kWarpInst || n_sl_w > (uint32_t)kWarpSl) {   // more lines than any tracer-written file can hold: flag, treat the range as empty
        if (lane == 0) note_error(ms, TI_EF_DENSE_CHUNK, (uint32_t)cb);
        n_inst_w = n_sl_w = n_tb_w = n_warp_w = 0;
      } else {
        uint32_t ai = (uint32_t)(inc - packed) & 0xffff, as = (uint32_t)((inc - packed) >> 16) & 0xffff;
        for (unsigned long long b = st; b; b &= b - 1) {
          const unsigned k = __ffsll((long long)b) - 1;
          if (isinst >> k & 1) wl_inst[ai++] = (uint16_t)(p + k); else wl_sl[as++] = (uint16_t)(p + k);
        }
      }
      __syncwarp();
    }
    TI_STAMP(3);   // line starts

    // A lane per instruction line: fast tokenizer (ti_fast.cuh), general parser for what it declines (it
    // also fills the memo); the record goes to the warp's record region.  Memory instructions get a slot
    // ("item") in file order; base+stride ones are parsed and counted right here, the others queue up for
    // the warp path.  The counters are fed from the registers the record was built in.
    if (lane == 0) bulk_store_wait_read();                  // the warp's previous record block has left shared memory
    __syncwarp();
    const bool all_general = P.trace_version < 3;           // four leading decimals per line: general parser only
    const bool stride_path = P.mem_warp_parts == 1;
    uint32_t mem_seen = 0, fast_seen = 0, txn_seen = 0;
    for (uint32_t i0 = 0; i0 < n_inst_w; i0 += 32) {
      const uint32_t i = i0 + lane;
      const bool valid = i < n_inst_w;
      uint4 rlo = make_uint4(0, 0, 0, 0xFFFFFFFFu), rhi = make_uint4(0, 0, 0, 0);
      unsigned kd = 0, le = 0;
      if (valid) {
        const unsigned o = wl_inst[i];
        le = line_end(nlmap32, o);
        if (!all_general && le - o <= (unsigned)TI_MAX_LINE && parse_inst_fast(text, dlmap32, memo, o, le, P, tbase + (uint32_t)P.off_text, rlo, rhi, kd)) {
          ++fast_lines;
        } else {
          ti_inst_rec r;
          uint64_t facts;
          uint32_t fl = parse_inst_line(text, dlmap32, (int)o, le, P, tbase + (uint32_t)P.off_text, r, kd, facts);
          if (le - o > (unsigned)TI_MAX_LINE) fl |= TI_EF_LINE_TOO_LONG;
          if (fl) { note_error(ms, fl, tbase + o); kd = 0; r.mem_idx = 0xFFFFFFFFu; r.pack_c &= 0x7f; }
          else if (facts) {
            uint64_t k0, k1, k2;
            memo_key(text32, r.opc_off - (tbase + (uint32_t)P.off_text), r.opcode_len, k0, k1, k2);
            memo_insert(memo, k0, k1, k2, facts);
          }
          const uint4 *src = reinterpret_cast<const uint4 *>(&r);
          rlo = src[0]; rhi = src[1];
        }
        uint4 *dst = reinterpret_cast<uint4 *>(recs + wid * kWarpInst + i);
        dst[0] = rlo; dst[1] = rhi;
      }
      const bool has_mem = valid && (kd & 1);
      const uint32_t mb = __ballot_sync(kFull, has_mem);
      const bool fastp = has_mem && stride_path && (kd & 2);
      const uint32_t fb = __ballot_sync(kFull, fastp);
      const unsigned item = wid * kWarpInst + mem_seen + __popc(mb & lt);
      unsigned n_txn = 0;
      bool to_warp_path = has_mem && !fastp;
      if (has_mem) m_io[item] = (uint8_t)i;
      if (fastp) {
        const unsigned fo = fast_seen + __popc(fb & lt);
        const uint32_t mask = rlo.y, tmp = rlo.w;
        const unsigned ds = rhi.z >> 16 & 0xff, op = rhi.y >> 16 & 0xff;
        StrideDesc d;
        uint32_t fl = 0;
        bool parsed;
        if ((tmp >> 16 & 3) == 2) {
          // one active lane in base+delta form: "0x<base> <immediate>".  Whatever the fast token reader does not take
          // (no immediate, a long one, ...) is left to the warp path, which reads it the way the reference's stream does.
          uint64_t b;
          int64_t imm;
          parsed = stride_tokens_fast(text, text32, dlmap32, tmp & 0xffff, le, b, imm);
          d.base = b; d.stride = (int64_t)(ds ? ds : 1); d.first = 0; d.n = 0;   // a run of one: the stride is never applied
          parsed = parsed && stride_finish(mask, ds ? ds : 1, d);
        } else parsed = stride_parse_fast(text, dlmap32, tmp & 0xffff, le, mask, ds ? ds : 1, d, fl);
        if (fo >= (unsigned)kNotesPerWarp || !parsed) {
          // no room for a descriptor, or the progression wraps around 2^64 (not monotone): warp path
          to_warp_path = true;
          if (fo < (unsigned)kNotesPerWarp) notes[fo].meta = 1u << 28; else atomicAdd(&P.g->rare[3], 1u);
        } else {
          const unsigned seg = segment_size(P, ds, rhi.z >> 12 & 7);
          if (!fl && mask && !seg) fl = TI_EF_BAD_WIDTH;
          if (fl) { note_error(ms, fl, rlo.z); d.n = 0; }
          const bool dense = !fl && mask && stride_is_dense(d, ds);
          n_txn = (fl || !mask) ? 0u : dense ? dense_count(d, ds, seg) : stride_count(d, ds, seg);
          const unsigned is_write = op == OP_STORE ? 1u : 0u;
          StrideNote nt;
          nt.base = d.base; nt.stride = (int32_t)d.stride;
          nt.meta = d.first | d.n << 5 | (unsigned)(__ffs(ds ? ds : 1) - 1) << 11 | (seg == 128 ? 2u : seg == 64 ? 1u : 0u) << 14 | is_write << 16 |
                    (dense ? 1u : 0u) << 17 | item << 18;
          notes[fo] = nt;
          txn_cnt[item] = (uint8_t)n_txn; txn_slot[item] = (uint16_t)(0x8000u | ((wid * kNotesPerWarp + fo)));
        }
      }
      if (to_warp_path) {
        txn_cnt[item] = 0; txn_slot[item] = 0xFFFFu;
        slowq[atomicAdd(&ms->n_slowq, 1u)] = (uint16_t)item;
      }
      uint32_t inc = n_txn;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
      if (has_mem) txn_loc[item] = (uint16_t)(txn_seen + inc - n_txn);
      txn_seen += __shfl_sync(kFull, inc, 31);
      mem_seen += __popc(mb); fast_seen += __popc(fb);
      // counters: instructions with equal (sp_op, active lanes, op, space, store, const) share one atomic.  The
      // space of a warp-path instruction may still change (generic LD / ST): that one is counted there.
      {
        const unsigned op = rhi.y >> 16 & 0xff, sp = rhi.y >> 24 & 0x1f, act = __popc(rlo.y), pack_b = rhi.z & 0xff, pack_c = rhi.z >> 8 & 0xff;
        const bool ldst = (op == OP_LOAD || op == OP_STORE || op == OP_MEMBAR || op == OP_TLOAD || op == OP_TSTORE) && !to_warp_path;
        const unsigned stw = (op == OP_STORE || op == OP_TSTORE || (pack_b >> 4 & 3) == 2) ? 1u : 0u;
        const unsigned sidx = (pack_c & 0xf) * 2 + stw, cc = pack_b >> 7;
        const uint32_t key = valid ? (sp * 33 + act) | fold_op(op) << 10 | (ldst ? 1u : 0u) << 15 | sidx << 16 | cc << 21 : 0xffffffffu;
        const uint32_t grp = __match_any_sync(kFull, key);
        if (valid && (unsigned)(__ffs(grp) - 1) == lane) {
          const uint32_t k = __popc(grp);
          atomicAdd(&cnt[kCntHist + sp * 33 + act], k);
          atomicAdd(&cnt[kCntOp + fold_op(op)], k);
          if (ldst && act) atomicAdd(&cnt[kCntSpace + sidx], act * k);
          if (cc) atomicAdd(&cnt[kCntConst], k);
        }
      }
    }
    const uint32_t n_notes_w = fast_seen < (uint32_t)kNotesPerWarp ? fast_seen : (uint32_t)kNotesPerWarp;
    if (lane == 0) {
      WarpTot t;
      t.n_inst = (uint16_t)n_inst_w; t.n_sl = (uint16_t)n_sl_w; t.n_tb = (uint16_t)n_tb_w; t.n_warp = (uint16_t)n_warp_w;
      t.n_mem = (uint16_t)mem_seen; t.n_notes = (uint16_t)n_notes_w; t.n_txn = txn_seen;
      ms->wt[wid] = t;
    }
    __syncthreads();
    TI_STAMP(4);   // W

    // ---------------------------------------------------------------- S offsets of the warps inside the chunk; chains 0 / 1 / 2
    // (lane k takes warp k's totals; sums and the sums over the warps before this one by warp reductions)
    uint32_t n_inst, n_sl, n_tb, n_warp, n_mem, n_txn_chunk, inst0, tb0, warp0, mem0, txn0, sl0;
    {
      const bool in = lane < (unsigned)kWarps, before = lane < wid;
      WarpTot t = ms->wt[in ? lane : 0];
      if (!in) { t.n_inst = t.n_sl = t.n_tb = t.n_warp = t.n_mem = 0; t.n_txn = 0; }
      // two 16-bit counts per reduction: no field of a chunk exceeds 16 bits
      const uint32_t a = t.n_inst | (uint32_t)t.n_tb << 16, b = t.n_warp | (uint32_t)t.n_mem << 16;
      const uint32_t sa = __reduce_add_sync(kFull, a), sb = __reduce_add_sync(kFull, b), pa = __reduce_add_sync(kFull, before ? a : 0u), pb = __reduce_add_sync(kFull, before ? b : 0u);
      n_inst = sa & 0xffff; n_tb = sa >> 16; n_warp = sb & 0xffff; n_mem = sb >> 16;
      inst0 = pa & 0xffff; tb0 = pa >> 16; warp0 = pb & 0xffff; mem0 = pb >> 16;
      n_sl = __reduce_add_sync(kFull, (uint32_t)t.n_sl); sl0 = __reduce_add_sync(kFull, before ? (uint32_t)t.n_sl : 0u);
      n_txn_chunk = __reduce_add_sync(kFull, t.n_txn); txn0 = __reduce_add_sync(kFull, before ? t.n_txn : 0u);
    }
    const uint32_t n_warp_path = ms->n_slowq;               // CTA-uniform
    if (lane == 0) { WarpPre q; q.inst0 = (uint16_t)inst0; q.mem0 = (uint16_t)mem0; q.n_mem = (uint16_t)mem_seen; q.n_notes = (uint16_t)n_notes_w; q.txn0 = txn0; ms->pw[par][wid] = q; }
    const unsigned long long aggA0 = chain_pack(n_inst, n_tb), aggA1 = chain_pack(n_warp, 0);
    if (tid == 0) {
      chain_publish(P.chain[0], chunk, aggA0); chain_publish(P.chain[1], chunk, aggA1); chain_publish(P.chain[2], chunk, chain_pack(n_mem, 0));
      if (!n_warp_path) chain_publish(P.chain[3], chunk, chain_pack(n_txn_chunk, 0));   // every transaction count is known already
    }

    // the warps list their structural lines for the CTA-wide pass H (if the chunk's lines fit the list)
    const bool sl_listed = n_sl <= (uint32_t)kSlList;          // CTA-uniform
    if (sl_listed) {
      uint32_t tb_seen = tb0, warp_seen = warp0;
      for (uint32_t i0 = 0; i0 < n_sl_w; i0 += 32) {
        const uint32_t i = i0 + lane;
        const bool valid = i < n_sl_w;
        const int o = valid ? wl_sl[i] : 0;
        const unsigned c = valid ? text[o] : 0u;
        const uint32_t tbb = __ballot_sync(kFull, c == 't'), wb = __ballot_sync(kFull, c == 'w');
        const uint32_t ex_tb = tb_seen + __popc(tbb & lt), ex_w = warp_seen + __popc(wb & lt);
        tb_seen += __popc(tbb); warp_seen += __popc(wb);
        if (!valid) continue;
        uint32_t lo = 0, hi = n_inst_w;              // instruction lines before this one = its rank in the warp's sorted line list
        while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (wl_inst[mid] < o) lo = mid + 1; else hi = mid; }
        reinterpret_cast<uint4 *>(smem + oSlList)[sl0 + i] = make_uint4((uint32_t)o, inst0 + lo, ex_tb, ex_w);
      }
    }

    // ---------------------------------------------------------------- F' the previous chunk's rows and transactions
    // (before anything of this chunk goes into the pool)
    if (pend) {
      const LbPre preF = pend_have_mem ? cta_lookahead(P.chain[3], nullptr, pend_chunk) : cta_lookahead(P.chain[2], P.chain[3], pend_chunk);
      finish(pend_chunk, pend_par, pend_n_mem, pend_n_txn, pend_base_inst, pend_have_mem, pend_base_mem, pend_recompute, &preF);
      pend = false;
    } else __syncthreads();                                // Misc::pw visible to the warp path below
    TI_STAMP(5);   // S + F'

    // ---------------------------------------------------------------- pool layout of this chunk
    // Thread-block / warp headers are parked chunk-relative at the top of the pool until chains 0 / 1
    // resolve.  A chunk with more headers than the parking area holds (no tracer writes one) resolves
    // them first and writes the headers direct.
    const uint32_t park_slots = n_tb + ((n_warp + 1) >> 1);
    const bool late = park_slots <= (uint32_t)kParkMax;
    const uint32_t stage_cap = late ? (uint32_t)kTxnStage - park_slots : (uint32_t)kTxnStage;
    ti_tb_rec *park_tb = reinterpret_cast<ti_tb_rec *>(stage + stage_cap);
    ti_warp_rec *park_warp = reinterpret_cast<ti_warp_rec *>(stage + stage_cap + n_tb);
    uint64_t exA[4] = {0, 0, 0, 0};                          // chain 0 / 1 prefixes: instructions, thread blocks, warps
    if (!late) {
      cta_resolve(P.chain[0], P.chain[1], chunk, aggA0, aggA1, exA, ms->lb);
      if (tid == 0) atomicAdd(&P.g->rare[1], 1u);
      __syncthreads();                                     // Misc::lb may be written again by the next resolve
    }

    // ---------------------------------------------------------------- H structural lines
    // "#BEGIN_TB" / "#END_TB" / "thread block = x,y,z" / "warp = w" + look-ahead "insts = n".  There are few of them per
    // warp range (a lane per line would leave most lanes idle through a long dependent chain), so the warps only listed
    // them (above, Misc-side list `sllist`: offset, instruction lines before it, thread-block / warp ordinal) and the
    // CTA takes the list a thread per line; a chunk with more of them than the list holds keeps them per warp.
    {
      ti_tb_rec *tb_out = late ? park_tb : P.tb + exA[1];
      ti_warp_rec *warp_out = late ? park_warp : P.warp + exA[2];
      const uint64_t tb_room = late ? ~0ull : (P.cap_tb > exA[1] ? P.cap_tb - exA[1] : 0);
      const uint64_t warp_room = late ? ~0ull : (P.cap_warp > exA[2] ? P.cap_warp - exA[2] : 0);
      const uint64_t inst_base = late ? 0 : exA[0] + P.off_inst, warp_base = late ? 0 : exA[2] + P.off_warp;
      auto struct_line = [&](int o, uint32_t first_rel, uint32_t ex_tb, uint32_t ex_w) {
        const unsigned c = text[o];
        const uint64_t first_inst = inst_base + first_rel;
        SlParsed sp = parse_struct_line(text, text32, o);
        int32_t n = 0;
        if (sp.type == SL_WARP) {                     // trace_parser.cc:428-437: the next non-blank line must be "insts = n"
          int q = sp.q;
          sp.ok = sp.ok && sp.v[0] >= 0;
          while (text[q] == '\n' && q < 16 + kChunk + kSlack) ++q;
          sp.ok = sp.ok && load8(text32, (unsigned)q) == kLitInsts0;
          q += 8;
          sp.ok = sp.ok && parse_i32(text, q, n) && text[q] == '\n' && n >= 0;
        }
        if (sp.type == SL_INSTS) sp.ok = sp.ok && sp.v[0] >= 0;
        uint32_t fl = sp.ok ? 0u : (uint32_t)TI_EF_BAD_STRUCTURE;
        if (sp.ok && sp.type != SL_TB && sp.type != SL_WARP)
          atomicAdd(&cnt[kCntConst + (sp.type == SL_BEGIN ? 1 : sp.type == SL_END ? 2 : 4)], 1u);
        if (c == 't') {                               // trace_parser.cc:423-427
          ti_tb_rec r;
          r.text_off = cb - 16 + (uint64_t)o + (sp.ok ? P.off_text : 0); r.line_len = sp.ok ? (uint32_t)(sp.q - o) : 0u;
          r.x = sp.ok ? sp.v[0] : 0; r.y = sp.ok ? sp.v[1] : 0; r.z = sp.ok ? sp.v[2] : 0;
          r.first_warp = sp.ok ? (uint32_t)(warp_base + ex_w) : 0u; r.first_inst = sp.ok ? (uint32_t)first_inst : 0u;
          if (ex_tb < tb_room) tb_out[ex_tb] = r;
        } else if (c == 'w') {
          ti_warp_rec r;  // written even when malformed so the table never holds uninitialised rows
          r.warp_id = sp.ok ? (uint32_t)sp.v[0] : 0u; r.n_insts = sp.ok ? (uint32_t)n : 0u; r.first_inst = (uint32_t)first_inst; r.reserved = 0;
          if (ex_w < warp_room) warp_out[ex_w] = r;
        }
        if (line_end(nlmap32, (unsigned)o) - (unsigned)o > (unsigned)TI_MAX_LINE) fl |= TI_EF_LINE_TOO_LONG;
        note_error(ms, fl, tbase + (uint32_t)o);
      };
      if (sl_listed) {
        for (uint32_t e = tid; e < n_sl; e += kThreads) {
          const uint4 ent = reinterpret_cast<const uint4 *>(smem + oSlList)[e];
          struct_line((int)ent.x, ent.y, ent.z, ent.w);
        }
      } else {
        uint32_t tb_seen = tb0, warp_seen = warp0;
        for (uint32_t i0 = 0; i0 < n_sl_w; i0 += 32) {
          const uint32_t i = i0 + lane;
          const bool valid = i < n_sl_w;
          const int o = valid ? wl_sl[i] : 0;
          const unsigned c = valid ? text[o] : 0u;
          const uint32_t tbb = __ballot_sync(kFull, c == 't'), wb = __ballot_sync(kFull, c == 'w');
          const uint32_t ex_tb = tb_seen + __popc(tbb & lt), ex_w = warp_seen + __popc(wb & lt);
          tb_seen += __popc(tbb); warp_seen += __popc(wb);
          if (!valid) continue;
          uint32_t lo = 0, hi = n_inst_w;            // instruction lines before this one = its rank in the warp's sorted line list
          while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (wl_inst[mid] < o) lo = mid + 1; else hi = mid; }
          struct_line(o, inst0 + lo, ex_tb, ex_w);
        }
      }
    }
    TI_STAMP(6);   // H

    // ---------------------------------------------------------------- M everything but base+stride: warp per instruction
    // (skipped with its barriers when the chunk has no such instruction: the count is CTA-uniform).  It writes the
    // 256 B rows itself, so a chunk that has any resolves chain 2 now instead of one iteration later.
    uint64_t base_mem = 0;
    if (n_warp_path) {
      uint64_t exM[4];
      cta_resolve(P.chain[2], nullptr, chunk, chain_pack(n_mem, 0), 0, exM, ms->lb);
      base_mem = exM[0];
      if (tid == 0) atomicAdd(&P.g->rare[2], 1u);
      for (;;) {
        uint32_t so = 0;
        if (lane == 0) so = atomicAdd(&ms->next_mem, 1u);
        so = __shfl_sync(kFull, so, 0);
        if (so >= n_warp_path) break;
        const uint32_t item = slowq[so];
        const unsigned ow = item / kWarpInst;                // the warp whose range holds the instruction
        ti_inst_rec *rp = recs + ow * kWarpInst + m_io[item];
        const uint32_t mo = ms->pw[par][ow].mem0 + (item - ow * kWarpInst);
        const uint32_t mask = rp->mask, tmp = rp->mem_idx;
        const unsigned mode = tmp >> 16 & 3, resolve = tmp >> 18 & 1;
        unsigned op = rp->op, pack_c = rp->pack_c;
        const unsigned ds = rp->data_size, pack_b = rp->pack_b;
        __syncwarp();
        uint32_t fl = 0;
        const uint64_t addr = decode_addresses(text, dlmap32, tmp & 0xffff, line_end(nlmap32, tmp & 0xffff), mode, mask, tokpos, fl);
        fl = __reduce_or_sync(kFull, fl);              // a bad token is seen by one lane only: make it warp-uniform
        const uint64_t row = base_mem + mo;
        if (row < P.cap_mem) P.rows[row * 32 + lane] = addr;
        if (P.desc && lane == 0 && row < P.cap_mem) { ti_mem_desc dd; dd.base = 0; dd.stride = 0; dd.meta = TI_DESC_FULL_ROW; P.desc[row] = dd; }
        if (resolve) {                                 // trace_driven.cc:331-351
          const uint64_t a0 = __shfl_sync(kFull, addr, __ffs(mask) - 1);
          unsigned space = SPACE_GLOBAL, cache_op = CACHE_ALL;
          if (a0 >= P.shmem_base && a0 < P.local_base) { space = SPACE_SHARED; cache_op = CACHE_UNDEF; }
          else if (a0 >= P.local_base && a0 < P.local_base + (1u << 14)) space = SPACE_LOCAL;
          pack_c = (pack_c & 0x80) | space | cache_op << 4;
        }
        const unsigned space = pack_c & 0xf, cache_op = pack_c >> 4 & 7;
        unsigned n_txn = 0, cycles = 0;
        Txn t2[2];
        unsigned pos[2] = {~0u, ~0u}, seg = 0;
        const bool gated = (op == OP_LOAD || op == OP_STORE || op == OP_TLOAD || op == OP_TSTORE) && mask != 0;
        const unsigned is_write = (op == OP_STORE || op == OP_TSTORE || (pack_b >> 4 & 3) == 2) ? 1u : 0u;
        if (gated && !fl) {
          if (space == SPACE_SHARED || space == SPACE_SSTARR) cycles = shared_cycles(P, mask, addr);
          else if (space == SPACE_GLOBAL) {
            seg = segment_size(P, ds, cache_op);
            if (!seg) fl |= TI_EF_BAD_WIDTH;
            else if (pack_b >> 6 & 1) n_txn = coalesce_atomic(P, mask, addr, ds, seg, t2[0], pos[0], fl);
            else n_txn = coalesce_global(P, mask, addr, ds, seg, t2, pos, wscratch);
          } else if (space != SPACE_LOCAL) fl |= TI_EF_UNDEF_SPACE;
        }
        if (fl) n_txn = 0;                              // fl is warp-uniform here
        // The pool is handed out in 16 B units.  When every transaction of the instruction is one run of bytes they are
        // staged in the compact form (one unit each, ti_mem.cuh), else as finished 32 B records (two units each).
        unsigned rp0[2] = {0, 0}, rlen[2] = {0, 0};
        bool runs = true;
        if (pos[0] < n_txn) runs = txn_is_run(t2[0], rp0[0], rlen[0]);
        if (pos[1] < n_txn) runs = txn_is_run(t2[1], rp0[1], rlen[1]) && runs;
        const bool compact_form = __all_sync(kFull, runs) && n_txn != 0;
        const uint32_t n_units = compact_form ? n_txn : 2 * n_txn;
        uint32_t slot = 0;
        if (lane == 0) {
          note_error(ms, fl, rp->opc_off);
          rp->pack_c = (uint8_t)pack_c; rp->cycles = (uint8_t)cycles;
          slot = n_txn ? atomicAdd(&ms->stage_used, n_units) : 0;
          if (slot + n_units > 2 * stage_cap) slot = 0xFFFFu;   // staging full: recomputed when the chunk's transactions leave
          txn_cnt[item] = (uint8_t)n_txn; txn_slot[item] = (uint16_t)(slot == 0xFFFFu ? 0xFFFFu : slot | (compact_form ? 0x4000u : 0u));
          // the counters this instruction was left out of in W (its space is final only now)
          const unsigned act = __popc(mask);
          if ((op == OP_LOAD || op == OP_STORE || op == OP_MEMBAR || op == OP_TLOAD || op == OP_TSTORE) && act)
            atomicAdd(&cnt[kCntSpace + (pack_c & 0xf) * 2 + is_write], act);
          if (cycles) atomicAdd(&cnt[kCntConst + 3], cycles);
        }
        slot = __shfl_sync(kFull, slot, 0);
        if (slot != 0xFFFFu) {
          const unsigned type = is_write ? 4u : 0u;       // GLOBAL_ACC_W / GLOBAL_ACC_R
          uint4 *unit = reinterpret_cast<uint4 *>(stage) + slot;
          if (compact_form) {
            if (pos[0] < n_txn) unit[pos[0]] = txn_pack_run(t2[0], rp0[0], rlen[0], seg, is_write);
            if (pos[1] < n_txn) unit[pos[1]] = txn_pack_run(t2[1], rp0[1], rlen[1], seg, is_write);
          } else {
            if (pos[0] < n_txn) txn_store(reinterpret_cast<ti_txn_rec *>(unit + 2 * pos[0]), reduce_and_send(t2[0], seg, type, is_write));
            if (pos[1] < n_txn) txn_store(reinterpret_cast<ti_txn_rec *>(unit + 2 * pos[1]), reduce_and_send(t2[1], seg, type, is_write));
          }
        }
      }
      __syncthreads();
      // the transaction counts of the warp-path instructions are known now: offsets inside every warp's range again
      uint32_t seen = 0;
      for (uint32_t j0 = 0; j0 < mem_seen; j0 += 32) {
        const uint32_t j = j0 + lane;
        const uint32_t v = j < mem_seen ? (uint32_t)txn_cnt[wid * kWarpInst + j] : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
        if (j < mem_seen) txn_loc[wid * kWarpInst + j] = (uint16_t)(seen + inc - v);
        seen += __shfl_sync(kFull, inc, 31);
      }
      if (lane == 0) ms->wt[wid].n_txn = seen;
      __syncthreads();
      {
        const uint32_t t = lane < (unsigned)kWarps ? ms->wt[lane].n_txn : 0u;
        n_txn_chunk = __reduce_add_sync(kFull, t); txn0 = __reduce_add_sync(kFull, lane < wid ? t : 0u);
      }
      if (lane == 0) ms->pw[par][wid].txn0 = txn0;
    }
    const uint32_t stage_used = ms->stage_used;        // final
    TI_STAMP(7);   // M

    // ---------------------------------------------------------------- T chain 3; next ticket; chains 0 / 1 resolve
    if (tid == 0 && n_warp_path) chain_publish(P.chain[3], chunk, chain_pack(n_txn_chunk, 0));
    if (tid == kThreads - 32) ms->chunk_next = atomicAdd(&P.g->ticket, 1u);   // visible after the barriers of the resolve below
    if (late) {
      const LbPre preA = cta_lookahead(P.chain[0], P.chain[1], chunk);
      cta_resolve(P.chain[0], P.chain[1], chunk, aggA0, aggA1, exA, ms->lb, &preA);
    }
    __syncthreads();   // (the resolve of the shard's first chunk has no barrier of its own) parked headers, chunk_next visible
    // every warp is through with the text: the next chunk's bulk copy may land in the buffer
    const unsigned nbuf = TI_TEXT_BUFS == 2 ? (buf ^ 1) : 0u;
    const uint32_t chunk_next = ms->chunk_next;
    if (tid == kThreads - 32 && chunk_next < P.n_chunks) issue_bulk(chunk_next, nbuf);
    const uint64_t base_inst = exA[0], base_tb = exA[1], base_warp = exA[2];
    TI_STAMP(8);   // chains 0 / 1 resolved
    if (tid == 0) {
      uint32_t ov = 0;
      if (base_inst + n_inst > P.cap_inst) ov |= 1;
      if (base_tb + n_tb > P.cap_tb) ov |= 8;
      if (base_warp + n_warp > P.cap_warp) ov |= 16;
      if (base_inst + n_inst > 0x7ffffffeull) ov |= 32;  // 31-bit prefix fields
      if (ov) atomicOr(&P.g->overflow, ov);
      if (chunk == P.n_chunks - 1) {                   // last chunk: totals (memory instructions, transactions: F)
        P.counters[kCntNInst + 0] = base_inst + n_inst;
        P.counters[kCntNInst + 3] = base_tb + n_tb; P.counters[kCntNInst + 4] = base_warp + n_warp;
        P.counters[kCntNInst + 7] = P.len;
      }
      if (ms->err) {
        atomicOr(&P.g->error_flags, ms->err);
        atomicMin(&P.g->first_error_off, ms->first_err);
      }
      ms->next_mem = 0; ms->stage_used = 0; ms->err = 0; ms->first_err = 0xffffffffu; ms->n_slowq = 0;   // for the next chunk (published by the barrier at the loop top)
    }

    // ---------------------------------------------------------------- parked headers out
    if (late) {
      for (uint32_t i = tid; i < n_tb; i += kThreads) {
        ti_tb_rec r = park_tb[i];
        if (r.line_len) { r.first_warp += (uint32_t)(base_warp + P.off_warp); r.first_inst += (uint32_t)(base_inst + P.off_inst); }
        if (base_tb + i < P.cap_tb) P.tb[base_tb + i] = r;
      }
      for (uint32_t i = tid; i < n_warp; i += kThreads) {
        ti_warp_rec r = park_warp[i];
        r.first_inst += (uint32_t)(base_inst + P.off_inst);
        if (base_warp + i < P.cap_warp) P.warp[base_warp + i] = r;
      }
    }
    // ---------------------------------------------------------------- records out: one TMA bulk store per warp
    // (mem_idx still holds the tokenizer's scratch value; F patches it in global memory.)  Only a block that would
    // run past the output capacity is copied by the lanes, record by record.
    {
      const uint64_t at = base_inst + inst0;
      if (TI_TMA_STORE && base_inst + n_inst <= P.cap_inst) {
        if (lane == 0 && n_inst_w) {
          fence_proxy_async();                         // the records were written through the generic proxy
          bulk_s2g(P.inst + at, recs + wid * kWarpInst, n_inst_w * (unsigned)sizeof(ti_inst_rec));
        }
      } else {
        const uint4 *src = reinterpret_cast<const uint4 *>(recs + wid * kWarpInst);
        uint4 *dst = reinterpret_cast<uint4 *>(P.inst + at);
        for (uint32_t k = lane; k < 2 * n_inst_w; k += 32)
          if (at + (k >> 1) < P.cap_inst) dst[k] = src[k];
      }
    }
    if (chunk_next < P.n_chunks) fill_edges(chunk_next, nbuf);
    TI_STAMP(9);   // headers, records out

    // ---------------------------------------------------------------- F: parked for the next iteration
    // (a chunk whose transactions overflowed the staging area included: what was spilled is coalesced again there)
    if (stage_used > 2 * stage_cap && tid == 0) atomicAdd(&P.g->rare[0], 1u);    // (stage_used counts 16 B units)
    pend = true; pend_chunk = chunk; pend_par = par; pend_n_mem = n_mem; pend_n_txn = n_txn_chunk;
    pend_base_inst = base_inst; pend_have_mem = n_warp_path != 0; pend_base_mem = base_mem; pend_recompute = stage_used > 2 * stage_cap;
    TI_STAMP(10);
    if (++done_chunks % kFlushEvery == 0) { __syncthreads(); flush_counters(cnt, P.counters); }
  }
  if (pend) finish(pend_chunk, pend_par, pend_n_mem, pend_n_txn, pend_base_inst, pend_have_mem, pend_base_mem, pend_recompute, nullptr);
  if (lane == 0) bulk_store_wait_all();
#pragma unroll
  for (int d = 16; d; d >>= 1) fast_lines += __shfl_xor_sync(kFull, fast_lines, d);
  if (lane == 0 && fast_lines) atomicAdd(P.counters + kCntNInst + 13, fast_lines);
  __syncthreads();
  flush_counters(cnt, P.counters);
}

// Framing and structure check over the finished tables (the reference would trip an assert or index past
// a vector, trace_parser.cc:414-443): every "insts = n" must be followed by exactly n instruction lines,
// every "warp = w" by its "insts = n", and "#BEGIN_TB" / "thread block" / "#END_TB" lines must pair up.
// The second check is also made on the host by the entry points that bring the counters there; here it
// reaches the callers that read only the device-side status block (ti_decode_noalloc).
__global__ void ti_check_kernel(const ti_warp_rec *warp, uint64_t cap_warp, const unsigned long long *counters, Globals *g, uint64_t off_inst = 0) {
  const uint64_t n_warp = counters[kCntNInst + 4], n_inst = counters[kCntNInst + 0] + off_inst;
  const uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (uint64_t)gridDim.x * blockDim.x;
  if (t0 == 0) {
    const uint64_t n_tb = counters[kCntNInst + 3], n_begin = counters[kCntNInst + 5], n_end = counters[kCntNInst + 6], n_insts_lines = counters[kCntNInst + 9];
    if (n_begin != n_end || n_begin != n_tb || n_insts_lines != n_warp) atomicOr(&g->error_flags, (uint32_t)TI_EF_BAD_STRUCTURE);
  }
  if (n_warp > cap_warp) return;  // the table overflowed: the host decodes again with exact capacities
  if (t0 == 0 && (n_warp ? warp[0].first_inst != off_inst : n_inst != off_inst)) atomicOr(&g->error_flags, (uint32_t)TI_EF_FRAMING);
  for (uint64_t k = t0; k < n_warp; k += stride) {
    const uint64_t next = k + 1 < n_warp ? warp[k + 1].first_inst : n_inst;
    if (next - warp[k].first_inst != warp[k].n_insts) atomicOr(&g->error_flags, (uint32_t)TI_EF_FRAMING);
  }
}

// What a launch needs at zero: the look-back words it will use, the status block (first_error_off: all ones, it
// is an atomicMin target), the counter block and the CSR sentinel.  One kernel instead of five memsets: a memset
// may be queued on a copy engine behind a large transfer of another stream, which stalls the decode behind it.
__global__ void ti_reset_kernel(unsigned long long *chains, uint64_t n_words, Globals *g, unsigned long long *counters, uint32_t *txn_off) {
  const uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = t0; i < n_words; i += stride) chains[i] = 0;
  for (uint64_t i = t0; i < sizeof(ti_counters) / 8; i += stride) counters[i] = 0;
  if (t0 == 0) {
    Globals z;
    z.ticket = z.error_flags = z.overflow = 0; z.first_error_off = 0xffffffffu;
    z.rare[0] = z.rare[1] = z.rare[2] = z.rare[3] = 0;
    *g = z;
    txn_off[0] = 0;
  }
}

// A shard's status, counters and (compact results) kept-row / kept-transaction totals, written by the GPU straight
// into pinned host memory: no copy engine involved, so they are not queued behind the previous shard's results.
__global__ void ti_publish_kernel(const Globals *g, const unsigned long long *counters, const uint32_t *keep, unsigned char *host) {
  static_assert(sizeof(Globals) % 8 == 0 && sizeof(ti_counters) % 8 == 0, "publish layout");
  const unsigned t = threadIdx.x;                    // one CTA
  if (t < sizeof(Globals) / 4) reinterpret_cast<uint32_t *>(host)[t] = reinterpret_cast<const uint32_t *>(g)[t];
  for (unsigned i = t; i < sizeof(ti_counters) / 8; i += blockDim.x) reinterpret_cast<unsigned long long *>(host + sizeof(Globals))[i] = counters[i];
  if (t < 2) reinterpret_cast<uint32_t *>(host + sizeof(Globals) + sizeof(ti_counters))[t] = keep ? keep[t] : 0u;
#ifndef TI_EMU
  __threadfence_system();
#endif
}

}  // namespace ti
#endif
