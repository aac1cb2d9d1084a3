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
  StrideNote *notes = reinterpret_cast<StrideNote *>(smem + oPool);      // descriptors from the bottom of the pool
  uint16_t *inst_off = reinterpret_cast<uint16_t *>(smem + oInstOff);
  uint16_t *sl_off = reinterpret_cast<uint16_t *>(smem + oSlOff);
  uint16_t *fast_list = reinterpret_cast<uint16_t *>(smem + oFastList);
  uint16_t *slow_list = reinterpret_cast<uint16_t *>(smem + oSlowList);
  uint16_t *txn_slot = reinterpret_cast<uint16_t *>(smem + oTxnSlot);
  uint16_t *txn_loc = reinterpret_cast<uint16_t *>(smem + oTxnLoc);
  uint8_t *txn_cnt = smem + oTxnCnt;
  uint8_t *kind = smem + oKind;
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
  for (int i = tid; i < kSmCnt; i += kThreads) cnt[i] = 0;
  for (int i = tid; i < kMemoSlots * 4; i += kThreads) reinterpret_cast<unsigned long long *>(memo)[i] = 0;
  unsigned done_chunks = 0;
  unsigned long long fast_lines = 0;   // instruction lines the fast pass of P2 decoded (ti_counters.reserved[4], a path statistic)
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
  // everything it needs was left in shared memory by P6a / P6b / P7: the base+stride descriptors, the
  // warp path's staged transactions, txn_cnt / txn_loc / txn_slot and the chunk's own copy of mem_list.
  // The records left earlier (P8) with chunk-relative mem_idx; their row index is patched here.
  // `recompute` (CTA-uniform): the staging area overflowed, so the spilled warp-path instructions are
  // coalesced again from their address rows, which needs the chunk's records: same iteration only.
  auto finish = [&](uint32_t c, uint32_t n_mem, uint32_t n_txn_chunk, uint32_t n_notes, uint64_t base_inst, const uint16_t *mlist,
                    bool have_mem, uint64_t base_mem, bool recompute, const LbPre *pre) {
    if (tid == 0) bulk_store_wait_all();               // the chunk's record block is in global memory: the patch below lands on top of it
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
    // address rows of the base+stride instructions: half a warp per row, two addresses (16 B) per lane,
    // so every store instruction covers two whole 256 B rows
    for (uint32_t f0 = wid * 2; f0 < n_notes; f0 += kWarps * 2) {
      const uint32_t fo = f0 + (lane >> 4);
      if (fo < n_notes) {
        const StrideNote nt = notes[fo];
        const uint32_t meta = nt.meta;
        const uint64_t row = base_mem + (meta >> 18 & 0x3ff);
        if (!(meta >> 28 & 1) && row < P.cap_mem) {    // bit 28: the warp path took it and wrote its row
          const unsigned first = meta & 31, n = meta >> 5 & 63, l = (lane & 15) * 2;
          const uint64_t a0 = l - first < n ? nt.base + (uint64_t)(int64_t)nt.stride * (uint64_t)(l - first) : 0;
          const uint64_t a1 = l + 1 - first < n ? nt.base + (uint64_t)(int64_t)nt.stride * (uint64_t)(l + 1 - first) : 0;
          reinterpret_cast<ulonglong2 *>(P.rows + row * 32)[lane & 15] = make_ulonglong2(a0, a1);
        }
      }
    }
    // per memory instruction: CSR offset and the row index in its record
    for (uint32_t mo = tid; mo < n_mem; mo += kThreads) {
      const uint64_t row = base_mem + mo;
      if (row < P.cap_mem) P.txn_off[row] = (uint32_t)(base_txn + txn_loc[mo] + P.off_txn);
      const uint64_t ri = base_inst + mlist[mo];
      if (ri < P.cap_inst) P.inst[ri].mem_idx = (uint32_t)(row + P.off_mem);
    }
    // its transactions: four lanes per instruction.  A dense base+stride progression has every transaction in
    // closed form (lane q writes transactions q, q+4, ...); the streaming merge of other progressions and the
    // copy of what the warp path staged are split the same way where they can be.
    for (uint32_t m0 = tid >> 2; m0 < n_mem; m0 += kThreads / 4) {
      const uint32_t mo = m0, sub = tid & 3;
      const uint32_t n = txn_cnt[mo], slot = txn_slot[mo];
      const uint64_t at = base_txn + txn_loc[mo];
      if (!n || slot == 0xFFFFu || at >= P.cap_txn) continue;
      const unsigned room = (unsigned)(P.cap_txn - at < n ? P.cap_txn - at : n);
      if (slot & 0x8000u) {                            // base+stride: straight from the descriptor
        const StrideNote nt = notes[slot & 0x7fffu];
        const uint32_t meta = nt.meta;
        StrideDesc d;
        d.base = nt.base; d.stride = nt.stride; d.first = meta & 31; d.n = meta >> 5 & 63;
        const unsigned ds = 1u << (meta >> 11 & 7), seg = 32u << (meta >> 14 & 3), is_write = meta >> 16 & 1;
        if (meta >> 17 & 1) {
          for (unsigned k = sub; k < room; k += 4) {
            txn_store(P.txn + at + k, dense_txn(d, ds, seg, is_write ? 4u : 0u, is_write, k));
          }
        } else if (sub == 0) stride_emit(d, ds, seg, is_write ? 4u : 0u, is_write, P.txn + at, room);
      } else {                                         // warp path: staged in the pool
        const uint4 *src = reinterpret_cast<const uint4 *>(stage + slot);
        uint4 *dst = reinterpret_cast<uint4 *>(P.txn + at);
        for (uint32_t k = sub; k < room; k += 4) { dst[2 * k] = src[2 * k]; dst[2 * k + 1] = src[2 * k + 1]; }
      }
    }
    if (recompute) {                                   // CTA-uniform
      for (uint32_t mo = wid; mo < n_mem; mo += kWarps) {
        if (txn_slot[mo] != 0xFFFFu || !txn_cnt[mo]) continue;
        const uint64_t at = base_txn + txn_loc[mo];
        const ti_inst_rec *rp = recs + mlist[mo];
        const uint64_t row = base_mem + mo;
        const uint64_t addr = row < P.cap_mem ? __ldcg(P.rows + row * 32 + lane) : 0;   // written in P6b
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
    __syncthreads();   // the pool and the txn_* lists may be reused
  };
  bool pend = false, pend_have_mem = false;   // a chunk whose rows and transactions are still parked in shared memory (CTA-uniform)
  uint32_t pend_chunk = 0, pend_n_mem = 0, pend_n_txn = 0, pend_n_notes = 0;
  uint64_t pend_base_inst = 0, pend_base_mem = 0;
  const uint16_t *pend_mlist = nullptr;

  // prologue: first ticket, first bulk copy
  if (tid == 0) {
    mbar_init(&ms->mbar[0], 1);
    mbar_init(&ms->mbar[1], 1);
    ms->chunk_next = atomicAdd(&P.g->ticket, 1u);
    if (ms->chunk_next < P.n_chunks) issue_bulk(ms->chunk_next, 0);
    ms->next_mem = 0; ms->stage_used = 0; ms->err = 0; ms->first_err = 0xffffffffu; ms->n_tb = 0; ms->n_warp = 0; ms->n_extra = 0; ms->n_p2slow = 0;
  }
  __syncthreads();
  if (ms->chunk_next < P.n_chunks) fill_edges(ms->chunk_next, 0);

  for (unsigned iter = 0;; ++iter) {
    if (tid == 0) bulk_store_wait_read();   // the previous chunk's record block has left shared memory
    __syncthreads();  // previous chunk fully drained; edge fills of this chunk's buffer visible
    const unsigned buf = TI_TEXT_BUFS == 2 ? (iter & 1) : 0u;
    const uint32_t chunk = ms->chunk_next;
    if (chunk >= P.n_chunks) break;
    uint8_t *text = smem + oText + buf * kTextBytes;
    const uint32_t *text32 = reinterpret_cast<const uint32_t *>(text);
    uint4 *text_v = reinterpret_cast<uint4 *>(text);
    uint16_t *mem_list = reinterpret_cast<uint16_t *>(smem + oMemList) + (iter & 1) * (align16(kRecCap * 2) / 2);
    const uint64_t cb = (uint64_t)chunk * kChunk;            // text offset of the chunk
    const uint32_t tbase = (uint32_t)(cb - 16);              // text offset = tbase + smem offset (mod 2^32)
    TI_STAMP(0);   // top-of-loop bookkeeping, ticket, edge fills
    mbar_wait(&ms->mbar[buf], TI_TEXT_BUFS == 2 ? ((iter >> 1) & 1) : (iter & 1));  // P0: this chunk's bulk copy has landed
    TI_STAMP(1);

    // ---------------------------------------------------------------- P1 line index
    // A line start is a byte that follows '\n' and is not '\n' itself (blank lines
    // never become lines, trace_parser.cc:408).  The newline bitmap of the whole
    // staged region is kept (nlmap) so any thread can find the end of its line.
    // Counts are packed 2 x 16 bit: instruction lines | structural lines.
    const uint32_t lim = (uint32_t)(P.len - cb < (uint64_t)kChunk ? P.len - cb : (uint64_t)kChunk);
    for (int v = tid; v < kTextBytes / 16; v += kThreads)
      if (v == 0 || v > kChunk / 16) {
        uint32_t nl16, dl16;
        cls_masks16(text_v[v], nl16, dl16);
        nlmap[v] = (uint16_t)nl16; dlmap[v] = (uint16_t)dl16;
      }
    // Each lane takes kVecIters CONSECUTIVE 16 B vectors (48 B apart from its neighbours': the 16 B
    // loads stay bank-conflict free), so that its line starts are contiguous in file order and one
    // shuffle scan per warp orders them all.
    uint32_t starts[kVecIters], m[kVecIters];
    const uint32_t v0 = wid * kVecPerWarp + lane * kVecIters;
#pragma unroll
    for (int j = 0; j < kVecIters; ++j) {
      const uint32_t v = v0 + j;
      uint32_t dl16 = 0;
      m[j] = 0;
      if (v < (uint32_t)(kChunk / 16)) {
        cls_masks16(text_v[1 + v], m[j], dl16);
        nlmap[1 + v] = (uint16_t)m[j]; dlmap[1 + v] = (uint16_t)dl16;
      }
    }
    const uint32_t up = __shfl_up_sync(kFull, m[kVecIters - 1], 1);
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < kVecIters; ++j) {
      const uint32_t v = v0 + j, s0 = v * 16;
      const bool inside = v < (uint32_t)(kChunk / 16);
      const uint32_t prev = j ? (m[j ? j - 1 : 0] >> 15) & 1 : lane ? (up >> 15) & 1 : (uint32_t)(inside && text[16 + s0 - 1] == '\n');
      uint32_t st = ((m[j] << 1) | prev) & ~m[j] & 0xffffu;
      if (!inside || s0 >= lim) st = 0;
      else if (s0 + 16 > lim) st &= (1u << (lim - s0)) - 1;
      for (uint32_t b = st; b; b &= b - 1) {
        const unsigned c = text[16 + s0 + __ffs(b) - 1];
        if (hexval(c) < 16) { packed += 1u; st |= (b & (0u - b)) << 16; }   // high half: which starts are instruction lines
        else {
          packed += 1u << 16;
          if (c == 't') atomicAdd(&ms->n_tb, 1u);
          if (c == 'w') atomicAdd(&ms->n_warp, 1u);
        }
      }
      starts[j] = st;
    }
    uint32_t inc = packed;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
    const uint32_t excl = inc - packed;
    const uint32_t run = __shfl_sync(kFull, inc, 31);
    if (lane == 0) ms->warp_tot[wid] = run;
    __syncthreads();
    uint32_t pbase = 0, ptot = 0;
    for (unsigned k = 0; k < (unsigned)kWarps; ++k) { const uint32_t t = (uint32_t)ms->warp_tot[k]; if (k < wid) pbase += t; ptot += t; }
    uint32_t n_inst = ptot & 0xffff, n_sl = ptot >> 16, n_tb = ms->n_tb, n_warp = ms->n_warp;
    const bool dense = n_inst > (uint32_t)kRecCap || n_sl > (uint32_t)kSlCap;
    if (dense) {  // more lines than any tracer-written file can hold: flag, publish an empty chunk
      if (tid == 0) note_error(ms, TI_EF_DENSE_CHUNK, (uint32_t)cb);
      n_inst = n_sl = n_tb = n_warp = 0;
    } else {
      uint32_t at = pbase + excl;
#pragma unroll
      for (int j = 0; j < kVecIters; ++j) {
        const uint32_t s0 = (v0 + j) * 16;
        for (uint32_t b = starts[j] & 0xffffu; b; b &= b - 1) {
          const uint32_t o = 16 + s0 + __ffs(b) - 1;
          if ((starts[j] >> 16) & b & (0u - b)) { inst_off[at & 0xffff] = (uint16_t)o; at += 1; }
          else { sl_off[at >> 16] = (uint16_t)o; at += 1u << 16; }
        }
      }
    }
    __syncthreads();


    // chains 0 / 1: everything they carry is known as soon as the lines are classified
    const unsigned long long aggA0 = chain_pack(n_inst, n_tb), aggA1 = chain_pack(n_warp, 0);
    if (tid == 0) { chain_publish(P.chain[0], chunk, aggA0); chain_publish(P.chain[1], chunk, aggA1); }
    TI_STAMP(2);   // P1
    // ---------------------------------------------------------------- P2 tokenize
    // Fast pass (ti_fast.cuh): tracer-shaped lines whose opcode string is in the CTA's memo.  Whatever
    // it declines goes on a list (in the not yet used slow_list area) for the general parser below,
    // which also fills the memo.  The list is empty for all but a CTA's first chunk or two.
    const bool all_general = P.trace_version < 3;          // four leading decimals per line: general parser only
    uint16_t *p2slow = slow_list;
    if (!all_general) {
      for (uint32_t io = tid; io < n_inst; io += kThreads) {
        uint4 rlo, rhi;
        unsigned kd;
        const unsigned o = inst_off[io];
        const unsigned le = line_end(nlmap32, o);
        if (le - o <= (unsigned)TI_MAX_LINE && parse_inst_fast(text, dlmap32, memo, o, le, P, tbase + (uint32_t)P.off_text, rlo, rhi, kd)) {
          uint4 *dst = reinterpret_cast<uint4 *>(recs + io);
          dst[0] = rlo; dst[1] = rhi;
          kind[io] = (uint8_t)kd;
        } else {
          p2slow[atomicAdd(&ms->n_p2slow, 1u)] = (uint16_t)io;
        }
      }
      __syncthreads();
    }
    const uint32_t n_general = all_general ? n_inst : ms->n_p2slow;
    fast_lines += n_inst - n_general;
    if (n_general) {                                       // CTA-uniform
      for (uint32_t k = tid; k < n_general; k += kThreads) {
        const uint32_t io = all_general ? k : (uint32_t)p2slow[k];
        ti_inst_rec r;
        unsigned kd;
        uint64_t facts;
        const unsigned o = inst_off[io];
        const unsigned le = line_end(nlmap32, o);
        uint32_t fl = parse_inst_line(text, dlmap32, (int)o, le, P, tbase + (uint32_t)P.off_text, r, kd, facts);
        if (le - o > (unsigned)TI_MAX_LINE) fl |= TI_EF_LINE_TOO_LONG;
        if (fl) { note_error(ms, fl, tbase + o); kd = 0; r.mem_idx = 0xFFFFFFFFu; r.pack_c &= 0x7f; }
        else if (facts) {
          uint64_t k0, k1, k2;
          memo_key(text32, r.opc_off - (tbase + (uint32_t)P.off_text), r.opcode_len, k0, k1, k2);
          memo_insert(memo, k0, k1, k2, facts);
        }
        const uint4 *src = reinterpret_cast<const uint4 *>(&r);
        uint4 *dst = reinterpret_cast<uint4 *>(recs + io);
        dst[0] = src[0]; dst[1] = src[1];
        kind[io] = (uint8_t)kd;
      }
      __syncthreads();
    }


    TI_STAMP(3);   // P2
    // ---------------------------------------------------------------- P3 memory-instruction ordinals
    // low half: all memory instructions; high half: those taking the thread-per-instruction path
    const bool stride_path = P.mem_warp_parts == 1;
    LbPre preF = {kValid, kValid};                           // first round of the previous chunk's look-back, a phase ahead
    if (pend) preF = pend_have_mem ? cta_lookahead(P.chain[3], nullptr, pend_chunk) : cta_lookahead(P.chain[2], P.chain[3], pend_chunk);
    const uint32_t p3 = block_scan(
        n_inst, ms->warp_tot,
        [&](uint32_t i) { const unsigned kd = kind[i]; return (uint32_t)((kd & 1) | ((stride_path && (kd & 2)) ? 0x10000u : 0u)); },
        [&](uint32_t i, uint32_t ex, uint32_t v) {
          if (!v) return;
          const uint32_t mo = ex & 0xffff, fo = ex >> 16;
          mem_list[mo] = (uint16_t)i;
          if (v >> 16) fast_list[fo] = (uint16_t)mo; else slow_list[mo - fo] = (uint16_t)mo;
        });
    const uint32_t n_mem = p3 & 0xffff, n_fast = p3 >> 16, n_slow = n_mem - n_fast;
    if (tid == 0) chain_publish(P.chain[2], chunk, chain_pack(n_mem, 0));
    TI_STAMP(4);   // P3

    // ---------------------------------------------------------------- F' the previous chunk's rows and transactions
    // (before anything of this chunk goes into the pool or the txn_* lists)
    if (pend) {
      finish(pend_chunk, pend_n_mem, pend_n_txn, pend_n_notes, pend_base_inst, pend_mlist, pend_have_mem, pend_base_mem, false, &preF);
      pend = false;
    }
    TI_STAMP(5);   // F'

    // ---------------------------------------------------------------- pool layout of this chunk
    // Everything up to P7 works with chunk-relative positions: thread-block / warp headers are parked at
    // the top of the pool.  A chunk with more headers than the parking area holds (no tracer writes one)
    // resolves chains 0 / 1 first and writes them direct.
    const uint32_t n_notes = n_fast < (uint32_t)kNoteCap ? n_fast : (uint32_t)kNoteCap;
    const uint32_t park_slots = n_tb + ((n_warp + 1) >> 1);
    const bool late = park_slots <= (uint32_t)kParkMax;
    const uint32_t stage_lo = (n_notes + 1) >> 1;            // pool slots [stage_lo, stage_hi) stage the warp path's transactions
    const uint32_t stage_hi = late ? (uint32_t)kTxnStage - park_slots : (uint32_t)kTxnStage;
    const uint32_t stage_cap = stage_hi - stage_lo;
    ti_tb_rec *park_tb = reinterpret_cast<ti_tb_rec *>(stage + stage_hi);
    ti_warp_rec *park_warp = reinterpret_cast<ti_warp_rec *>(stage + stage_hi + n_tb);
    uint64_t exA[4] = {0, 0, 0, 0};                          // chain 0 / 1 prefixes: instructions, thread blocks, warps
    if (!late) {
      cta_resolve(P.chain[0], P.chain[1], chunk, aggA0, aggA1, exA, ms->lb);
      if (tid == 0) atomicAdd(&P.g->rare[1], 1u);
    }

    // ---------------------------------------------------------------- P5 structural lines
    // ordinals of "thread block" (low half) and "warp" (high half) lines among the structural lines
    ti_tb_rec *tb_out = late ? park_tb : P.tb + exA[1];
    ti_warp_rec *warp_out = late ? park_warp : P.warp + exA[2];
    const uint64_t tb_room = late ? ~0ull : (P.cap_tb > exA[1] ? P.cap_tb - exA[1] : 0);
    const uint64_t warp_room = late ? ~0ull : (P.cap_warp > exA[2] ? P.cap_warp - exA[2] : 0);
    const uint64_t inst0 = late ? 0 : exA[0] + P.off_inst, warp0 = late ? 0 : exA[2] + P.off_warp;
    auto struct_val = [&](uint32_t i) { const unsigned c = text[sl_off[i]]; return (uint32_t)(c == 't' ? 1u : c == 'w' ? 0x10000u : 0u); };
    auto struct_line = [&](uint32_t i, uint32_t ex, uint32_t) {
      const int o = sl_off[i];
      const unsigned c = text[o];
      // instruction lines before this one = its rank in the sorted inst_off list
      uint32_t lo = 0, hi = n_inst;
      while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (inst_off[mid] < o) lo = mid + 1; else hi = mid; }
      const uint64_t first_inst = inst0 + lo;
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
        r.first_warp = sp.ok ? (uint32_t)(warp0 + (ex >> 16)) : 0u; r.first_inst = sp.ok ? (uint32_t)first_inst : 0u;
        if ((ex & 0xffff) < tb_room) tb_out[ex & 0xffff] = r;
      } else if (c == 'w') {
        ti_warp_rec r;  // written even when malformed so the table never holds uninitialised rows
        r.warp_id = sp.ok ? (uint32_t)sp.v[0] : 0u; r.n_insts = sp.ok ? (uint32_t)n : 0u; r.first_inst = (uint32_t)first_inst; r.reserved = 0;
        if ((ex >> 16) < warp_room) warp_out[ex >> 16] = r;
      }
      if (line_end(nlmap32, (unsigned)o) - (unsigned)o > (unsigned)TI_MAX_LINE) fl |= TI_EF_LINE_TOO_LONG;
      note_error(ms, fl, tbase + (uint32_t)o);
    };

    // The structural lines are few (tens per chunk) but a long dependent chain: normally the last warp
    // takes them alone, beside P6a on the other warps (same barrier-delimited phase, below).  Many of
    // them, or headers that cannot be parked, go CTA-wide here first.
    const bool p5_beside = kWarps > 1 && late && n_sl <= (uint32_t)kSlSplitMax;
    if (!p5_beside) block_scan(n_sl, ms->warp_tot, struct_val, struct_line);
    TI_STAMP(6);   // P5 (CTA-wide variant) + early resolve

    // ---------------------------------------------------------------- P6a base+stride: thread per instruction
    // Parses base and stride, counts the transactions (closed form for dense progressions) and leaves a
    // 16 B descriptor in the pool; the address row and the transactions themselves are written from it
    // when the chunk's rows leave (F).
    const uint32_t p6a_threads = p5_beside ? (uint32_t)kThreads - 32u : (uint32_t)kThreads;
    if (p5_beside && wid == kWarps - 1) {              // P5, one warp: 32 lines per round, ordinals by shuffle scan
      uint32_t run = 0;
      for (uint32_t i0 = 0; i0 < n_sl; i0 += 32) {
        const uint32_t i = i0 + lane;
        const uint32_t v = i < n_sl ? struct_val(i) : 0u;
        uint32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t up = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += up; }
        if (i < n_sl) struct_line(i, run + inc - v, v);
        run += __shfl_sync(kFull, inc, 31);
      }
    }
    for (uint32_t fo = tid; fo < n_fast && tid < p6a_threads; fo += p6a_threads) {
      const uint32_t mo = fast_list[fo];
      ti_inst_rec *rp = recs + mem_list[mo];
      const uint32_t mask = rp->mask, tmp = rp->mem_idx;
      const unsigned ds = rp->data_size, op = rp->op;
      StrideDesc d;
      uint32_t fl = 0;
      if (fo >= n_notes || !stride_parse_fast(text, dlmap32, tmp & 0xffff, line_end(nlmap32, tmp & 0xffff), mask, ds ? ds : 1, d, fl)) {
        // no room for a descriptor, or the progression wraps around 2^64 (not monotone): warp path
        slow_list[n_slow + atomicAdd(&ms->n_extra, 1u)] = (uint16_t)mo;
        if (fo < n_notes) notes[fo].meta = 1u << 28; else atomicAdd(&P.g->rare[3], 1u);
        continue;
      }
      const unsigned seg = segment_size(P, ds, rp->pack_c >> 4 & 7);
      if (!fl && mask && !seg) fl = TI_EF_BAD_WIDTH;
      if (fl) { note_error(ms, fl, rp->opc_off); d.n = 0; }
      const bool dense = !fl && mask && stride_is_dense(d, ds);
      const unsigned n_txn = (fl || !mask) ? 0u : dense ? dense_count(d, ds, seg) : stride_count(d, ds, seg);
      const unsigned is_write = op == OP_STORE ? 1u : 0u;
      StrideNote nt;
      nt.base = d.base; nt.stride = (int32_t)d.stride;
      nt.meta = d.first | d.n << 5 | (unsigned)(__ffs(ds ? ds : 1) - 1) << 11 | (seg == 128 ? 2u : seg == 64 ? 1u : 0u) << 14 | is_write << 16 |
                (dense ? 1u : 0u) << 17 | mo << 18;
      notes[fo] = nt;
      txn_cnt[mo] = (uint8_t)n_txn; txn_slot[mo] = (uint16_t)(0x8000u | fo);
      rp->mem_idx = mo;                                // chunk-relative; F patches the record in global memory
    }
    __syncthreads();
    const uint32_t n_warp_path = n_slow + ms->n_extra;
    TI_STAMP(7);   // P6a

    // ---------------------------------------------------------------- P6b everything else: warp per instruction
    // (skipped with its barrier when the chunk has no such instruction: the count is CTA-uniform).  It writes
    // the 256 B rows itself, so a chunk that has any resolves chain 2 now instead of one iteration later.
    uint64_t base_mem = 0;
    if (n_warp_path) {
      uint64_t exM[4];
      cta_resolve(P.chain[2], nullptr, chunk, chain_pack(n_mem, 0), 0, exM, ms->lb);
      base_mem = exM[0];
      if (tid == 0) atomicAdd(&P.g->rare[2], 1u);
    }
    TI_STAMP(8);   // chain 2 resolved early
    if (n_warp_path) for (;;) {
      uint32_t so = 0;
      if (lane == 0) so = atomicAdd(&ms->next_mem, 1u);
      so = __shfl_sync(kFull, so, 0);
      if (so >= n_warp_path) break;
      const uint32_t mo = slow_list[so];
      const uint32_t io = mem_list[mo];
      ti_inst_rec *rp = recs + io;
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
      uint32_t slot = 0;
      if (lane == 0) {
        note_error(ms, fl, rp->opc_off);
        rp->mem_idx = mo; rp->pack_c = (uint8_t)pack_c; rp->cycles = (uint8_t)cycles;
        slot = n_txn ? atomicAdd(&ms->stage_used, n_txn) : 0;
        slot = slot + n_txn > stage_cap ? 0xFFFFu : slot + stage_lo;   // staging full: recomputed when the chunk's transactions leave
        txn_cnt[mo] = (uint8_t)n_txn; txn_slot[mo] = (uint16_t)slot;   // 0xFFFF: recomputed when the chunk's transactions leave
      }
      slot = __shfl_sync(kFull, slot, 0);
      if (slot != 0xFFFFu) {
        const unsigned type = is_write ? 4u : 0u;       // GLOBAL_ACC_W / GLOBAL_ACC_R
        if (pos[0] < n_txn) txn_store(stage + slot + pos[0], reduce_and_send(t2[0], seg, type, is_write));
        if (pos[1] < n_txn) txn_store(stage + slot + pos[1], reduce_and_send(t2[1], seg, type, is_write));
      }
    }
    if (n_warp_path) __syncthreads();
    TI_STAMP(9);   // P6b

    // ---------------------------------------------------------------- P7 transaction offsets; chain 3; chains 0 / 1 resolve
    LbPre preA = {kValid, kValid};
    if (late) preA = cta_lookahead(P.chain[0], P.chain[1], chunk);
    const uint32_t n_txn_chunk = block_scan(
        n_mem, ms->warp_tot, [&](uint32_t i) { return (uint32_t)txn_cnt[i]; },
        [&](uint32_t i, uint32_t ex, uint32_t) { txn_loc[i] = (uint16_t)ex; }, false);   // txn_loc is next read after later barriers
    if (tid == 0) chain_publish(P.chain[3], chunk, chain_pack(n_txn_chunk, 0));
    const uint32_t stage_used = ms->stage_used;        // final since P6b
    TI_STAMP(10);  // P7 scan + publish
    if (late) cta_resolve(P.chain[0], P.chain[1], chunk, aggA0, aggA1, exA, ms->lb, &preA);
    const uint64_t base_inst = exA[0], base_tb = exA[1], base_warp = exA[2];
    TI_STAMP(11);  // chains 0 / 1 resolved
    // Next ticket + prefetch of that chunk's text (the text is dead after P6b, so a single buffer takes the
    // prefetch as well as a second one).  Tickets must be taken in the order the chunks really start, or a
    // chunk's predecessor could begin after the chunk itself and every look-back would wait for it: so
    // after the last wait of the iteration.  A lane of the last warp waits for the atomic's round trip.
    const unsigned nbuf = TI_TEXT_BUFS == 2 ? (buf ^ 1) : 0u;
    if (tid == kThreads - 32) {
      const uint32_t nxt = atomicAdd(&P.g->ticket, 1u);
      ms->chunk_next = nxt;
      if (nxt < P.n_chunks) issue_bulk(nxt, nbuf);
    }
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
    }

    // ---------------------------------------------------------------- P6c parked headers out
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

    // ---------------------------------------------------------------- P8 records out + counters
    // The record block is contiguous in smem and in HBM: one TMA bulk store by one thread (below, after
    // the barrier).  mem_idx is still the chunk-relative ordinal; F patches it in global memory.  Only a
    // block that would run past the output capacity is copied by the threads, record by record.
    const bool recs_bulk = TI_TMA_STORE && base_inst + n_inst <= P.cap_inst;
    if (!recs_bulk) {
      const uint4 *src = reinterpret_cast<const uint4 *>(recs);
      uint4 *dst = reinterpret_cast<uint4 *>(P.inst + base_inst);
      for (uint32_t k = tid; k < 2 * n_inst; k += kThreads)
        if (base_inst + (k >> 1) < P.cap_inst) dst[k] = src[k];
    }
    // counters: instructions with equal (sp_op, active lanes, op, space, store, const) share one atomic
    for (uint32_t i0 = wid * 32; i0 < n_inst; i0 += kThreads) {
      const uint32_t io = i0 + lane;
      const bool valid = io < n_inst;
      const ti_inst_rec *rp = recs + (valid ? io : 0);
      const unsigned op = rp->op, sp = rp->pack_a & 0x1f, act = __popc(rp->mask);
      const bool ldst = op == OP_LOAD || op == OP_STORE || op == OP_MEMBAR || op == OP_TLOAD || op == OP_TSTORE;
      const unsigned st = (op == OP_STORE || op == OP_TSTORE || (rp->pack_b >> 4 & 3) == 2) ? 1u : 0u;
      const unsigned sidx = (rp->pack_c & 0xf) * 2 + st, cc = rp->pack_b >> 7;
      const uint32_t key = valid ? (sp * 33 + act) | fold_op(op) << 10 | (ldst ? 1u : 0u) << 15 | sidx << 16 | cc << 21 : 0xffffffffu;
      const uint32_t grp = __match_any_sync(kFull, key);
      if (valid && (unsigned)(__ffs(grp) - 1) == lane) {
        const uint32_t k = __popc(grp);
        atomicAdd(&cnt[kCntHist + sp * 33 + act], k);
        atomicAdd(&cnt[kCntOp + fold_op(op)], k);
        if (ldst && act) atomicAdd(&cnt[kCntSpace + sidx], act * k);
        if (cc) atomicAdd(&cnt[kCntConst], k);
      }
      if (valid && rp->cycles) atomicAdd(&cnt[kCntConst + 3], (uint32_t)rp->cycles);
    }
    if (tid == 0 && ms->err) {
      atomicOr(&P.g->error_flags, ms->err);
      atomicMin(&P.g->first_error_off, ms->first_err);
    }

    __syncthreads();   // chunk_next visible; nobody reads the per-chunk scalars any more; records final
    if (tid == 0 && recs_bulk && n_inst) {
      fence_proxy_async();                             // the records were written through the generic proxy
      bulk_s2g(P.inst + base_inst, recs, n_inst * (unsigned)sizeof(ti_inst_rec));
    }
    if (tid == 0) {    // reset them for the next chunk here: the top-of-loop barrier publishes it
      ms->next_mem = 0; ms->stage_used = 0; ms->err = 0; ms->first_err = 0xffffffffu; ms->n_tb = 0; ms->n_warp = 0; ms->n_extra = 0; ms->n_p2slow = 0;
    }
    if (ms->chunk_next < P.n_chunks) fill_edges(ms->chunk_next, nbuf);
    TI_STAMP(12);  // P6c + P8

    // ---------------------------------------------------------------- F: now, or parked for the next iteration
    // The pool, txn_cnt / txn_slot / txn_loc and this iteration's mem_list stay untouched until the next
    // chunk's P5, so chains 2 / 3 can resolve then.  Not when staging overflowed: the recompute needs this
    // chunk's records, which the next chunk's P2 overwrites.
    if (stage_used > stage_cap) {
      if (tid == 0) atomicAdd(&P.g->rare[0], 1u);
      finish(chunk, n_mem, n_txn_chunk, n_notes, base_inst, mem_list, n_warp_path != 0, base_mem, true, nullptr);
    } else {
      pend = true; pend_chunk = chunk; pend_n_mem = n_mem; pend_n_txn = n_txn_chunk; pend_n_notes = n_notes;
      pend_base_inst = base_inst; pend_mlist = mem_list; pend_have_mem = n_warp_path != 0; pend_base_mem = base_mem;
    }
    TI_STAMP(13);  // F (staging overflow only)
    if (++done_chunks % kFlushEvery == 0) { __syncthreads(); flush_counters(cnt, P.counters); }
  }
  if (pend) finish(pend_chunk, pend_n_mem, pend_n_txn, pend_n_notes, pend_base_inst, pend_mlist, pend_have_mem, pend_base_mem, false, nullptr);
  if (tid == 0) bulk_store_wait_all();
  if (tid == 0 && fast_lines) atomicAdd(P.counters + kCntNInst + 13, fast_lines);
  __syncthreads();
  flush_counters(cnt, P.counters);
}

// Framing check over the finished warp-header table (the reference would trip an
// assert or index past a vector, trace_parser.cc:435-443): every "insts = n" must be
// followed by exactly n instruction lines.
__global__ void ti_check_kernel(const ti_warp_rec *warp, uint64_t cap_warp, const unsigned long long *counters, Globals *g, uint64_t off_inst = 0) {
  const uint64_t n_warp = counters[kCntNInst + 4], n_inst = counters[kCntNInst + 0] + off_inst;
  if (n_warp > cap_warp) return;  // the table overflowed: the host decodes again with exact capacities
  const uint64_t t0 = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (uint64_t)gridDim.x * blockDim.x;
  if (t0 == 0 && (n_warp ? warp[0].first_inst != off_inst : n_inst != off_inst)) atomicOr(&g->error_flags, (uint32_t)TI_EF_FRAMING);
  for (uint64_t k = t0; k < n_warp; k += stride) {
    const uint64_t next = k + 1 < n_warp ? warp[k + 1].first_inst : n_inst;
    if (next - warp[k].first_inst != warp[k].n_insts) atomicOr(&g->error_flags, (uint32_t)TI_EF_FRAMING);
  }
}

}  // namespace ti
#endif
