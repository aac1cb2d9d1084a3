// The fused trace-ingest kernel (see ti_device.cuh for the data layout, DESIGN.md 3 for the plan).
//
// A chunk is cut into one byte range per warp; a warp owns the lines that START in its range and takes
// them through tokenizing, base+stride parsing and transaction counting on its own, so that the CTA
// meets at a barrier only a few times per chunk:
//   P0  stage-in    TMA bulk copy of [chunk-16, chunk+CHUNK+slack) -> smem, prefetched one chunk ahead
//   P1  classify    newline + delimiter bitmaps of the staged text (SWAR), all threads          | barrier
//   W   per warp    line starts of the range -> per-warp line lists; a lane per instruction line:
//                   fast tokenizer (general parser for what it declines) -> 32 B record in the warp's
//                   record region; base+stride lines -> 16 B descriptor + transaction count; counters;
//                   warp totals                                                                 | barrier
//   S   prefixes    per-warp offsets inside the chunk from the eight totals; publish chains 0 / 1 / 2
//   F'  previous    the PREVIOUS chunk's address rows and transactions leave (chains 2 / 3 resolve)
//   H   headers     "thread block" / "warp"+"insts" / "#BEGIN_TB" / "#END_TB" lines, per warp, parked
//   M   memory      (only chunks that have any) warp-per-instruction: list / delta addresses -> 256 B
//                   row, coalescer -> smem staging; needs chain 2 now                            | barrier
//   T   publish chain 3; next ticket + prefetch; resolve chains 0 / 1; parked headers out; every warp
//       sends its record region with one TMA bulk store
#ifndef TI_KERNELS_CUH
#define TI_KERNELS_CUH

#include "ti_device.cuh"
#include "ti_mem.cuh"
#include "ti_parse.cuh"
#include "ti_fast.cuh"

#ifdef TI_EMU
#define TI_DYN_SMEM(p) unsigned char *p = emu::cur_smem()
static inline unsigned __vcmpeq4(unsigned a, unsigned b) {
  unsigned r = 0;
  for (int i = 0; i < 4; ++i)
    if (((a >> (8 * i)) & 0xff) == ((b >> (8 * i)) & 0xff)) r |= 0xffu << (8 * i);
  return r;
}
#else
#define TI_DYN_SMEM(p) extern __shared__ __align__(16) unsigned char ti_smem_raw_[]; unsigned char *p = ti_smem_raw_
#endif

namespace ti {

// ------------------------------------------------------------ shared-memory plan
struct LbWarp { uint32_t L[2]; uint32_t s[4]; unsigned long long inc[2]; };  // see cta_resolve
// what a warp found in its byte range / where its results start inside the chunk
struct WarpTot { uint16_t n_inst, n_sl, n_tb, n_warp, n_mem, n_notes; uint32_t n_txn; };
struct WarpPre { uint16_t inst0, mem0, n_mem, n_notes; uint32_t txn0; };   // kept for the iteration in which the chunk's rows leave
struct Misc {
  WarpTot wt[kWarps];
  WarpPre pw[2][kWarps];
  uint32_t next_mem, stage_used, err, first_err, n_slowq, chunk_next, n_general;
  uint64_t mbar[2];  // one mbarrier per text buffer (completion of the bulk copy)
  LbWarp lb[kWarps];  // per-warp partial results of the CTA-wide look-back
};
constexpr int align16(int x) { return (x + 15) & ~15; }
constexpr int kRange = kChunk / kWarps;                      // bytes of the chunk a warp owns
constexpr int kLaneBits = kRange / 32;                       // ... and a lane scans for line starts
constexpr int kWarpInst = kRange / 24;                       // instruction lines a range can hold (no tracer line is shorter than 24 bytes)
constexpr int kWarpSl = kRange / 8;                          // structural lines ("#END_TB\n")
constexpr int kItems = kWarps * kWarpInst;                   // memory-instruction slots of a chunk: warp w owns [w * kWarpInst, ...)
constexpr int kNotesPerWarp = 32;                            // base+stride descriptors per warp and chunk (more: warp path)
constexpr int kParkMax = 128;                                // pool slots (32 B) lent to parked headers
static_assert(kChunk % (kWarps * 32) == 0 && kLaneBits <= 64 && kLaneBits >= 16 && kRange % 16 == 0, "range geometry");
static_assert(kWarpInst <= 255 && kItems <= 1024, "local line indices are 8 bit, item ids 10 bit");
constexpr int oText = 0;                                     // text buffer(s): the chunk being decoded (+ the prefetched next one)
constexpr int oRecs = oText + TI_TEXT_BUFS * kTextBytes;     // ti_inst_rec[kWarps][kWarpInst]
constexpr int oPool = oRecs + kItems * 32;                   // 32 B slots: warp-path transactions from the bottom, parked headers at the top
constexpr int oNotes = oPool + kTxnStage * 32;               // StrideNote[2][kWarps][kNotesPerWarp], by iteration parity
constexpr int oWlInst = oNotes + 2 * kWarps * kNotesPerWarp * 16;   // u16[kWarps][kWarpInst] smem offset of each instruction line
constexpr int oWlSl = oWlInst + align16(kItems * 2);         // u16[kWarps][kWarpSl]   smem offset of each structural line
constexpr int oTxnSlot = oWlSl + align16(kWarps * kWarpSl * 2);  // u16[2][kItems] 0x8000 | descriptor index, or staging slot, or 0xFFFF = recompute later
constexpr int oTxnLoc = oTxnSlot + 2 * align16(kItems * 2);  // u16[2][kItems] exclusive txn offset inside the warp's range
constexpr int oTxnCnt = oTxnLoc + 2 * align16(kItems * 2);   // u8 [2][kItems]
constexpr int oMIo = oTxnCnt + 2 * align16(kItems);          // u8 [2][kItems] index of the instruction among the warp's records
constexpr int oSlowQ = oMIo + 2 * align16(kItems);           // u16[kItems]    items waiting for the warp path
constexpr int kSlList = 128;                                 // structural lines of a chunk the CTA-wide pass takes (more: per warp)
constexpr int oSlList = oSlowQ + align16(kItems * 2);        // uint4[kSlList] offset | instruction lines before | thread-block ordinal | warp ordinal
constexpr int oCnt = oSlList + kSlList * 16;                 // u32[kSmCnt]    counter mirror
constexpr int kWarpScratch = 32 * 6 * 4;                     // per warp: token offsets (decode), then B-group merge slots (coalesce)
constexpr int oTok = oCnt + align16(kSmCnt * 4);
constexpr int oNlMap = oTok + kWarps * kWarpScratch;      // u16[kTextBytes/16] newline bitmap of the staged text
constexpr int oDlMap = oNlMap + align16(kTextBytes / 16 * 2 + 4);  // u16[kTextBytes/16] delimiter bitmap (' ', 0x09..0x0d)
constexpr int oMemo = oDlMap + align16(kTextBytes / 16 * 2 + 4);   // MemoEnt[kMemoSlots]: opcode string -> facts, kept for the CTA's life
constexpr int oMisc = oMemo + kMemoSlots * (int)sizeof(MemoEnt);
constexpr int kSmemBytes = oMisc + align16((int)sizeof(Misc));
static_assert(kTextBytes % 16 == 0 && kChunk % 16 == 0 && kThreads % 32 == 0 && kWarps <= 32, "chunk geometry");
static_assert(kTextBytes < 65536, "smem offsets are 16 bit");
static_assert(TI_MIN_BLOCKS * (kSmemBytes + 1024) <= 233472, "TI_MIN_BLOCKS CTAs no longer fit into an SM's 228 KB of shared memory");
// 16 B descriptor of a base+stride instruction: enough to write its address row and its transactions later
//   meta: [4:0] first lane  [10:5] run length (0 = no lane carries an address)  [13:11] log2(data size)
//         [15:14] segment (0: 32 B, 1: 64 B, 2: 128 B)  [16] write  [17] dense progression  [27:18] item (memory-instruction slot)
//         [28] void: the warp path took the instruction after all
struct StrideNote { uint64_t base; int32_t stride; uint32_t meta; };
static_assert(kParkMax < kTxnStage && oPool % 16 == 0 && oNotes % 16 == 0 && sizeof(StrideNote) == 16, "pool tenants");

// ---- TMA 1-D bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier, used to prefetch the
// next chunk's text while the current one is decoded.  Under the emulator the copy is a
// synchronous memcpy and the barrier a no-op, so the control flow stays identical.
#ifdef TI_EMU
TI_D void mbar_init(uint64_t *, unsigned) {}
TI_D void mbar_expect_tx(uint64_t *, unsigned) {}
TI_D void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *) { memcpy(dst, src, bytes); }
TI_D void mbar_wait(uint64_t *, unsigned) {}
TI_D void fence_proxy_async() {}
TI_D void bulk_s2g(void *dst, const void *src, unsigned bytes) { memcpy(dst, src, bytes); }
TI_D void bulk_store_wait_read() {}
TI_D void bulk_store_wait_all() {}
#else
TI_D uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
TI_D void mbar_init(uint64_t *b, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
TI_D void mbar_expect_tx(uint64_t *b, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
TI_D void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(b))
               : "memory");
}
TI_D void mbar_wait(uint64_t *b, unsigned parity) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok)
                 : "r"(smem_u32(b)), "r"(parity)
                 : "memory");
}
TI_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// shared -> global bulk copy (TMA store, SASS UBLKCP.S2G); completion tracked by the thread's bulk group
TI_D void bulk_s2g(void *dst, const void *src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
TI_D void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }   // the source may be overwritten
TI_D void bulk_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }          // the writes have been performed
#endif


TI_D uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
  for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
  return v;
}

// ---- chained scans with decoupled look-back (single pass over the text).  A chunk publishes its
// aggregate as early as it knows it and resolves -- sums the aggregates of the predecessors that have
// not resolved yet plus the nearest published inclusive prefix -- as late as it can: a chunk resolves
// only once EVERY chunk before it has published, i.e. at the running maximum of ~450 in-flight
// predecessors' publish times, so everything that does not need the prefix sits between the two.
// Chunks are handed out in order by the ticket, so every predecessor is already running.
TI_D unsigned long long ld_vol64(const unsigned long long *p) { return *reinterpret_cast<const volatile unsigned long long *>(p); }
TI_D void st_vol64(unsigned long long *p, unsigned long long v) { *reinterpret_cast<volatile unsigned long long *>(p) = v; }
TI_HD unsigned long long chain_pack(uint32_t f0, uint32_t f1) { return (unsigned long long)f1 << 31 | f0; }

TI_D void chain_publish(unsigned long long *ch, uint32_t c, unsigned long long payload) {   // one thread
  st_vol64(&ch[c], kAggValid | payload);
}

// CTA-wide resolve of two chains at once (Y may be null): thread t inspects chunk c-1-t, so one round
// covers kThreads predecessors (the persistent CTAs run phase-locked, so the nearest inclusive prefix is
// typically ~100 chunks back: a single warp needed several dependent rounds of L2 round trips).
// excl[0..1] = fields of X, excl[2..3] = fields of Y.  Publishes the inclusive prefixes.
// The first round's loads can be issued ahead of time (cta_lookahead, a phase earlier) so that their L2
// round trip is over when the resolve starts; words that were not ready then are simply polled again.
struct LbPre { unsigned long long x, y; };
TI_D LbPre cta_lookahead(const unsigned long long *X, const unsigned long long *Y, uint32_t c) {
  LbPre p = {kValid, kValid};
  const int64_t j = (int64_t)c - 1 - (int64_t)threadIdx.x;
  if (j >= 0) { p.x = ld_vol64(&X[j]); if (Y) p.y = ld_vol64(&Y[j]); }
  return p;
}
TI_D void cta_resolve(unsigned long long *X, unsigned long long *Y, uint32_t c, unsigned long long aggX, unsigned long long aggY,
                      uint64_t (&excl)[4], LbWarp *lb, const LbPre *pre = nullptr) {
  const unsigned lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  excl[0] = excl[1] = excl[2] = excl[3] = 0;
#ifdef TI_EXP_NOWAIT  // tuning experiment only (wrong output): what the kernel costs with no look-back at all
  return;
#endif
  bool found_x = false, found_y = Y == nullptr;
  for (int64_t idx = (int64_t)c - 1; idx >= 0; idx -= kThreads) {
    const int64_t j = idx - (int64_t)threadIdx.x;
    unsigned long long x = kValid, y = kValid;             // before the first chunk: inclusive prefix 0
    if (pre && idx == (int64_t)c - 1) { x = pre->x; y = pre->y; }
    else if (j >= 0) { x = ld_vol64(&X[j]); if (Y) y = ld_vol64(&Y[j]); }
    // every warp polls its own 32 predecessors: a warp whose words are all there goes on to the barrier
    // below and costs no issue slots while another one still waits for a publish
    for (;;) {
      const bool ok = (x >> 62) != 0 && (y >> 62) != 0;
      if (__all_sync(0xffffffffu, ok)) break;
      __nanosleep(32);
      if (!ok) { x = ld_vol64(&X[j]); if (Y) y = ld_vol64(&Y[j]); }
    }
    const uint32_t have_x = __ballot_sync(0xffffffffu, (x >> 63) != 0), have_y = __ballot_sync(0xffffffffu, (y >> 63) != 0);
    const unsigned Lx = have_x ? (unsigned)(__ffs(have_x) - 1) : 32u, Ly = have_y ? (unsigned)(__ffs(have_y) - 1) : 32u;
    const uint32_t s0 = __reduce_add_sync(0xffffffffu, lane < Lx ? (uint32_t)x & 0x7fffffffu : 0u);
    const uint32_t s1 = __reduce_add_sync(0xffffffffu, lane < Lx ? (uint32_t)(x >> 31) & 0x7fffffffu : 0u);
    uint32_t s2 = 0, s3 = 0;
    if (Y) {
      s2 = __reduce_add_sync(0xffffffffu, lane < Ly ? (uint32_t)y & 0x7fffffffu : 0u);
      s3 = __reduce_add_sync(0xffffffffu, lane < Ly ? (uint32_t)(y >> 31) & 0x7fffffffu : 0u);
    }
    const unsigned long long ix = __shfl_sync(0xffffffffu, x, Lx & 31), iy = __shfl_sync(0xffffffffu, y, Ly & 31);
    if (lane == 0) { lb[w].L[0] = Lx; lb[w].L[1] = Ly; lb[w].s[0] = s0; lb[w].s[1] = s1; lb[w].s[2] = s2; lb[w].s[3] = s3; lb[w].inc[0] = ix; lb[w].inc[1] = iy; }
    __syncthreads();
    // combine the warps' partial results (warp k looked at nearer chunks than warp k+1): lane k takes warp k's,
    // every warp does the same few collectives instead of every thread walking the list
    {
      const bool in = lane < (unsigned)kWarps;
      const LbWarp *q = lb + (in ? lane : 0);
      const uint32_t hx = __ballot_sync(0xffffffffu, in && q->L[0] < 32), hy = __ballot_sync(0xffffffffu, in && q->L[1] < 32);
      const unsigned kx = hx ? (unsigned)(__ffs(hx) - 1) : 31u, ky = hy ? (unsigned)(__ffs(hy) - 1) : 31u;   // first warp that saw an inclusive prefix
      if (!found_x) {
        excl[0] += __reduce_add_sync(0xffffffffu, in && lane <= kx ? q->s[0] : 0u);
        excl[1] += __reduce_add_sync(0xffffffffu, in && lane <= kx ? q->s[1] : 0u);
        if (hx) { const unsigned long long inc = lb[kx].inc[0]; excl[0] += inc & 0x7fffffffull; excl[1] += inc >> 31 & 0x7fffffffull; found_x = true; }
      }
      if (!found_y) {
        excl[2] += __reduce_add_sync(0xffffffffu, in && lane <= ky ? q->s[2] : 0u);
        excl[3] += __reduce_add_sync(0xffffffffu, in && lane <= ky ? q->s[3] : 0u);
        if (hy) { const unsigned long long inc = lb[ky].inc[1]; excl[2] += inc & 0x7fffffffull; excl[3] += inc >> 31 & 0x7fffffffull; found_y = true; }
      }
    }
    if (found_x && found_y) break;  // (the next use of lb is behind other barriers)
    __syncthreads();                // lb is reused by the next round
  }
  if (threadIdx.x == 0) {
    st_vol64(&X[c], kValid | chain_pack((uint32_t)(excl[0] + (aggX & 0x7fffffffu)), (uint32_t)(excl[1] + (aggX >> 31 & 0x7fffffffu))));
    if (Y) st_vol64(&Y[c], kValid | chain_pack((uint32_t)(excl[2] + (aggY & 0x7fffffffu)), (uint32_t)(excl[3] + (aggY >> 31 & 0x7fffffffu))));
  }
}

TI_D void note_error(Misc *ms, uint32_t flags, uint32_t off) {
  if (!flags) return;
  atomicOr(&ms->err, flags);
  atomicMin(&ms->first_err, off);
}

// newline bitmap of 16 bytes
TI_D uint32_t nl_mask16(const uint4 x) {
  const uint32_t a = __vcmpeq4(x.x, 0x0a0a0a0au) & 0x01010101u, b = __vcmpeq4(x.y, 0x0a0a0a0au) & 0x01010101u;
  const uint32_t c = __vcmpeq4(x.z, 0x0a0a0a0au) & 0x01010101u, d = __vcmpeq4(x.w, 0x0a0a0a0au) & 0x01010101u;
  return ((a * 0x00204081u >> 21) & 0xf) | ((b * 0x00204081u >> 21) & 0xf) << 4 |
         ((c * 0x00204081u >> 21) & 0xf) << 8 | ((d * 0x00204081u >> 21) & 0xf) << 12;
}

// Exclusive scan of small per-item values over items [0, n), in item order, by the
// whole CTA (two ballot-free passes, one barrier between them, one after unless the caller's
// next reader of emit()'s output sits behind another barrier anyway).  val(i) is evaluated twice.
template <class V, class E>
TI_D uint32_t block_scan(uint32_t n, uint64_t *warp_tot, V val, E emit, bool sync_after = true) {
  const unsigned lane = lane_id(), w = threadIdx.x >> 5;
  const uint32_t seg = ((n + kWarps - 1) / kWarps + 31) & ~31u;
  const uint32_t lo = w * seg, hi = lo + seg < n ? lo + seg : n;
  uint32_t tot = 0;
  for (uint32_t i = lo + lane; i < lo + seg && lo < n; i += 32) {
    uint32_t v = i < hi ? val(i) : 0;
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(kFull, v, d);
    tot += v;
  }
  if (lane == 0) warp_tot[w] = tot;
  __syncthreads();
  uint32_t base = 0, total = 0;
  for (unsigned k = 0; k < (unsigned)kWarps; ++k) { const uint32_t t = (uint32_t)warp_tot[k]; if (k < w) base += t; total += t; }
  for (uint32_t i = lo + lane; i < lo + seg && lo < n; i += 32) {
    const uint32_t v = i < hi ? val(i) : 0;
    uint32_t inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t up = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += up; }
    if (i < hi) emit(i, base + inc - v, v);
    base += __shfl_sync(kFull, inc, 31);
  }
  if (sync_after) __syncthreads();   // what emit() wrote is visible to the whole CTA
  return total;
}

// smem offset of the first '\n' at or after offset o (the sentinel guarantees one)
TI_D unsigned line_end(const uint32_t *nlmap32, unsigned o) {
  unsigned w = o >> 5;
  uint32_t bits = nlmap32[w] >> (o & 31);
  if (bits) return o + __ffs(bits) - 1;
  do bits = nlmap32[++w]; while (!bits);
  return (w << 5) + __ffs(bits) - 1;
}

TI_D bool parse_i32(const uint8_t *t, int &q, int32_t &v) {  // optional '-', <= 10 digits, fits int32
  const bool neg = t[q] == '-';
  if (neg) ++q;
  int64_t m = 0;
  int nd = 0;
  unsigned d;
  while ((d = t[q] - '0') <= 9) { m = m * 10 + d; ++q; if (++nd > 10) return false; }
  if (!nd || m > 0x7fffffffLL + (neg ? 1 : 0)) return false;
  v = (int32_t)(neg ? -m : m);
  return true;
}

// ---- structural lines ("#BEGIN_TB", "#END_TB", "thread block = x,y,z", "warp = w",
// "insts = n"; trace_parser.cc:414-437): one convergent routine for all five shapes -- the
// literal is compared as two 8-byte words under a length mask, then up to three
// comma-separated integers follow -- so a warp holding a mix of them does not run one path
// per shape.
enum { SL_BEGIN = 0, SL_END = 1, SL_TB = 2, SL_WARP = 3, SL_INSTS = 4, SL_BAD = 5 };
constexpr uint64_t lit_word(const char *s, int from, int len) {
  uint64_t v = 0;
  for (int i = 0; i < 8; ++i) if (from + i < len) v |= (uint64_t)(uint8_t)s[from + i] << (8 * i);
  return v;
}
constexpr uint64_t kLitBegin0 = lit_word("#BEGIN_TB", 0, 9), kLitBegin1 = lit_word("#BEGIN_TB", 8, 9);
constexpr uint64_t kLitEnd0 = lit_word("#END_TB", 0, 7);
constexpr uint64_t kLitTb0 = lit_word("thread block = ", 0, 15), kLitTb1 = lit_word("thread block = ", 8, 15);
constexpr uint64_t kLitWarp0 = lit_word("warp = ", 0, 7);
constexpr uint64_t kLitInsts0 = lit_word("insts = ", 0, 8);
struct SlParsed { unsigned type; int32_t v[3]; int q; bool ok; };

TI_D SlParsed parse_struct_line(const uint8_t *text, const uint32_t *t32, int o) {
  SlParsed r;
  const unsigned c = text[o];
  r.type = c == '#' ? (text[o + 1] == 'B' ? SL_BEGIN : SL_END) : c == 't' ? SL_TB : c == 'w' ? SL_WARP : c == 'i' ? SL_INSTS : SL_BAD;
  const unsigned t = r.type;
  const uint64_t l0 = t == SL_BEGIN ? kLitBegin0 : t == SL_END ? kLitEnd0 : t == SL_TB ? kLitTb0 : t == SL_WARP ? kLitWarp0 : kLitInsts0;
  const uint64_t l1 = t == SL_BEGIN ? kLitBegin1 : t == SL_TB ? kLitTb1 : 0ull;
  const unsigned len = t == SL_BEGIN ? 9u : t == SL_TB ? 15u : t == SL_INSTS ? 8u : 7u;
  const unsigned nints = t == SL_TB ? 3u : (t == SL_WARP || t == SL_INSTS) ? 1u : 0u;
  const uint64_t m0 = len >= 8 ? ~0ull : (1ull << (8 * len)) - 1, m1 = len > 8 ? (1ull << (8 * (len - 8))) - 1 : 0ull;
  const uint64_t a0 = load8(t32, (unsigned)o), a1 = load8(t32, (unsigned)o + 8);
  r.ok = t != SL_BAD && (((a0 ^ l0) & m0) | ((a1 ^ l1) & m1)) == 0;
  r.q = o + (int)len;
  r.v[0] = r.v[1] = r.v[2] = 0;
#pragma unroll
  for (unsigned k = 0; k < 3; ++k)
    if (k < nints && r.ok) {
      if (k) r.ok = text[r.q++] == ',';
      r.ok = r.ok && parse_i32(text, r.q, r.v[k]);
    }
  r.ok = r.ok && text[r.q] == '\n';
  return r;
}

}  // namespace ti

#include "ti_kernel_main.cuh"

#endif
