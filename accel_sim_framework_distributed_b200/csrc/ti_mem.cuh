// Warp-per-instruction memory phase of the fused trace-ingest kernel: lane i owns
// trace lane i of one memory instruction.
//   * address decompression (trace_parser.cc:86-125,181-209): list / base+stride /
//     base+delta, tokens located with ballots over the address text in smem
//   * shared-memory bank conflicts (abstract_hardware_model.cc:329-420)
//   * global coalescing into 32 B sectors / 64 B / 128 B segments incl. the
//     straddle rule and the ascending emit order (:475-586), atomics (:588-692) and
//     the size reduction (:694-748)
#ifndef TI_MEM_CUH
#define TI_MEM_CUH

#include "ti_device.cuh"
#include "ti_parse.cuh"

#ifndef TI_SWAR_ADDR
#define TI_SWAR_ADDR 1
#endif

namespace ti {

constexpr unsigned kFull = 0xffffffffu;

struct Txn {  // a transaction_info (abstract_hardware_model.h:1131-1141) in registers
  uint64_t block;
  uint32_t lanes, bytes[4], chunks;
};

TI_D unsigned lane_id() { return threadIdx.x & 31; }
TI_D unsigned lanemask_lt() { return (1u << lane_id()) - 1; }

// 64-bit warp minimum from two 32-bit redux steps
TI_D uint64_t warp_min_u64(uint64_t v) {
  const uint32_t hi = __reduce_min_sync(kFull, (uint32_t)(v >> 32));
  const uint32_t lo = __reduce_min_sync(kFull, (uint32_t)(v >> 32) == hi ? (uint32_t)v : 0xffffffffu);
  return (uint64_t)hi << 32 | lo;
}

// bits [idx, idx+n) of a 128-bit mask, clipped at 128, as four words
TI_D void byte_words(unsigned idx, unsigned n, uint32_t w[4]) {
  w[0] = w[1] = w[2] = w[3] = 0;
  const uint64_t v = (uint64_t)((1u << n) - 1) << (idx & 31);
  const unsigned k = idx >> 5;
  w[k] = (uint32_t)v;
  if (k < 3) w[k + 1] = (uint32_t)(v >> 32);
}

// abstract_hardware_model.cc:482-507
TI_D unsigned segment_size(const DecodeParams &P, unsigned data_size, unsigned cache_op) {
  bool sector = false;
  if (P.coalesce_arch >= 20 && P.coalesce_arch < 39) sector = P.gmem_skip_L1D || cache_op == CACHE_GLOBAL;
  else if (P.coalesce_arch >= 40) sector = true;
  switch (data_size) {
    case 1: return 32;
    case 2: return sector ? 32 : 64;
    case 4: case 8: case 16: return sector ? 32 : 128;
  }
  return 0;
}

// abstract_hardware_model.cc:694-748.  The finished mem_access_t as the two 16-byte halves of a
// ti_txn_rec (addr | lane_mask | byte_mask[0] and byte_mask[1..3] | size, sector_mask, type, is_write),
// built in registers: a struct with byte fields read back as vectors would live in local memory.
struct TxnWords { uint4 a, b; };
TI_D TxnWords reduce_and_send(const Txn &t, unsigned seg, unsigned type, unsigned is_write) {
  uint64_t addr = t.block;
  unsigned size = seg, q = t.chunks, h = 0;
  if (seg == 128) {
    const bool lo = q & 3, hi = q & 12;
    if (lo && !hi) { size = 64; h = q & 3; }
    else if (!lo && hi) { addr += 64; size = 64; h = (q >> 2) & 3; }
  } else if (seg == 64) {
    h = (addr % 128 == 0) ? (q & 3) : ((q >> 2) & 3);
  }
  if (size == 64) {
    const bool lo = h & 1, hi = h & 2;
    if (lo && !hi) size = 32;
    else if (!lo && hi) { addr += 32; size = 32; }
  }
  TxnWords r;
  r.a = make_uint4((uint32_t)addr, (uint32_t)(addr >> 32), t.lanes, t.bytes[0]);
  r.b = make_uint4(t.bytes[1], t.bytes[2], t.bytes[3], size | (t.chunks & 0xff) << 8 | type << 16 | is_write << 24);
  return r;
}
TI_D void txn_store(ti_txn_rec *dst, const TxnWords &w) {
  uint4 *d = reinterpret_cast<uint4 *>(dst);
  d[0] = w.a; d[1] = w.b;
}

// ---- compact staging form (16 B instead of 32): a transaction whose bytes are ONE run [p0, p0 + len) of its
// 128 B line -- every access of a lane on its own, and most merged ones -- is kept as block | lanes | run, and
// becomes a ti_txn_rec only when the chunk's transactions leave shared memory.
TI_D void run_bytes(unsigned p0, unsigned len, uint32_t (&bytes)[4], uint32_t &chunks) {
  const unsigned p1 = p0 + len;
  chunks = 0;
#pragma unroll
  for (unsigned w = 0; w < 4; ++w) {
    const unsigned a = p0 > 32 * w ? p0 : 32 * w, b = p1 < 32 * w + 32 ? p1 : 32 * w + 32;
    const bool hit = b > a;
    bytes[w] = hit ? ((b - a == 32 ? 0xffffffffu : ((1u << (b - a)) - 1)) << (a - 32 * w)) : 0u;
    chunks |= hit ? 1u << w : 0u;
  }
}
TI_D bool txn_is_run(const Txn &t, unsigned &p0, unsigned &len) {
  len = __popc(t.bytes[0]) + __popc(t.bytes[1]) + __popc(t.bytes[2]) + __popc(t.bytes[3]);
  p0 = t.bytes[0] ? __ffs(t.bytes[0]) - 1 : t.bytes[1] ? 31 + __ffs(t.bytes[1]) : t.bytes[2] ? 63 + __ffs(t.bytes[2]) : 95 + __ffs(t.bytes[3]);
  if (!len) return false;
  uint32_t b[4], c;
  run_bytes(p0, len, b, c);
  return b[0] == t.bytes[0] && b[1] == t.bytes[1] && b[2] == t.bytes[2] && b[3] == t.bytes[3] && c == t.chunks;
}
TI_D uint4 txn_pack_run(const Txn &t, unsigned p0, unsigned len, unsigned seg, unsigned is_write) {
  return make_uint4((uint32_t)t.block, (uint32_t)(t.block >> 32), t.lanes,
                    p0 | (len - 1) << 7 | (seg == 128 ? 2u : seg == 64 ? 1u : 0u) << 14 | is_write << 16);
}
TI_D TxnWords txn_unpack_run(const uint4 c) {
  Txn x;
  x.block = (uint64_t)c.y << 32 | c.x;
  x.lanes = c.z;
  run_bytes(c.w & 127, (c.w >> 7 & 127) + 1, x.bytes, x.chunks);
  const unsigned is_write = c.w >> 16 & 1;
  return reduce_and_send(x, 32u << (c.w >> 14 & 3), is_write ? 4u : 0u, is_write);
}

// ---- address decompression.  `q` = smem offset of the first token after the mode field, `le` =
// offset of the line's '\n'; tokpos = 33 u16 of per-warp scratch.  Returns this lane's address.
TI_D_COLD uint64_t decode_addresses(const uint8_t *text, const uint32_t *dlmap32, unsigned q, unsigned le, unsigned mode, uint32_t mask,
                               uint16_t *tokpos, uint32_t &flags) {
  const unsigned lane = lane_id();
  const bool active = mask >> lane & 1;
  if (mode == 1) {  // every lane reads the same two tokens (smem broadcast)
    Reader rd = {text, (int)q};
    uint64_t base = 0;
    int64_t stride = 0;
    rd.skip_ws();
    if (rd.peek() != '\n') {
      if (!rd.hex(16, base)) { flags |= TI_EF_BAD_NUMBER; return 0; }
      rd.skip_ws();
      if (rd.peek() != '\n' && !rd.sdec(0x7fffffffull, stride)) { flags |= TI_EF_BAD_NUMBER; return 0; }
    }
    if (!mask) return 0;
    const unsigned first = __ffs(mask) - 1;
    const uint32_t inv = ~(mask >> first);          // zero bit = still inside the first run
    const unsigned run = inv ? __ffs(inv) - 1 : 32; // length of the first contiguous run
    if (lane < first || lane >= first + run) return 0;  // incl. active lanes after the gap
    return base + (uint64_t)stride * (uint64_t)(lane - first);
  }
  if (mode != 0 && mode != 2) return 0;             // value-initialised addresses stay 0
  // ---- list / delta: token k belongs to the k-th active lane.  Token starts come from the
  // delimiter bitmap P1 built: lane l owns bitmap word (q >> 5) + l, i.e. 32 characters, so a
  // 650-character list-mode line is one round (a ballot per 32 characters took twenty).
  const unsigned nact = __popc(mask);
  const unsigned rank = __popc(mask & lanemask_lt());
  const unsigned wq = q >> 5, wle = le >> 5;
  unsigned ntok = 0;
  uint32_t carry = 1;                                // the character before q is a delimiter
  for (unsigned wb = wq;; wb += 32) {
    const unsigned w = wb + lane;
    const uint32_t dl = dlmap32[w < wle ? w : wle];
    uint32_t valid = w > wle ? 0u : kFull;
    if (w == wq) valid &= ~0u << (q & 31);
    if (w == wle) valid &= (1u << (le & 31)) - 1;
    const uint32_t up = __shfl_up_sync(kFull, dl, 1);
    uint32_t before = (dl << 1) | (lane ? up >> 31 : carry);     // bit b = character b-1 is a delimiter
    if (w == wq) before |= 1u << (q & 31);
    const uint32_t start = ~dl & before & valid;
    const unsigned cnt = __popc(start);
    unsigned inc = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const unsigned u = __shfl_up_sync(kFull, inc, d); if ((int)lane >= d) inc += u; }
    unsigned k = ntok + inc - cnt;
    for (uint32_t b = start; b; b &= b - 1, ++k)
      if (k < 33) tokpos[k] = (uint16_t)((w << 5) + __ffs(b) - 1);
    ntok += __shfl_sync(kFull, inc, 31);
    carry = __shfl_sync(kFull, dl, 31) >> 31;
    if (wb + 32 > wle || ntok >= nact) break;        // warp-uniform
  }
  __syncwarp();
  uint64_t val = 0;
  if (active && rank < ntok) {
    Reader rd = {text, (int)tokpos[rank]};
    bool ok;
    const uint32_t *t32 = reinterpret_cast<const uint32_t *>(text);
    const unsigned tp = tokpos[rank];
    // the tracer writes every address as 0x%016lx: 18 characters, converted four digits at a time
    uint32_t h0, h1, h2, h3;
    if (TI_SWAR_ADDR && (mode == 0 || rank == 0) && text[tp] == '0' && (text[tp + 1] | 0x20) == 'x' && is_end(text[tp + 18]) &&
        hex4(load4(t32, tp + 2), 4, h0) & hex4(load4(t32, tp + 6), 4, h1) & hex4(load4(t32, tp + 10), 4, h2) & hex4(load4(t32, tp + 14), 4, h3)) {
      val = (uint64_t)h0 << 48 | (uint64_t)h1 << 32 | (uint64_t)h2 << 16 | h3;
      ok = true;
    } else if (mode == 0 || rank == 0) ok = rd.hex(16, val);
    else { int64_t d; ok = rd.sdec(0x7fffffffffffffffull, d); val = (uint64_t)d; }
    if (!ok) flags |= TI_EF_BAD_NUMBER;
  }
  __syncwarp();
  if (mode == 0) return active ? val : 0;
  // base + running sum of deltas over the active lanes (wraps mod 2^64)
  uint64_t sum = active ? val : 0;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint64_t up = __shfl_up_sync(kFull, sum, d);
    if ((int)lane >= d) sum += up;
  }
  return active ? sum : 0;
}

// =====================================================================================
// Thread-per-instruction path for base+stride lines (the tracer's format for every
// regular access, tracer_tool.cu:599-636).  The addresses of the first contiguous run
// of active lanes are an arithmetic progression, so one thread can walk them in
// ascending order and keep at most two open transactions (the block of the current
// address and, for a straddling access, the next one): no sort, no cross-lane traffic,
// 32 instructions per warp in flight instead of one.
// =====================================================================================
struct StrideDesc {
  uint64_t base;
  int64_t stride;
  unsigned first, n;  // lanes [first, first + n) carry base + i * stride; every other lane reads 0
};

// lanes of the first contiguous run of active lanes; false when the progression wraps around 2^64
// (then it is not monotone and the warp path handles the instruction)
TI_D bool stride_finish(uint32_t mask, unsigned ds, StrideDesc &d) {
  if (!mask) return true;
  d.first = __ffs(mask) - 1;
  const uint32_t inv = ~(mask >> d.first);
  d.n = inv ? __ffs(inv) - 1 : 32;
  const uint64_t span = (uint64_t)(d.stride < 0 ? -d.stride : d.stride) * (d.n - 1);
  if (d.stride >= 0) return d.base + span + (ds - 1) >= d.base;  // no overflow past 2^64
  return d.base >= span && d.base + (ds - 1) >= d.base;
}

// Parses "0x<base> <stride>" at smem offset q (character loops: every shape the reference's stream
// extraction accepts).  Returns stride_finish's verdict.
TI_D bool stride_parse(const uint8_t *text, unsigned q, uint32_t mask, unsigned ds, StrideDesc &d, uint32_t &flags) {
  Reader rd = {text, (int)q};
  d.base = 0; d.stride = 0; d.first = 0; d.n = 0;
  rd.skip_ws();
  if (rd.peek() != '\n') {
    if (!rd.hex(16, d.base)) { flags |= TI_EF_BAD_NUMBER; return true; }
    rd.skip_ws();
    if (rd.peek() != '\n' && !rd.sdec(0x7fffffffull, d.stride)) { flags |= TI_EF_BAD_NUMBER; return true; }
  }
  return stride_finish(mask, ds, d);
}

TI_D uint64_t stride_addr(const StrideDesc &d, unsigned lane) {
  return lane >= d.first && lane < d.first + d.n ? d.base + (uint64_t)d.stride * (uint64_t)(lane - d.first) : 0;
}

// number of transactions: distinct blocks over all touches, visited in ascending order
TI_D unsigned stride_count(const StrideDesc &d, unsigned ds, unsigned seg) {
  const uint64_t segmask = ~(uint64_t)(seg - 1);
  unsigned cnt = 0;
  uint64_t hi = 0;
  bool any = false;
  for (unsigned k = 0; k < d.n; ++k) {
    const unsigned i = d.stride >= 0 ? k : d.n - 1 - k;
    const uint64_t a = d.base + (uint64_t)d.stride * i;
    if (a == 0) continue;                                // abstract_hardware_model.cc:534-537
    const uint64_t A = a & segmask, B = (a + ds - 1) & segmask;
    if (!any || A > hi) { ++cnt; hi = A; any = true; }
    if (B > hi) { ++cnt; hi = B; }
  }
  return cnt;
}

// ---- dense progression: stride == data size and an aligned base.  The run then covers
// one contiguous byte range with no access crossing a segment, so every transaction
// follows in closed form from its block: O(transactions), not O(lanes).  This is the
// shape of every unit-stride access of a real trace (and all of BASELINE config 2).
TI_D bool stride_is_dense(const StrideDesc &d, unsigned ds) {
  return d.n > 0 && d.stride == (int64_t)ds && d.base != 0 && (d.base & (ds - 1)) == 0;
}
TI_D unsigned dense_count(const StrideDesc &d, unsigned ds, unsigned seg) {
  const uint64_t segmask = ~(uint64_t)(seg - 1), end = d.base + (uint64_t)d.n * ds;
  return (unsigned)((((end - 1) & segmask) - (d.base & segmask)) >> (__ffs(seg) - 1)) + 1;   // seg is 32, 64 or 128: no 64-bit division
}
TI_D unsigned dense_emit(const StrideDesc &d, unsigned ds, unsigned seg, unsigned type, unsigned is_write, ti_txn_rec *out, unsigned room = ~0u) {
  const uint64_t segmask = ~(uint64_t)(seg - 1), end = d.base + (uint64_t)d.n * ds;
  const unsigned sh = ds == 1 ? 0 : ds == 2 ? 1 : ds == 4 ? 2 : ds == 8 ? 3 : 4;
  unsigned n = 0;
  for (uint64_t S = d.base & segmask; S < end; S += seg, ++n) {
    const uint64_t lo = S > d.base ? S : d.base, hi = S + seg < end ? S + seg : end;
    const unsigned i0 = (unsigned)((lo - d.base) >> sh), i1 = (unsigned)((hi - d.base) >> sh);  // lanes [i0, i1) of the run
    const unsigned p0 = (unsigned)(lo & 127), p1 = p0 + (unsigned)(hi - lo);                 // bytes [p0, p1) of the 128 B line
    Txn x;
    x.block = S;
    x.lanes = (i1 - i0 == 32 ? 0xffffffffu : ((1u << (i1 - i0)) - 1)) << (d.first + i0);
    x.chunks = 0;
#pragma unroll
    for (unsigned w = 0; w < 4; ++w) {
      const unsigned a = p0 > 32 * w ? p0 : 32 * w, b = p1 < 32 * w + 32 ? p1 : 32 * w + 32;
      const bool hit = b > a;
      x.bytes[w] = hit ? ((b - a == 32 ? 0xffffffffu : ((1u << (b - a)) - 1)) << (a - 32 * w)) : 0u;
      x.chunks |= hit ? 1u << w : 0u;
    }
    if (n < room) txn_store(out + n, reduce_and_send(x, seg, type, is_write));
  }
  return n;
}

// the k-th transaction of a dense progression, in closed form (one lane of a quad emits one transaction)
TI_D TxnWords dense_txn(const StrideDesc &d, unsigned ds, unsigned seg, unsigned type, unsigned is_write, unsigned k) {
  const uint64_t segmask = ~(uint64_t)(seg - 1), end = d.base + (uint64_t)d.n * ds;
  const unsigned sh = __ffs(ds) - 1;
  const uint64_t S = (d.base & segmask) + (uint64_t)k * seg;
  const uint64_t lo = S > d.base ? S : d.base, hi = S + seg < end ? S + seg : end;
  const unsigned i0 = (unsigned)((lo - d.base) >> sh), i1 = (unsigned)((hi - d.base) >> sh);  // lanes [i0, i1) of the run
  const unsigned p0 = (unsigned)(lo & 127), p1 = p0 + (unsigned)(hi - lo);                 // bytes [p0, p1) of the 128 B line
  Txn x;
  x.block = S;
  x.lanes = (i1 - i0 == 32 ? 0xffffffffu : ((1u << (i1 - i0)) - 1)) << (d.first + i0);
  x.chunks = 0;
#pragma unroll
  for (unsigned w = 0; w < 4; ++w) {
    const unsigned a = p0 > 32 * w ? p0 : 32 * w, b = p1 < 32 * w + 32 ? p1 : 32 * w + 32;
    const bool hit = b > a;
    x.bytes[w] = hit ? ((b - a == 32 ? 0xffffffffu : ((1u << (b - a)) - 1)) << (a - 32 * w)) : 0u;
    x.chunks |= hit ? 1u << w : 0u;
  }
  return reduce_and_send(x, seg, type, is_write);
}

struct OpenTxn { uint32_t lanes, b0, b1, b2, b3, chunks; };
TI_D void open_clear(OpenTxn &t) { t.lanes = t.b0 = t.b1 = t.b2 = t.b3 = t.chunks = 0; }
TI_D void open_touch(OpenTxn &t, uint64_t a, unsigned ds, unsigned lane) {
  const unsigned idx = (unsigned)(a & 127), k = idx >> 5;
  const uint64_t v = (uint64_t)((1u << ds) - 1) << (idx & 31);
  const uint32_t lo = (uint32_t)v, hi = (uint32_t)(v >> 32);
  t.b0 |= k == 0 ? lo : 0u;
  t.b1 |= (k == 1 ? lo : 0u) | (k == 0 ? hi : 0u);
  t.b2 |= (k == 2 ? lo : 0u) | (k == 1 ? hi : 0u);
  t.b3 |= (k == 3 ? lo : 0u) | (k == 2 ? hi : 0u);
  t.chunks |= 1u << k;
  t.lanes |= 1u << lane;
}
TI_D void open_flush(const OpenTxn &t, uint64_t block, unsigned seg, unsigned type, unsigned is_write, ti_txn_rec *out, unsigned &n, unsigned room) {
  if (!t.lanes) return;
  Txn x;
  x.block = block; x.lanes = t.lanes; x.bytes[0] = t.b0; x.bytes[1] = t.b1; x.bytes[2] = t.b2; x.bytes[3] = t.b3; x.chunks = t.chunks;
  if (n < room) txn_store(out + n, reduce_and_send(x, seg, type, is_write));
  ++n;
}

// writes the transactions in ascending block order (the std::map order of
// abstract_hardware_model.cc:576-584); returns how many
TI_D unsigned stride_emit(const StrideDesc &d, unsigned ds, unsigned seg, unsigned type, unsigned is_write, ti_txn_rec *out, unsigned room = ~0u) {
  const uint64_t segmask = ~(uint64_t)(seg - 1);
  OpenTxn c0, c1;             // blocks b and b + seg
  open_clear(c0); open_clear(c1);
  uint64_t b = 0;
  unsigned n = 0;
  for (unsigned k = 0; k < d.n; ++k) {
    const unsigned i = d.stride >= 0 ? k : d.n - 1 - k;
    const uint64_t a = d.base + (uint64_t)d.stride * i;
    if (a == 0) continue;
    const uint64_t A = a & segmask, last = a + ds - 1;
    if (A != b) {               // every later touch is >= A: close what lies below it
      open_flush(c0, b, seg, type, is_write, out, n, room);
      if (A == b + seg) { c0 = c1; }
      else { open_flush(c1, b + seg, seg, type, is_write, out, n, room); open_clear(c0); }
      open_clear(c1);
      b = A;
    }
    open_touch(c0, a, ds, d.first + i);
    if ((last & segmask) != A) open_touch(c1, last, ds, d.first + i);   // the straddle rule, :559-571
  }
  open_flush(c0, b, seg, type, is_write, out, n, room);
  open_flush(c1, b + seg, seg, type, is_write, out, n, room);
  return n;
}

// ---- shared-memory bank conflicts: returns `cycles` (warp-uniform)
TI_D_COLD unsigned shared_cycles(const DecodeParams &P, uint32_t mask, uint64_t addr) {
  const unsigned lane = lane_id();
  const unsigned parts = (unsigned)P.mem_warp_parts, sub = 32 / parts;
  const uint64_t word = addr & ~3ull;
  const unsigned bank = (unsigned)((addr / 4) % (unsigned)P.num_shmem_bank);
  unsigned total = 0;
  for (unsigned p = 0; p < parts; ++p) {
    const uint32_t pm = (sub == 32 ? kFull : ((1u << sub) - 1) << (p * sub)) & mask;
    if (!pm) continue;                                   // warp-uniform
    const bool in = pm >> lane & 1;
    unsigned grp = 0, nword = 0;
    bool leader = false;
    if (in) {
      grp = __match_any_sync(pm, word);                  // lanes touching the same word
      leader = (unsigned)(__ffs(grp) - 1) == lane;
      nword = __popc(grp);
    }
    unsigned acc = 0;
    if (!P.shmem_limited_broadcast) {
      const uint32_t lm = __ballot_sync(kFull, leader);
      if (leader) acc = __popc(__match_any_sync(lm, bank));  // distinct words per bank
    } else {
      // first (bank, word) in ascending order that more than one lane reads
      const bool multi = leader && nword > 1;
      const unsigned bbank = __reduce_min_sync(kFull, multi ? bank : 0xffffffffu);
      uint64_t bword = 0;
      unsigned bn = 0;
      if (bbank != 0xffffffffu) {
        bword = warp_min_u64(multi && bank == bbank ? word : ~0ull);
        const uint32_t who = __ballot_sync(kFull, multi && bank == bbank && word == bword);
        bn = __shfl_sync(kFull, nword, __ffs(who) - 1);
      }
      if (in) {
        acc = __popc(__match_any_sync(pm, bank));        // lanes per bank
        if (bbank != 0xffffffffu && bank == bbank) acc -= bn - 1;
      }
    }
    total += __reduce_max_sync(kFull, acc);
  }
  return total;
}

// ---- global coalescing (non-atomic), sort-free.
// Every active lane makes one "touch" A = (block of its address) and, when its access
// crosses into the next segment, a second touch B (abstract_hardware_model.cc:559-571).
// Transactions = distinct blocks over all touches, emitted in ascending block order
// (std::map, :511,576-584).  Lanes with equal blocks are grouped with match.any, their
// byte / chunk masks OR-reduced with redux over the group mask, and each group leader
// ranks its block by counting the leaders with a smaller one (one shuffle broadcast per
// leader) -- no serial extraction of minima, so 32 scattered sectors cost about as much
// as 4.  A B-group whose block some A-group also touches is merged into it through a
// small per-warp smem scratch; the others ("pure B") are transactions of their own.
// Outputs: out[0]/pos[0] = this lane's A-group if it leads one, out[1]/pos[1] = its pure
// B-group; pos = ~0u when absent.  Returns the number of transactions.
TI_D_COLD unsigned coalesce_global(const DecodeParams &P, uint32_t mask, uint64_t addr, unsigned ds, unsigned seg,
                              Txn out[2], unsigned pos[2], uint32_t *scratch) {
  const unsigned lane = lane_id();
  const unsigned parts = (unsigned)P.mem_warp_parts, sub = 32 / parts;
  const uint64_t segmask = ~(uint64_t)(seg - 1);
  const uint64_t last = addr + ds - 1;
  const uint64_t keyA = addr & segmask, keyB = last & segmask;
  const bool act = (mask >> lane & 1) && addr != 0;      // abstract_hardware_model.cc:534-537
  uint32_t wa[4], wb[4];
  byte_words((unsigned)(addr & 127), ds, wa);
  byte_words((unsigned)(last & 127), ds, wb);
  const unsigned ca = 1u << ((addr & 127) / 32), cb = 1u << ((last & 127) / 32);
  pos[0] = pos[1] = ~0u;
  unsigned n = 0;
  for (unsigned p = 0; p < parts; ++p) {
    const bool in = sub == 32 || (lane / sub) == p;
    const bool actp = act && in, stradp = actp && keyA != keyB;
    const uint32_t mA = __ballot_sync(kFull, actp);
    if (!mA) continue;                                   // warp-uniform
    const uint32_t mB = __ballot_sync(kFull, stradp);
    Txn ta, tb;
    bool leadA = false, leadB = false;
    if (actp) {
      const uint32_t g = __match_any_sync(mA, keyA);
      leadA = (unsigned)(__ffs(g) - 1) == lane;
      ta.block = keyA; ta.lanes = g;
      ta.bytes[0] = wa[0]; ta.bytes[1] = wa[1]; ta.bytes[2] = wa[2]; ta.bytes[3] = wa[3]; ta.chunks = ca;
      // a redux with a partial mask issues once per distinct mask: a fully divergent access would
      // run it 32 times with one lane each, so lanes alone in their block skip it
      if (g & (g - 1)) {
#pragma unroll
        for (int w = 0; w < 4; ++w) ta.bytes[w] = __reduce_or_sync(g, wa[w]);
        ta.chunks = __reduce_or_sync(g, ca);
      }
    }
    if (stradp) {
      const uint32_t g = __match_any_sync(mB, keyB);
      leadB = (unsigned)(__ffs(g) - 1) == lane;
      tb.block = keyB; tb.lanes = g;
      tb.bytes[0] = wb[0]; tb.bytes[1] = wb[1]; tb.bytes[2] = wb[2]; tb.bytes[3] = wb[3]; tb.chunks = cb;
      if (g & (g - 1)) {
#pragma unroll
        for (int w = 0; w < 4; ++w) tb.bytes[w] = __reduce_or_sync(g, wb[w]);
        tb.chunks = __reduce_or_sync(g, cb);
      }
    }
    const uint32_t LA = __ballot_sync(kFull, leadA);
    unsigned cntA = 0, cntB = 0;
    int owner = -1;
    // every leader's block goes through the warp's scratch once: a broadcast LDS.64 per leader instead
    // of two shuffles (a divergent access has 32 leaders)
    uint64_t *ks = reinterpret_cast<uint64_t *>(scratch);
    const bool many = !mB && __popc(LA) > 12;            // warp-uniform
    __syncwarp();
    ks[lane] = (many && !leadA) ? ~0ull : keyA;          // (all ones is no block address: never below a key)
    __syncwarp();
    if (many) {
      // a divergent access (up to 32 leaders): all 32 slots, two per 128-bit broadcast load, no loop over a mask
      const ulonglong2 *k2 = reinterpret_cast<const ulonglong2 *>(ks);
#pragma unroll
      for (int i = 0; i < 16; ++i) { const ulonglong2 v = k2[i]; cntA += (v.x < keyA) + (v.y < keyA); }
    } else if (!mB) {                                    // warp-uniform: nobody straddles, only A ranks are needed
      for (uint32_t m = LA; m; m &= m - 1) cntA += ks[__ffs(m) - 1] < keyA;
    } else {
      for (uint32_t m = LA; m; m &= m - 1) {
        const int j = __ffs(m) - 1;
        const uint64_t kj = ks[j];
        cntA += kj < keyA;
        cntB += kj < keyB;
        if (leadB && kj == keyB) owner = j;
      }
    }
    uint32_t LBp = 0;
    if (mB) {                                            // warp-uniform
      const bool pureB = leadB && owner < 0, owned = leadB && owner >= 0;
      LBp = __ballot_sync(kFull, pureB);
      __syncwarp();
      ks[lane] = keyB;
      __syncwarp();
      for (uint32_t m = LBp; m; m &= m - 1) {
        const int j = __ffs(m) - 1;
        const uint64_t kj = ks[j];
        cntA += kj < keyA;
        cntB += kj < keyB;
      }
      __syncwarp();                                      // the scratch is reused for the merge slots below
      const uint32_t ownedmask = __reduce_or_sync(kFull, owned ? 1u << owner : 0u);
      if (ownedmask) {
        if (owned) {
          uint32_t *s = scratch + owner * 6;
          s[0] = tb.bytes[0]; s[1] = tb.bytes[1]; s[2] = tb.bytes[2]; s[3] = tb.bytes[3]; s[4] = tb.chunks; s[5] = tb.lanes;
        }
        __syncwarp();
        if (leadA && (ownedmask >> lane & 1)) {
          const uint32_t *s = scratch + lane * 6;
          ta.bytes[0] |= s[0]; ta.bytes[1] |= s[1]; ta.bytes[2] |= s[2]; ta.bytes[3] |= s[3]; ta.chunks |= s[4]; ta.lanes |= s[5];
        }
        __syncwarp();
      }
      if (pureB) { out[1] = tb; pos[1] = n + cntB; }
    }
    if (leadA) { out[0] = ta; pos[0] = n + cntA; }
    n += __popc(LA) + __popc(LBp);
  }
  return n;
}

// ---- atomics: order-dependent first fit per block (abstract_hardware_model.cc:626-673).
// Transaction k lives in lane k; `pos` receives its emit position.  Returns the count,
// or sets TI_EF_ATOMIC_STRADDLE.
TI_D_COLD unsigned coalesce_atomic(const DecodeParams &P, uint32_t mask, uint64_t addr, unsigned ds, unsigned seg, Txn &mine,
                              unsigned &pos, uint32_t &flags) {
  const unsigned lane = lane_id();
  const unsigned parts = (unsigned)P.mem_warp_parts, sub = 32 / parts;
  const uint64_t segmask = ~(uint64_t)(seg - 1);
  unsigned n = 0;
  mine.block = 0; mine.lanes = 0; mine.chunks = 0; mine.bytes[0] = mine.bytes[1] = mine.bytes[2] = mine.bytes[3] = 0;
  pos = ~0u;
  for (unsigned p = 0; p < parts; ++p) {
    const unsigned first_of_part = n;
    for (unsigned t = p * sub; t < (p + 1) * sub; ++t) {
      if (!(mask >> t & 1)) continue;                    // warp-uniform
      const uint64_t a = __shfl_sync(kFull, addr, t);
      const uint64_t blk = a & segmask;
      if (blk != ((a + ds - 1) & segmask)) { flags |= TI_EF_ATOMIC_STRADDLE; return 0; }
      uint32_t tw[4];
      byte_words((unsigned)(a & 127), ds, tw);
      const bool cand = lane >= first_of_part && lane < n && mine.block == blk &&
                        !((mine.bytes[0] & tw[0]) | (mine.bytes[1] & tw[1]) | (mine.bytes[2] & tw[2]) | (mine.bytes[3] & tw[3]));
      const uint32_t cbm = __ballot_sync(kFull, cand);
      const unsigned k = cbm ? (unsigned)(__ffs(cbm) - 1) : n;
      if (lane == k) {
        if (!cbm) mine.block = blk;
        mine.bytes[0] |= tw[0]; mine.bytes[1] |= tw[1]; mine.bytes[2] |= tw[2]; mine.bytes[3] |= tw[3];
        mine.lanes |= 1u << t;
        mine.chunks |= 1u << ((a & 127) / 32);
      }
      if (!cbm) ++n;
    }
    // emit order inside the part: ascending block, then creation order
    unsigned rank = 0;
    for (unsigned j = first_of_part; j < n; ++j) {
      const uint64_t bj = __shfl_sync(kFull, mine.block, j);
      rank += (bj < mine.block) || (bj == mine.block && j < lane);
    }
    if (lane >= first_of_part && lane < n) pos = first_of_part + rank;
  }
  return n;
}

}  // namespace ti
#endif
