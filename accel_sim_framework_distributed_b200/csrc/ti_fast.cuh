// Fast path of P2 for lines in the shape the tracer writes (tracer_tool.cu:685-756):
//
//     <pc: 1-8 hex> <mask: 1-8 hex> <0|1> [R<n>] <opcode> <0-4> [R<n>]* <mem width> [<mode digit> ...]
//
// Token boundaries come from 32-bit windows of the delimiter bitmap P1 built (one window
// at the line start, one at the opcode, one behind it), so finding a token is a
// find-first-set, not a character loop; numbers are converted with SWAR; and everything
// that depends on the opcode STRING alone (ISA lookup, modifier scan, the opcode switch of
// trace_driven.cc:254-378) comes from a per-CTA memo in shared memory keyed on the whole
// string -- a kernel has a few dozen distinct opcode strings, so after the first chunk
// every line hits.  A line that is not in this shape, or whose opcode is not in the memo
// yet, is NOT decoded here: it goes on a list and the general parser (ti_parse.cuh, which
// defines the semantics) takes it afterwards and fills the memo.  The two paths share
// apply_opcode_facts(), so a record does not depend on which one produced it.
#ifndef TI_FAST_CUH
#define TI_FAST_CUH

#include "ti_mem.cuh"
#include "ti_parse.cuh"

namespace ti {

struct MemoEnt { unsigned long long k0, k1, k2, facts; };   // opcode string, zero padded to 24 bytes; facts == 0: empty
constexpr int kMemoSlots = 64;
constexpr int kMemoProbes = 4;
static_assert(sizeof(MemoEnt) == 32 && (kMemoSlots & (kMemoSlots - 1)) == 0, "memo layout");

TI_HD uint32_t memo_hash(uint64_t k0, uint64_t k1, uint64_t k2) {
  uint32_t h = (uint32_t)k0 * 0x9E3779B1u ^ (uint32_t)(k0 >> 32) * 0x85EBCA77u ^ (uint32_t)k1 * 0xC2B2AE3Du ^
               (uint32_t)(k1 >> 32) * 0x27D4EB2Fu ^ (uint32_t)k2 * 0x165667B1u ^ (uint32_t)(k2 >> 32) * 0xD3A2646Du;
  h ^= h >> 15;
  return h & (uint32_t)(kMemoSlots - 1);
}

// the opcode token [ob, ob + len), len <= kMemoMaxLen, as three zero-padded words
TI_D void memo_key(const uint32_t *t32, unsigned ob, unsigned len, uint64_t &k0, uint64_t &k1, uint64_t &k2) {
  k0 = keep_bytes(load8(t32, ob), (int)len);
  k1 = k2 = 0;
  if (len > 8) k1 = keep_bytes(load8(t32, ob + 8), (int)len - 8);
  if (len > 16) k2 = keep_bytes(load8(t32, ob + 16), (int)len - 16);
}

TI_D uint64_t memo_find(const MemoEnt *memo, uint64_t k0, uint64_t k1, uint64_t k2, unsigned len) {
  uint32_t h = memo_hash(k0, k1, k2);
  uint64_t found = 0;
#pragma unroll 1
  for (int p = 0; p < kMemoProbes; ++p, h = (h + 1) & (kMemoSlots - 1)) {
    const ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(&memo[h].k0);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2 *>(&memo[h].k2);
    if (!b.y) break;                                                  // empty slot: the string was never inserted
    if (a.x == k0 && a.y == k1 && b.x == k2 && ((unsigned)(b.y >> 56) & 31u) == len) { found = b.y; break; }
  }
  return found;
}

// Called by the thread that ran the general parser on a line (between two CTA barriers that
// separate it from every memo_find): claims the first free slot of the probe sequence.  Two
// threads inserting the same string in the same pass may both succeed; lookups then find the first.
TI_D void memo_insert(MemoEnt *memo, uint64_t k0, uint64_t k1, uint64_t k2, uint64_t facts) {
  uint32_t h = memo_hash(k0, k1, k2);
  for (int p = 0; p < kMemoProbes; ++p, h = (h + 1) & (kMemoSlots - 1)) {
    if (memo[h].facts == facts && memo[h].k0 == k0 && memo[h].k1 == k1 && memo[h].k2 == k2) return;   // already there
    // the high word of a valid facts value is never zero (kOpfValid), so it doubles as the claim flag
    uint32_t *claim = reinterpret_cast<uint32_t *>(&memo[h].facts) + 1;
    if (atomicCAS(claim, 0u, (uint32_t)(facts >> 32)) == 0u) {
      memo[h].k0 = k0; memo[h].k1 = k1; memo[h].k2 = k2;
      *reinterpret_cast<uint32_t *>(&memo[h].facts) = (uint32_t)facts;
      return;
    }
  }
}

// 32 bits of a bitmap starting at bit q
TI_D uint32_t win32(const uint32_t *m, unsigned q) { return __funnelshift_r(m[q >> 5], m[(q >> 5) + 1], q & 31); }

// "R<1-3 digits>" in [b, b + n), value <= 255 (any other n is rejected)
TI_D bool fast_reg(const uint32_t *t32, unsigned b, unsigned n, unsigned &v) {
  const uint32_t x = load4(t32, b);
  const bool shape = n - 2 < 3;                             // 2 <= n <= 4
  const unsigned k = shape ? n - 1 : 1u;                    // digits
  const unsigned sh = 8 * (3 - k);
  uint32_t z = ((x >> 8) << sh) & 0xffffffu;                // digits right-aligned in three bytes, first digit lowest
  z |= 0x303030u & ((1u << sh) - 1);                        // padded with '0'
  const uint32_t dig = (z + 0x505050u) & ~(z + 0x464646u) & ~z & 0x808080u;   // '0' <= c <= '9'
  const uint32_t d = z - 0x303030u;
  v = (d & 0xff) * 100 + (d >> 8 & 0xff) * 10 + (d >> 16 & 0xff);
  return shape && (x & 0xff) == 'R' && dig == 0x808080u && v <= 255;
}

// hexadecimal token [b, b + n), 1 <= n <= 8, no prefix
TI_D bool fast_hex32(const uint32_t *t32, unsigned b, unsigned n, uint32_t &v) {
  uint32_t hv;
  bool ok = true;
  v = 0;
  unsigned k = n & 3;
  if (k) { ok = hex4(load4(t32, b), k, hv); v = hv; }
  for (; k < n; k += 4) { ok = hex4(load4(t32, b + k), 4, hv) && ok; v = v << 16 | hv; }
  return ok;
}

// n (1..4) decimal characters in the low bytes of x (first character lowest) -> value; false if any is not [0-9]
TI_D bool dec4(uint32_t x, unsigned n, uint32_t &v) {
  const unsigned sh = 8 * (4 - n);
  x = (x << sh) | (0x30303030u & ((1u << sh) - 1));       // right-align, pad with '0'
  const uint32_t dig = (x + 0x50505050u) & ~(x + 0x46464646u) & ~x & 0x80808080u;
  const uint32_t d = x - 0x30303030u;
  const uint32_t t = (d & 0x00ff00ffu) * 10 + ((d >> 8) & 0x00ff00ffu);   // two-digit pairs: first pair in the low half
  v = (t & 0xffffu) * 100 + (t >> 16);
  return dig == 0x80808080u;
}

// The two tokens "0x<1-16 hex digits> <[-]1-8 decimal digits>" behind the address-mode field of a base+stride
// line (q = smem offset of the delimiter after the mode token, le = the line's '\n'), located with one window of
// the delimiter bitmap and converted four characters at a time.  false = some other shape: the character
// loops of stride_parse (ti_mem.cuh), which define the semantics, take the line.
TI_D bool stride_tokens_fast(const uint8_t *text, const uint32_t *t32, const uint32_t *dlmap32, unsigned q, unsigned le, uint64_t &base, int64_t &stride) {
  const uint32_t D = win32(dlmap32, q);
  uint32_t S = ~D & (D << 1), E = D & ~(D << 1) & ~1u;      // bit 0 is the delimiter at q itself
  const unsigned s0 = __ffs(S) - 1, e0 = __ffs(E) - 1;
  S &= S - 1; E &= E - 1;
  const unsigned s1 = __ffs(S) - 1, e1 = __ffs(E) - 1;
  if (!S || !E || q + e1 > le) return false;                // two whole tokens inside the window and the line
  const unsigned nd = e0 - s0 - 2, neg = text[q + s1] == '-' ? 1u : 0u, m = e1 - s1 - neg;
  if (nd - 1 >= 16 || m - 1 >= 8) return false;
  if (((load4(t32, q + s0) | 0x2000u) & 0xffffu) != 0x7830u) return false;   // "0x" / "0X"
  const unsigned b0 = q + s0 + 2, b1 = q + s1 + neg;
  uint32_t hv;
  bool ok = true;
  uint64_t v = 0;
  unsigned k = nd & 3;
  if (k) { ok = hex4(load4(t32, b0), k, hv); v = hv; }
  for (; k < nd; k += 4) { ok = hex4(load4(t32, b0 + k), 4, hv) && ok; v = v << 16 | hv; }
  uint32_t sv = 0, dv;
  k = m & 3;
  if (k) { ok = dec4(load4(t32, b1), k, dv) && ok; sv = dv; }
  for (; k < m; k += 4) { ok = dec4(load4(t32, b1 + k), 4, dv) && ok; sv = sv * 10000 + dv; }
  base = v;
  stride = neg ? -(int64_t)sv : (int64_t)sv;
  return ok;
}

// stride_parse with the fast token reader in front (same results: the fast reader only accepts what the
// character loops would read to the same values)
TI_D bool stride_parse_fast(const uint8_t *text, const uint32_t *dlmap32, unsigned q, unsigned le, uint32_t mask, unsigned ds, StrideDesc &d, uint32_t &flags) {
  uint64_t b;
  int64_t st;
  if (!stride_tokens_fast(text, reinterpret_cast<const uint32_t *>(text), dlmap32, q, le, b, st)) return stride_parse(text, q, mask, ds, d, flags);
  d.base = b; d.stride = st; d.first = 0; d.n = 0;
  return stride_finish(mask, ds, d);
}

// Returns true and fills the record (as its two 16-byte halves) and `kind` when the line at smem offset
// o is in the fast shape and its opcode is memoised; false = leave it to the general parser.
TI_D bool parse_inst_fast(const uint8_t *text, const uint32_t *dlmap32, const MemoEnt *memo, unsigned o, unsigned le, const DecodeParams &P,
                          uint32_t chunk_text_base, uint4 &rec_lo, uint4 &rec_hi, unsigned &kind) {
  const uint32_t *t32 = reinterpret_cast<const uint32_t *>(text);
  // ---- window A at the line start: pc, mask, ndst, [Rd], start of the opcode.  text[o] is a hex digit
  // (that is what made this an instruction line), so bit 0 is clear.
  const uint32_t D = win32(dlmap32, o);
  uint32_t E = D & ~(D << 1), S = ~D & (D << 1);            // token ends (first delimiter after a token) / token starts
  bool ok = __popc(E) >= 3;
  const unsigned e0 = __ffs(E) - 1; E &= E - 1;
  const unsigned s1 = __ffs(S) - 1; S &= S - 1;
  const unsigned e1 = __ffs(E) - 1; E &= E - 1;
  const unsigned s2 = __ffs(S) - 1; S &= S - 1;
  const unsigned e2 = __ffs(E) - 1; E &= E - 1;
  ok = ok && e0 <= 8 && e1 - s1 <= 8 && e2 - s2 == 1;
  uint32_t pc = 0, mask = 0;
  if (ok) ok = fast_hex32(t32, o, e0, pc) & fast_hex32(t32, o + s1, e1 - s1, mask);
  const unsigned c_nd = text[o + (s2 & 31)] - '0';
  ok = ok && c_nd <= 1;
  const unsigned ndst = c_nd & 1;
  unsigned dst = 0;
  if (ndst) {
    ok = ok && E != 0;
    const unsigned s3 = __ffs(S) - 1; S &= S - 1;
    const unsigned e3 = __ffs(E) - 1; E &= E - 1;
    ok = ok && fast_reg(t32, o + (s3 & 31), e3 - s3, dst);
  }
  ok = ok && S != 0;
  const unsigned ob = o + ((__ffs(S) - 1) & 31);
  // ---- window B at the opcode: its length
  const uint32_t DB = win32(dlmap32, ob);
  const unsigned len = __ffs(DB) - 1;                       // DB == 0: 0xffffffff
  ok = ok && len <= (unsigned)kMemoMaxLen;
  if (!ok) return false;
  uint64_t k0, k1, k2;
  memo_key(t32, ob, len, k0, k1, k2);
  const uint64_t facts = memo_find(memo, k0, k1, k2, len);
  if (!facts) return false;
  // ---- window C at the delimiter behind the opcode: nsrc, [Rs]*, mem width, [mode]
  const unsigned oe = ob + len;
  const uint32_t DC = win32(dlmap32, oe);
  E = DC & ~(DC << 1) & ~1u;
  S = ~DC & (DC << 1);
  ok = __popc(E) >= 2;
  unsigned s = __ffs(S) - 1, e = __ffs(E) - 1;
  S &= S - 1; E &= E - 1;
  const unsigned c_ns = text[oe + (s & 31)] - '0';
  ok = ok && e - s == 1 && c_ns <= 4;
  const unsigned nsrc = c_ns & 7;
  uint32_t srcs = 0;
  for (unsigned i = 0; i < nsrc && ok; ++i) {
    ok = E != 0 && (E & (E - 1)) != 0;                      // this register and the mem-width token both end inside the window
    s = __ffs(S) - 1; e = __ffs(E) - 1;
    S &= S - 1; E &= E - 1;
    unsigned rv;
    ok = ok && fast_reg(t32, oe + (s & 31), e - s, rv);
    srcs |= rv << (8 * i);
  }
  s = __ffs(S) - 1; e = __ffs(E) - 1;                       // mem width: 1-2 decimal digits
  S &= S - 1; E &= E - 1;
  const unsigned nw = e - s;
  const unsigned w_a = text[oe + (s & 31)] - '0', w_b = text[oe + ((s + 1) & 31)] - '0';
  ok = ok && (nw == 1 || nw == 2) && w_a <= 9 && (nw == 1 || w_b <= 9);
  const bool has_mem = (w_a | (nw == 2 ? w_b : 0u)) != 0;
  unsigned mode = 0, q = oe + e;
  if (has_mem) {
    ok = ok && E != 0;
    s = __ffs(S) - 1; e = __ffs(E) - 1;
    const unsigned c_m = text[oe + (s & 31)] - '0';
    ok = ok && e - s == 1 && c_m <= 9;
    mode = c_m > 3 ? 3u : c_m;
    q = oe + e;
  }
  if (!ok || q > le) return false;                          // a token taken from behind the line's '\n'
  OpWords w;
  if (opcode_words(w, facts, mask, has_mem, mode, q, ndst, nsrc, len, P)) return false;
  rec_lo = make_uint4(pc, mask, chunk_text_base + ob, w.w3);
  rec_hi = make_uint4(dst | srcs << 8, srcs >> 24 | w.w5hi, w.w6, w.w7);
  kind = w.kind;
  return true;
}

}  // namespace ti
#endif
