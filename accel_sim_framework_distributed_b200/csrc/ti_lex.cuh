// Bitmap-driven lexer for instruction lines.
//
// P1 of the kernel classifies every staged byte once with SWAR compares and keeps two
// bitmaps in shared memory: newlines (line ends) and delimiters (every character the
// reference's stream extraction treats as white space: ' ' and 0x09..0x0d).  A thread
// that parses a line then finds each token with two find-first-set steps on the
// delimiter bitmap instead of testing characters one by one, fetches the token as 8-byte
// windows (three aligned 32-bit smem loads + funnel shifts) and converts numbers four
// characters at a time (SWAR hex) -- about a quarter of the instructions of a
// character loop, with no data-dependent inner loops for the warp to diverge on.
// Tokens outside the common shapes (more than 8 hex digits for a 32-bit field, more
// than 16 for an address, long decimals, ...) fall back to the character loops in
// ti_parse.cuh, which define the exact semantics.
#ifndef TI_LEX_CUH
#define TI_LEX_CUH

#include "ti_device.cuh"

#ifdef TI_EMU
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned s) {
  s &= 31;
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
}
#endif

namespace ti {

// 0x80 in every byte of y that is zero (exact, no cross-byte carries)
TI_HD uint32_t zbytes(uint32_t y) { return ~(((y & 0x7f7f7f7fu) + 0x7f7f7f7fu) | y | 0x7f7f7f7fu); }
TI_HD uint32_t gather4(uint32_t m80) { return ((m80 >> 7) * 0x00204081u >> 21) & 0xfu; }  // bit 7 of each byte -> 4 bits

// newline and delimiter flags (bit 7 of each byte) of four text bytes
TI_HD void cls_word(uint32_t x, uint32_t &nl, uint32_t &dl) {
  nl = zbytes(x ^ 0x0a0a0a0au);
  const uint32_t y = x & 0x7f7f7f7fu;
  const uint32_t ctl = (y + 0x77777777u) & ~(y + 0x72727272u) & ~x & 0x80808080u;  // 9 <= byte <= 13
  dl = zbytes(x ^ 0x20202020u) | ctl;
}
TI_HD void cls_masks16(const uint4 x, uint32_t &nl16, uint32_t &dl16) {
  uint32_t n0, n1, n2, n3, d0, d1, d2, d3;
  cls_word(x.x, n0, d0); cls_word(x.y, n1, d1); cls_word(x.z, n2, d2); cls_word(x.w, n3, d3);
  nl16 = gather4(n0) | gather4(n1) << 4 | gather4(n2) << 8 | gather4(n3) << 12;
  dl16 = gather4(d0) | gather4(d1) << 4 | gather4(d2) << 8 | gather4(d3) << 12;
}

// first set bit at or after position q (the sentinel guarantees one)
TI_D unsigned next_one(const uint32_t *m, unsigned q) {
  unsigned w = q >> 5;
  uint32_t bits = m[w] >> (q & 31);
  if (bits) return q + __ffs(bits) - 1;
  do bits = m[++w]; while (!bits);
  return (w << 5) + __ffs(bits) - 1;
}
// first clear bit in [q, lim), or lim
TI_D unsigned next_zero(const uint32_t *m, unsigned q, unsigned lim) {
  unsigned w = q >> 5;
  uint32_t bits = ~m[w] >> (q & 31);
  if (bits) { const unsigned r = q + __ffs(bits) - 1; return r < lim ? r : lim; }
  for (;;) {
    ++w;
    if ((w << 5) >= lim) return lim;
    bits = ~m[w];
    if (bits) { const unsigned r = (w << 5) + __ffs(bits) - 1; return r < lim ? r : lim; }
  }
}

// eight text bytes starting at smem offset o (little endian: byte k = character k)
TI_D uint64_t load8(const uint32_t *t32, unsigned o) {
  const unsigned w = o >> 2, s = (o & 3) * 8;
  const uint32_t a = t32[w], b = t32[w + 1], c = t32[w + 2];
  return (uint64_t)__funnelshift_r(b, c, s) << 32 | __funnelshift_r(a, b, s);
}
TI_D uint32_t load4(const uint32_t *t32, unsigned o) {
  const unsigned w = o >> 2, s = (o & 3) * 8;
  return __funnelshift_r(t32[w], t32[w + 1], s);
}

// n (1..4) hex characters in the low bytes of x (first character lowest) -> value.
// Returns false if any of them is not [0-9a-fA-F].
TI_D bool hex4(uint32_t x, unsigned n, uint32_t &v) {
  const unsigned sh = 8 * (4 - n);
  x = (x << sh) | (0x30303030u & ((1u << sh) - 1));       // right-align, pad with '0'
  const uint32_t y = x | 0x20202020u;
  const uint32_t dig = (x + 0x50505050u) & ~(x + 0x46464646u);  // '0' <= c <= '9'
  const uint32_t alp = (y + 0x1f1f1f1fu) & ~(y + 0x19191919u);  // 'a' <= c|0x20 <= 'f'
  const bool ok = ((dig | alp) & ~x & 0x80808080u) == 0x80808080u;
  const uint32_t d = (x & 0x0f0f0f0fu) + ((x >> 6) & 0x01010101u) * 9u;  // nibble per byte
  const uint32_t t = ((d & 0x000f000fu) << 4) | ((d >> 8) & 0x000f000fu);
  v = ((t & 0xffu) << 8) | ((t >> 16) & 0xffu);
  return ok;
}

struct Lex {
  const uint8_t *t;     // staged text
  const uint32_t *t32;  // same, as aligned words
  const uint32_t *dl;   // delimiter bitmap
  unsigned q, le;       // cursor, offset of the line's '\n'
  // next token [b, b + n); false at the end of the line
  TI_D bool next(unsigned &b, unsigned &n) {
    b = next_zero(dl, q, le);
    const bool have = b < le;
    const unsigned e = have ? next_one(dl, b) : le;
    n = e - b;
    q = e;
    return have;
  }
};

}  // namespace ti
#endif
