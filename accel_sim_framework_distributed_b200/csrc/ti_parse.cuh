// Thread-per-line tokenizer + instruction decode of the fused trace-ingest kernel.
//
// One thread walks one instruction line held in shared memory and produces the
// 32-byte ti_inst_rec: every field of inst_trace_t::parse_from_string
// (trace_parser.cc:127-214) and of trace_warp_inst_t::parse_from_trace_struct
// (trace_driven.cc:151-381) that does not need the decompressed addresses.  Address
// lists are left for the warp-per-instruction memory phase (ti_mem.cuh): only the
// offset of the first address token and the compression mode are recorded.
//
// Thread-per-line (not warp-per-line) because a ~50-byte line is ~450 scalar
// instructions of inherently sequential work: 32 lines per warp keep all lanes busy,
// a warp per line would idle most of them.
#ifndef TI_PARSE_CUH
#define TI_PARSE_CUH

#include "ti_device.cuh"
#include "ti_lex.cuh"

#define TI_OPCODE_IDS_ONLY
#define TI_ISA_BEGIN
#define TI_ISA_ROW(...)
#define TI_ISA_END
#define TI_SPOP_BEGIN
#define TI_SPOP(...)
#define TI_SPOP_END
#include "isa_tables.inc"
#undef TI_ISA_BEGIN
#undef TI_ISA_ROW
#undef TI_ISA_END
#undef TI_SPOP_BEGIN
#undef TI_SPOP
#undef TI_SPOP_END

namespace ti {

// blank = ' ' or 0x09..0x0d without '\n' (what operator>> skips inside a line); end = blank or '\n'
TI_HD bool is_end(unsigned c) { return c == ' ' || c - 9u <= 4u; }
TI_HD bool is_ws(unsigned c) { return is_end(c) && c != '\n'; }
TI_HD unsigned hexval(unsigned c) {  // 0..15, or >= 16 when not a hex digit
  unsigned d = c - '0';
  if (d <= 9) return d;
  d = (c | 0x20) - 'a';
  return d <= 5 ? d + 10 : 255u;
}

constexpr uint64_t pack8(const char *s) {
  uint64_t v = 0;
  for (int i = 0; i < 8 && s[i]; ++i) v |= (uint64_t)(unsigned char)s[i] << (8 * i);
  return v;
}

// Per-instruction scratch handed to the memory phase through the record's mem_idx
// field (overwritten with the real row index before the record leaves smem).
//   [15:0] smem offset of the first token after the address-mode field
//   [17:16] address mode (0 list, 1 base+stride, 2 base+delta, 3 other)
//   [18] generic LD/ST whose space depends on the first active address
TI_HD uint32_t memtmp_make(unsigned q, unsigned mode, bool resolve) {
  return q | (mode > 3 ? 3u : mode) << 16 | (resolve ? 1u : 0u) << 18;
}

// All loops below have ONE exit and no function returns from inside them: the threads
// of a warp parse 32 different lines, and a `return` in the middle would make the end
// of the function the only reconvergence point -- after the first variable-length token
// every lane would then run alone (measured: 5 of 32 lanes active).  Errors are
// therefore recorded in a flag and parsing simply continues on the (bounded) text.
struct Reader {
  const uint8_t *t;
  int q;
  TI_D unsigned peek() const { return t[q]; }
  TI_D void skip_ws() { while (is_ws(t[q])) ++q; }
  TI_D void skip_token() { while (!is_end(t[q])) ++q; }
  // unsigned decimal, <= 19 digits, value <= max; the token must end right after
  TI_D bool dec(uint64_t max, uint64_t &v) {
    int nd = 0;
    uint64_t m = 0;
    unsigned d;
    while ((d = t[q] - '0') <= 9 && nd < 20) { m = m * 10 + d; ++q; ++nd; }
    v = m;
    return nd > 0 && nd < 20 && is_end(t[q]) && m <= max;
  }
  TI_D bool sdec(uint64_t max_pos, int64_t &v) {
    const bool neg = t[q] == '-';
    q += neg ? 1 : 0;
    uint64_t m;
    const bool ok = dec(max_pos + (neg ? 1 : 0), m);
    v = neg ? (int64_t)(0 - m) : (int64_t)m;
    return ok;
  }
  // hexadecimal with optional 0x, at most `maxd` significant digits
  TI_D bool hex(int maxd, uint64_t &v, bool prefix = true) {
    if (prefix && t[q] == '0' && (t[q + 1] | 0x20) == 'x' && !is_end(t[q + 2])) q += 2;
    uint64_t m = 0;
    int nsig = 0, ntot = 0;
    unsigned d;
    while ((d = hexval(t[q])) < 16 && nsig <= 16) {
      m = m << 4 | d;
      nsig += m ? 1 : 0;
      ++ntot; ++q;
    }
    v = m;
    return ntot > 0 && is_end(t[q]) && nsig <= maxd;
  }
  TI_D bool reg(unsigned &r) {  // "R<digits>..." as sscanf("R%d") reads it, 0..255
    const bool isR = t[q] == 'R';
    q += isR ? 1 : 0;
    unsigned d, v = 0;
    int nd = 0;
    while ((d = t[q] - '0') <= 9 && v <= 255) { v = v * 10 + d; ++q; ++nd; }
    skip_token();
    r = v;
    return isR && nd > 0 && v <= 255;
  }
};

// trace_driven.cc:128-149
TI_HD int oprnd_type_of(unsigned op, unsigned sp) {
  switch (op) {
    case OP_SP: case OP_SFU: case OP_SPEC0 + 1: case OP_SPEC0 + 2: case OP_DP: case OP_LOAD: case OP_STORE: return 1;
    case OP_INTP: case OP_SPEC0 + 3: return 0;
    case OP_ALU:
      if (sp == SP_FP || sp == SP_TEX || sp == SP_OTHER) return 1;
      if (sp == SP_INT) return 0;
      return -1;
  }
  return -1;
}

// ---- token decoders over (offset, length); the fast shapes use SWAR, anything else the
// character loops of Reader (same semantics, see the header comment of ti_lex.cuh)
TI_D bool tok_hex(const Lex &lx, unsigned b, unsigned n, int maxd, uint64_t &v) {
  if (n > 2 && lx.t[b] == '0' && (lx.t[b + 1] | 0x20) == 'x') { b += 2; n -= 2; }
  if (n > (unsigned)maxd) {              // leading zeros or too many digits: exact loop
    Reader rd = {lx.t, (int)b};
    const bool ok = rd.hex(maxd, v, false);
    return ok && rd.q == (int)(b + n);
  }
  uint32_t hv;
  bool ok = true;
  v = 0;
  unsigned k = n & 3;                    // leading partial group, then groups of four
  if (k) { ok = hex4(load4(lx.t32, b), k, hv); v = hv; }
  for (; k < n; k += 4) { ok = hex4(load4(lx.t32, b + k), 4, hv) && ok; v = v << 16 | hv; }
  return ok;
}
TI_D bool tok_dec(const Lex &lx, unsigned b, unsigned n, uint64_t max, uint64_t &v) {
  if (n > 8) {                           // long decimals: exact loop
    Reader rd = {lx.t, (int)b};
    const bool ok = rd.dec(max, v);
    return ok && rd.q == (int)(b + n);
  }
  uint64_t x = load8(lx.t32, b);
  uint32_t m = 0;
  bool ok = n > 0;
  for (unsigned k = 0; k < n; ++k, x >>= 8) { const unsigned d = (unsigned)(x & 0xff) - '0'; ok = ok && d <= 9; m = m * 10 + d; }
  v = m;
  return ok && m <= max;
}
TI_D bool tok_reg(const Lex &lx, unsigned b, unsigned n, unsigned &r) {  // "R<digits>[suffix]", 0..255
  uint64_t x = load8(lx.t32, b);
  const bool isR = (x & 0xff) == 'R';
  x >>= 8;
  unsigned v = 0, nd = 0;
  const unsigned lim = n - 1 < 7 ? n - 1 : 7;
  unsigned d;
  while (nd < lim && (d = (unsigned)(x & 0xff) - '0') <= 9 && v <= 255) { v = v * 10 + d; ++nd; x >>= 8; }
  if (nd == 7) {                         // more than seven digits: exact loop
    Reader rd = {lx.t, (int)b};
    return rd.reg(r);
  }
  r = v;
  return isR && nd > 0 && v <= 255;
}

// 0x80 in every byte of a 64-bit word that equals '.'
TI_HD uint64_t dots64(uint64_t w) {
  const uint64_t y = w ^ 0x2e2e2e2e2e2e2e2eull;
  return ~(((y & 0x7f7f7f7f7f7f7f7full) + 0x7f7f7f7f7f7f7f7full) | y | 0x7f7f7f7f7f7f7f7full);
}
TI_HD uint64_t keep_bytes(uint64_t w, int n) { return n >= 8 ? w : n <= 0 ? 0ull : (w & ((1ull << (8 * n)) - 1)); }

// Modifier scan of a memory instruction's opcode (trace_parser.cc:62-76 for the width,
// trace_driven.cc:272-275 for STRONG + GPU).  Dots are located with a SWAR compare per
// 8-byte window; every dot-separated token of up to 8 characters is tested as a word.
TI_D void opcode_modifiers(const Lex &lx, unsigned b, unsigned n, uint32_t &width, bool &wbad, bool &strong, bool &gpu) {
  width = 4; wbad = strong = gpu = false;
  bool wfound = false;
  unsigned ts = 0;                                      // start of the current token
  for (unsigned w0 = 0; w0 < n; w0 += 8) {
    const uint64_t win = load8(lx.t32, b + w0);
    uint64_t dm = dots64(keep_bytes(win, (int)(n - w0)));
    const bool last_window = w0 + 8 >= n;
    for (;;) {                                          // every token that ends inside this window
      unsigned te;
      if (dm) { te = w0 + ((__ffsll((long long)dm) - 1) >> 3); dm &= dm - 1; }
      else if (last_window) te = n;
      else break;
      const unsigned tl = te - ts;
      if (tl > 0 && tl <= 8) {
        const uint64_t tok = keep_bytes(load8(lx.t32, b + ts), (int)tl);
        if (tl == 6 && tok == pack8("STRONG")) strong = true;
        if (tl == 3 && tok == pack8("GPU")) gpu = true;
        if (!wfound) {
          const unsigned skip = (tok & 0xff) == 'U' ? 1u : 0u;   // "U<digits>" counts like "<digits>"
          uint64_t x = tok >> (8 * skip);
          uint32_t val = 0;
          bool dig = tl > skip;
          for (unsigned k = skip; k < tl; ++k, x >>= 8) { const unsigned d = (unsigned)(x & 0xff) - '0'; dig = dig && d <= 9; val = val * 10 + d; }
          if (dig) { wfound = true; width = val / 8; }
        }
      } else if (tl > 8 && !wfound) {                   // a long token is a width only if all digits: exact loop
        const unsigned skip = lx.t[b + ts] == 'U' ? 1u : 0u;
        bool dig = true;
        uint32_t val = 0;
        for (unsigned k = skip; k < tl; ++k) { const unsigned d = lx.t[b + ts + k] - '0'; dig = dig && d <= 9; if (k - skip < 10) val = val * 10 + d; }
        if (dig) { wfound = true; wbad = tl - skip > 9; width = val / 8; }
      }
      ts = te + 1;
      if (te >= n) break;
    }
  }
}

// ---- everything about an instruction that is a function of its opcode STRING alone (and of the
// launch's constant model parameters), packed into 64 bits so that it can be memoised per CTA
// (ti_fast.cuh) and applied to a line in a few instructions:
//   [7:0] opcode id  [15:8] op_type (after the atomics flip)  [23:16] pack_a (sp_op | oprnd_type+1 << 5)
//   [25:24] memory_op  [26] atomic  [27] const-cache operand  [31:28] space  [34:32] cache_op
//   [39:35] latency slot (folded op_type before the flip)  [40] half-precision initiation /2
//   [41] a width is required  [42] generic LD/ST resolved by its first address  [43] may take the stride path
//   [55:44] width from the modifiers, clipped at 4095  [60:56] token length, clipped at 31  [63] opcode known
constexpr uint64_t kOpfValid = 1ull << 63;
constexpr int kMemoMaxLen = 24;
TI_HD uint32_t opf_width_raw(uint64_t f) { return (uint32_t)(f >> 44) & 0xfffu; }

TI_D uint64_t decode_opcode(const Lex &lx, unsigned ob, int len, const DecodeParams &P, bool &wbad) {
  const uint64_t w0 = keep_bytes(load8(lx.t32, ob), len), w1 = keep_bytes(load8(lx.t32, ob + 8), len - 8);
  int dot = len;
  {
    const uint64_t d0 = dots64(w0), d1 = dots64(w1);
    if (d0) dot = (__ffsll((long long)d0) - 1) >> 3;
    else if (d1) dot = 8 + ((__ffsll((long long)d1) - 1) >> 3);
    else if (len > 16) dot = 16;                       // a mnemonic this long is in no table
  }
  uint32_t width = 4;
  bool strong = false, gpu = false;
  opcode_modifiers(lx, ob, (unsigned)len, width, wbad, strong, gpu);
  // ---- mnemonic -> (opcode id, op_type), trace_driven.cc:188-206
  const bool keyable = dot > 0 && dot <= TI_ISA_MAX_NAME;
  if (!keyable) dot = 1;  // keep the shifts below defined; the lookup result is discarded
  const uint64_t klo = dot >= 8 ? w0 : (w0 & ((1ull << (8 * dot)) - 1));
  const uint32_t khi = dot <= 8 ? 0u : (uint32_t)(w1 & ((1ull << (8 * (dot - 8))) - 1));
  unsigned id = 0, cat = 0;
  for (uint32_t h = isa_hash(klo, khi), n = 0; n < (uint32_t)kIsaSlots; ++n, h = (h + 1) & (kIsaSlots - 1)) {
    const uint4 s = __ldg(reinterpret_cast<const uint4 *>(P.isa) + h);
    const unsigned sid = s.w & 0xff;
    if (!sid) break;
    if (s.x == (uint32_t)klo && s.y == (uint32_t)(klo >> 32) && s.z == khi) { id = sid; cat = (s.w >> 8) & 0xff; break; }
  }
  unsigned op = cat, sp = __ldg(P.spop + id);
  const int ot = oprnd_type_of(op, sp);  // before the overrides below (trace_driven.cc:201)
  if (id == TI_OP_MUFU && w1 == 0 && len == 8) {  // whole-string compares, trace_driven.cc:208-217
    if (w0 == pack8("MUFU.SIN") || w0 == pack8("MUFU.COS")) sp = SP_SIN;
    if (w0 == pack8("MUFU.EX2") || w0 == pack8("MUFU.RCP")) sp = SP_EXP;
    if (w0 == pack8("MUFU.RSQ")) sp = SP_SQRT;
    if (w0 == pack8("MUFU.LG2")) sp = SP_LOG;
  }
  if (id == TI_OP_IMAD && ((len == 8 && w1 == 0 && w0 == pack8("IMAD.MOV")) ||
                           (len == 9 && w0 == pack8("IMAD.IAD") && w1 == pack8("D"))))
    sp = SP_INT;                                  // trace_driven.cc:219-222
  unsigned memory_op = 0, space = SPACE_UNDEF, cache_op = CACHE_UNDEF;
  bool atomic = false, cconst = false, need_w = false, generic = false, half = false;
  switch (id) {                                   // trace_driven.cc:254-378
    case TI_OP_LDC: memory_op = 1; cconst = true; space = SPACE_CONST; cache_op = CACHE_ALL; break;   // data_size = 4
    case TI_OP_LDG: case TI_OP_LDL:
      need_w = true; memory_op = 1; cache_op = (strong && gpu) ? CACHE_GLOBAL : CACHE_ALL;
      space = id == TI_OP_LDL ? SPACE_LOCAL : SPACE_GLOBAL; break;
    case TI_OP_STG: case TI_OP_STL:
      need_w = true; memory_op = 2; cache_op = CACHE_ALL; space = id == TI_OP_STL ? SPACE_LOCAL : SPACE_GLOBAL; break;
    case TI_OP_ATOMG: case TI_OP_RED: case TI_OP_ATOM:
      need_w = true; memory_op = 1; op = OP_LOAD; space = SPACE_GLOBAL; atomic = true; cache_op = CACHE_GLOBAL; break;
    case TI_OP_LDS: need_w = true; memory_op = 1; space = SPACE_SHARED; break;
    case TI_OP_STS: need_w = true; memory_op = 2; space = SPACE_SHARED; break;
    case TI_OP_ATOMS: need_w = true; atomic = true; memory_op = 1; space = SPACE_SHARED; break;
    case TI_OP_LDSM: need_w = true; space = SPACE_SHARED; break;
    case TI_OP_LD: case TI_OP_ST:
      need_w = true; memory_op = id == TI_OP_LD ? 1 : 2;
      if (P.shmem_base == 0 || P.local_base == 0) space = SPACE_SHARED;
      else generic = true;                        // first active address decides (memory phase)
      break;
    case TI_OP_HADD2: case TI_OP_HADD2_32I: case TI_OP_HFMA2: case TI_OP_HFMA2_32I: case TI_OP_HMUL2_32I:
    case TI_OP_HSET2: case TI_OP_HSETP2: half = true; break;
    default: break;
  }
  const bool stridable = space == SPACE_GLOBAL && !atomic && (op == OP_LOAD || op == OP_STORE);
  const unsigned ulen = len > 31 ? 31u : (unsigned)len;
  return (uint64_t)id | (uint64_t)op << 8 | (uint64_t)(sp | (unsigned)(ot + 1) << 5) << 16 | (uint64_t)memory_op << 24 |
         (uint64_t)(atomic ? 1u : 0u) << 26 | (uint64_t)(cconst ? 1u : 0u) << 27 | (uint64_t)space << 28 | (uint64_t)cache_op << 32 |
         (uint64_t)fold_op(cat) << 35 | (uint64_t)(half ? 1u : 0u) << 40 | (uint64_t)(need_w ? 1u : 0u) << 41 |
         (uint64_t)(generic ? 1u : 0u) << 42 | (uint64_t)(stridable ? 1u : 0u) << 43 | (uint64_t)(width > 4095 ? 4095u : width) << 44 |
         (uint64_t)ulen << 56 | ((id && keyable) ? kOpfValid : 0ull);
}

// The per-line half, in record words (a record is eight 32-bit words, see ti_inst_rec):
//   w3 = mem_idx scratch (memtmp_make) or 0xFFFFFFFF   w5 = src[3] | opcode_id << 8 | op << 16 | pack_a << 24
//   w6 = pack_b | pack_c << 8 | data_size << 16 (cycles = 0)   w7 = latency | initiation_interval << 16 | opcode_len << 24
// Returns TI_EF_* for the two conditions that depend on the line as well as on the opcode.
struct OpWords { uint32_t w3, w5hi, w6, w7; unsigned kind; };
TI_D uint32_t opcode_words(OpWords &o, uint64_t f, uint32_t mask, bool has_mem, unsigned mode, unsigned q, unsigned ndst, unsigned nsrc,
                           unsigned opcode_len, const DecodeParams &P) {
  const unsigned lo = (unsigned)f, hi = (unsigned)(f >> 32);
  const unsigned op = lo >> 8 & 0xff, memory_op = lo >> 24 & 3, atomic = lo >> 26 & 1, cconst = lo >> 27 & 1, space = lo >> 28 & 0xf;
  const unsigned cache_op = hi & 7, slot = hi >> 3 & 31, half = hi >> 8 & 1, need_w = hi >> 9 & 1, generic = hi >> 10 & 1, stridable = hi >> 11 & 1;
  const unsigned width = hi >> 12 & 0xfff;
  const unsigned data_size = cconst ? 4u : has_mem ? width : 0u;
  uint32_t fl = 0;
  if (need_w && data_size == 0) fl = TI_EF_ZERO_WIDTH;
  // a ld/st-class opcode without addresses trips the reference's assert at
  // abstract_hardware_model.cc:296 as soon as one lane is active
  else if (!has_mem && mask && (op == OP_LOAD || op == OP_STORE || op == OP_TLOAD || op == OP_TSTORE)) fl = TI_EF_UNDEF_SPACE;
  const bool resolve = generic && has_mem && mask != 0;
  const unsigned lat = P.lat[slot][0], init = P.lat[slot][1] >> half;   // trace_driven.cc:243, 366-375
  o.w5hi = (lo & 0xffffffu) << 8;                                       // opcode_id, op, pack_a
  o.w6 = (nsrc | ndst << 3 | memory_op << 4 | atomic << 6 | cconst << 7) | (space | cache_op << 4 | (has_mem ? 1u : 0u) << 7) << 8 | (data_size & 0xff) << 16;
  o.w7 = (lat & 0xffff) | (init & 0xff) << 16 | (opcode_len & 0xff) << 24;
  o.w3 = has_mem ? memtmp_make(q, mode, resolve) : 0xFFFFFFFFu;
  // base+stride global loads/stores take the thread-per-instruction path (ti_mem.cuh), and so does the form the tracer
  // writes for ONE active lane: base+delta without deltas (its base+stride compressor needs a second lane,
  // tracer_tool.cu:609-614), which is a run of one
  const bool one_lane = mode == 2 && mask != 0 && (mask & (mask - 1)) == 0;
  o.kind = has_mem ? (((mode == 1 || one_lane) && stridable) ? 3u : 1u) : 0u;
  return fl;
}

// the same, into the fields of a record (r.mask, r.src[3] and r.opcode_len must be set)
TI_D uint32_t apply_opcode_facts(ti_inst_rec &r, unsigned &kind, uint64_t f, bool has_mem, unsigned mode, unsigned q, unsigned ndst,
                                 unsigned nsrc, const DecodeParams &P) {
  OpWords o;
  const uint32_t fl = opcode_words(o, f, r.mask, has_mem, mode, q, ndst, nsrc, r.opcode_len, P);
  r.opcode_id = (uint8_t)(o.w5hi >> 8); r.op = (uint8_t)(o.w5hi >> 16); r.pack_a = (uint8_t)(o.w5hi >> 24);
  r.pack_b = (uint8_t)o.w6; r.pack_c = (uint8_t)(o.w6 >> 8); r.data_size = (uint8_t)(o.w6 >> 16);
  r.latency = (uint16_t)o.w7; r.initiation_interval = (uint8_t)(o.w7 >> 16);
  if (has_mem) r.mem_idx = o.w3;
  kind = o.kind;
  return fl;
}

// Decodes the instruction line at smem offsets [q0, le).  Returns TI_EF_* flags.
// `kind` receives bit0 = has addresses (mem_width > 0), bit1 = eligible for the stride path.
TI_D uint32_t parse_inst_line(const uint8_t *text, const uint32_t *dlmap32, int q0, unsigned le, const DecodeParams &P,
                              uint32_t chunk_text_base, ti_inst_rec &r, unsigned &kind, uint64_t &memo_facts) {
#define TI_FAIL(cond, code) do { if ((cond) && !fl) fl = (code); } while (0)
  Lex lx = {text, reinterpret_cast<const uint32_t *>(text), dlmap32, (unsigned)q0, le};
  uint32_t fl = 0;  // first error wins, as in the sequential reference
  uint64_t u = 0;
  unsigned b, n;
  kind = 0;
  memo_facts = 0;
  r.pc = r.mask = r.opc_off = 0; r.mem_idx = 0xFFFFFFFFu;
  r.dst = 0; r.src[0] = r.src[1] = r.src[2] = r.src[3] = 0;
  r.opcode_id = 0; r.op = OP_ALU; r.pack_a = (uint8_t)(SP_OTHER | (0u << 5)); r.pack_b = r.pack_c = 0;
  r.data_size = r.cycles = 0; r.latency = 0; r.initiation_interval = 0; r.opcode_len = 0;

  if (P.trace_version < 3)  // trace_parser.cc:136-143 (warp-uniform)
    for (int i = 0; i < 4; ++i) { const bool have = lx.next(b, n); TI_FAIL(!have || !tok_dec(lx, b, n, 0xffffffffull, u), TI_EF_BAD_NUMBER); }
  bool have = lx.next(b, n);
  TI_FAIL(!have || !tok_hex(lx, b, n, 8, u), TI_EF_BAD_NUMBER);
  r.pc = (uint32_t)u;
  have = lx.next(b, n);
  TI_FAIL(!have || !tok_hex(lx, b, n, 8, u), TI_EF_BAD_NUMBER);
  r.mask = (uint32_t)u;
  have = lx.next(b, n);
  TI_FAIL(!have || !tok_dec(lx, b, n, 0xffffffffull, u), TI_EF_BAD_NUMBER);
  TI_FAIL(u > 1, TI_EF_TOO_MANY_REGS);
  const unsigned ndst = u ? 1u : 0u;
  unsigned rv = 0;
  if (ndst) { have = lx.next(b, n); TI_FAIL(!have || !tok_reg(lx, b, n, rv), TI_EF_BAD_REG); r.dst = (uint8_t)rv; }

  // ---- opcode token
  have = lx.next(b, n);
  TI_FAIL(!have, TI_EF_TRUNCATED);
  const unsigned ob = b;
  const int len = (int)n;
  TI_FAIL(len > 255, TI_EF_LONG_OPCODE);
  r.opc_off = chunk_text_base + ob;
  r.opcode_len = (uint8_t)len;

  have = lx.next(b, n);
  TI_FAIL(!have, TI_EF_TRUNCATED);
  TI_FAIL(!tok_dec(lx, b, n, 0xffffffffull, u), TI_EF_BAD_NUMBER);
  TI_FAIL(u > 4, TI_EF_TOO_MANY_REGS);
  const unsigned nsrc = u > 4 ? 4u : (unsigned)u;
  for (unsigned i = 0; i < nsrc; ++i) { have = lx.next(b, n); TI_FAIL(!have || !tok_reg(lx, b, n, rv), TI_EF_BAD_REG); r.src[i] = (uint8_t)rv; }
  have = lx.next(b, n);
  TI_FAIL(!have, TI_EF_TRUNCATED);
  TI_FAIL(!tok_dec(lx, b, n, 0xffffffffull, u), TI_EF_BAD_NUMBER);
  const bool has_mem = !fl && u != 0;
  unsigned mode = 0;
  // ---- opcode string -> everything that is a function of it alone (decode_opcode), then the
  // per-line part (apply_opcode_facts); the fast path (ti_fast.cuh) shares the second half
  bool wbad = false;
  const uint64_t facts = decode_opcode(lx, ob, len, P, wbad);
  TI_FAIL(has_mem && wbad, TI_EF_BAD_NUMBER);
  TI_FAIL(has_mem && opf_width_raw(facts) > 255, TI_EF_BAD_WIDTH);
  if (has_mem) {
    have = lx.next(b, n);
    TI_FAIL(!have, TI_EF_TRUNCATED);
    TI_FAIL(!tok_dec(lx, b, n, 0xffffffffull, u), TI_EF_BAD_NUMBER);
    mode = u > 3 ? 3u : (unsigned)u;
  }
  TI_FAIL(!(facts & kOpfValid), TI_EF_UNKNOWN_OPCODE);
  const uint32_t fl2 = apply_opcode_facts(r, kind, facts, has_mem, mode, lx.q, ndst, nsrc, P);
  TI_FAIL(fl2 != 0, fl2);
  memo_facts = (!wbad && opf_width_raw(facts) <= 255 && len <= kMemoMaxLen) ? facts : 0ull;
  if (fl) kind = 0;
  return fl;
#undef TI_FAIL
}

}  // namespace ti
#endif
