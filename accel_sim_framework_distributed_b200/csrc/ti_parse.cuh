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

TI_HD bool is_ws(unsigned c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
TI_HD bool is_end(unsigned c) { return c == '\n' || is_ws(c); }
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
  TI_D bool hex(int maxd, uint64_t &v) {
    if (t[q] == '0' && (t[q + 1] | 0x20) == 'x' && !is_end(t[q + 2])) q += 2;
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

// Decodes the instruction line starting at smem offset q0.  Returns TI_EF_* flags.
// `kind` receives bit0 = has addresses (mem_width > 0), bit1 = eligible for the stride path.
TI_D uint32_t parse_inst_line(const uint8_t *text, int q0, const DecodeParams &P, uint32_t chunk_text_base,
                              ti_inst_rec &r, unsigned &kind) {
#define TI_FAIL(cond, code) do { if ((cond) && !fl) fl = (code); } while (0)
  Reader rd = {text, q0};
  uint32_t fl = 0;  // first error wins, as in the sequential reference
  uint64_t u = 0;
  kind = 0;
  r.pc = r.mask = r.opc_off = 0; r.mem_idx = 0xFFFFFFFFu;
  r.dst = 0; r.src[0] = r.src[1] = r.src[2] = r.src[3] = 0;
  r.opcode_id = 0; r.op = OP_ALU; r.pack_a = (uint8_t)(SP_OTHER | (0u << 5)); r.pack_b = r.pack_c = 0;
  r.data_size = r.cycles = 0; r.latency = 0; r.initiation_interval = 0; r.opcode_len = 0;

  if (P.trace_version < 3)  // trace_parser.cc:136-143 (warp-uniform)
    for (int i = 0; i < 4; ++i) { rd.skip_ws(); TI_FAIL(!rd.dec(0xffffffffull, u), TI_EF_BAD_NUMBER); }
  rd.skip_ws(); TI_FAIL(!rd.hex(8, u), TI_EF_BAD_NUMBER);
  r.pc = (uint32_t)u;
  rd.skip_ws(); TI_FAIL(!rd.hex(8, u), TI_EF_BAD_NUMBER);
  r.mask = (uint32_t)u;
  rd.skip_ws(); TI_FAIL(!rd.dec(0xffffffffull, u), TI_EF_BAD_NUMBER);
  TI_FAIL(u > 1, TI_EF_TOO_MANY_REGS);
  const unsigned ndst = u ? 1u : 0u;
  unsigned rv = 0;
  if (ndst) { rd.skip_ws(); TI_FAIL(!rd.reg(rv), TI_EF_BAD_REG); r.dst = (uint8_t)rv; }

  // ---- opcode token: mnemonic key, first 16 bytes, width / STRONG / GPU modifiers
  rd.skip_ws();
  TI_FAIL(rd.peek() == '\n', TI_EF_TRUNCATED);
  const int qo = rd.q;
  uint64_t w0 = 0, w1 = 0, tok8 = 0;
  int len = 0, dot = -1, tlen = 0;
  bool alldig = true, restdig = true, firstU = false, wfound = false, wbad = false, strong = false, gpu = false;
  uint32_t tval = 0, width = 4;
  int tdig = 0;
  for (;; ++rd.q) {
    const unsigned c = rd.peek();
    const bool end = is_end(c);
    if (end || c == '.') {
      if (tlen > 0) {  // token boundary (empty tokens are dropped, trace_parser.cc:56-58)
        if (!wfound) {
          const bool num = alldig, unum = firstU && tlen > 1 && restdig;
          if (num || unum) { wfound = true; wbad = tdig > 9; width = tval / 8; }
        }
        if (tlen == 6 && tok8 == pack8("STRONG")) strong = true;
        if (tlen == 3 && tok8 == pack8("GPU")) gpu = true;
      }
      if (end) break;
      if (dot < 0) dot = len;
      tlen = 0; tok8 = 0; alldig = restdig = true; firstU = false; tval = 0; tdig = 0;
    } else {
      const unsigned d = c - '0';
      if (tlen == 0) { firstU = c == 'U'; alldig = d <= 9; }
      else { alldig = alldig && d <= 9; restdig = restdig && d <= 9; }
      if (d <= 9) { if (tdig < 10) tval = tval * 10 + d; ++tdig; }
      if (tlen < 8) tok8 |= (uint64_t)c << (8 * tlen);
      ++tlen;
    }
    if (len < 8) w0 |= (uint64_t)c << (8 * len);
    else if (len < 16) w1 |= (uint64_t)c << (8 * (len - 8));
    ++len;
  }
  if (dot < 0) dot = len;
  TI_FAIL(len > 255, TI_EF_LONG_OPCODE);
  r.opc_off = chunk_text_base + (uint32_t)qo;
  r.opcode_len = (uint8_t)len;

  rd.skip_ws();
  TI_FAIL(rd.peek() == '\n', TI_EF_TRUNCATED);
  TI_FAIL(!rd.dec(0xffffffffull, u), TI_EF_BAD_NUMBER);
  TI_FAIL(u > 4, TI_EF_TOO_MANY_REGS);
  const unsigned nsrc = u > 4 ? 4u : (unsigned)u;
  for (unsigned i = 0; i < nsrc; ++i) { rd.skip_ws(); TI_FAIL(!rd.reg(rv), TI_EF_BAD_REG); r.src[i] = (uint8_t)rv; }
  rd.skip_ws();
  TI_FAIL(rd.peek() == '\n', TI_EF_TRUNCATED);
  TI_FAIL(!rd.dec(0xffffffffull, u), TI_EF_BAD_NUMBER);
  const bool has_mem = !fl && u != 0;
  unsigned mode = 0;
  if (has_mem) {
    TI_FAIL(wbad, TI_EF_BAD_NUMBER);
    TI_FAIL(width > 255, TI_EF_BAD_WIDTH);
    rd.skip_ws();
    TI_FAIL(rd.peek() == '\n', TI_EF_TRUNCATED);
    TI_FAIL(!rd.dec(0xffffffffull, u), TI_EF_BAD_NUMBER);
    mode = u > 3 ? 3u : (unsigned)u;
    kind = 1;
  }

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
  TI_FAIL(!id || !keyable, TI_EF_UNKNOWN_OPCODE);
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
  unsigned lat = P.lat[fold_op(cat)][0], init = P.lat[fold_op(cat)][1];  // trace_driven.cc:243
  unsigned data_size = has_mem ? width : 0, memory_op = 0, space = SPACE_UNDEF, cache_op = CACHE_UNDEF;
  bool atomic = false, cconst = false, need_w = false, resolve = false;
  switch (id) {                                   // trace_driven.cc:254-378
    case TI_OP_LDC: data_size = 4; memory_op = 1; cconst = true; space = SPACE_CONST; cache_op = CACHE_ALL; break;
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
      else resolve = has_mem && r.mask != 0;      // first active address decides (memory phase)
      break;
    case TI_OP_HADD2: case TI_OP_HADD2_32I: case TI_OP_HFMA2: case TI_OP_HFMA2_32I: case TI_OP_HMUL2_32I:
    case TI_OP_HSET2: case TI_OP_HSETP2: init /= 2; break;
    default: break;
  }
  TI_FAIL(need_w && data_size == 0, TI_EF_ZERO_WIDTH);
  // a ld/st-class opcode without addresses trips the reference's assert at
  // abstract_hardware_model.cc:296 as soon as one lane is active
  TI_FAIL(!has_mem && r.mask && (op == OP_LOAD || op == OP_STORE || op == OP_TLOAD || op == OP_TSTORE), TI_EF_UNDEF_SPACE);
  r.opcode_id = (uint8_t)id; r.op = (uint8_t)op;
  r.pack_a = (uint8_t)(sp | (unsigned)(ot + 1) << 5);
  r.pack_b = (uint8_t)(nsrc | ndst << 3 | memory_op << 4 | (atomic ? 1u : 0u) << 6 | (cconst ? 1u : 0u) << 7);
  r.pack_c = (uint8_t)(space | cache_op << 4 | (has_mem ? 1u : 0u) << 7);
  r.data_size = (uint8_t)data_size;
  r.latency = (uint16_t)lat; r.initiation_interval = (uint8_t)init;
  if (has_mem) r.mem_idx = memtmp_make((unsigned)rd.q, mode, resolve);
  // base+stride global loads/stores take the thread-per-instruction path (ti_mem.cuh)
  if (has_mem && mode == 1 && space == SPACE_GLOBAL && !atomic && (op == OP_LOAD || op == OP_STORE)) kind |= 2;
  if (fl) kind = 0;
  return fl;
#undef TI_FAIL
}

}  // namespace ti
#endif
