// TEST INFRASTRUCTURE -- not a product path (see oracle/README.md).
//
// CPU restatement ("port") of the reference's trace-ingest path, sequential and
// deliberately plain, used (a) as the checker when oracle/_ref cannot be built
// (the GPU box has no /root/reference) and (b) as the spec the CUDA kernels are
// compared with, error flags included.  It is itself pinned against the
// reference-built oracle/_ref/ref_oracle by tests/test_oracle.py.
//
// Each function names the reference code it follows (paths relative to
// /root/reference/gpu-simulator):
//   TP  = trace-parser/trace_parser.cc      TD = trace-driven/trace_driven.cc
//   AHM = gpgpu-sim/src/abstract_hardware_model.cc
//
// Policy on malformed input (SURVEY 7 "hard parts"): every line the NVBit tracer can
// produce is decoded exactly as the reference's stream extraction would; anything
// else raises a TI_EF_* flag instead of reproducing undefined stream states.
#ifndef TI_ORACLE_PORT_CORE_H
#define TI_ORACLE_PORT_CORE_H

#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "../include/traceingest.h"
#include "counters_restate.h"
#include "dump_format.h"

namespace ti_port {

#define TI_ISA_BEGIN static const struct IsaRow { const char *name; uint8_t id[5]; uint8_t cat[5]; } kIsa[] = {
#define TI_ISA_ROW(n, a, b, c, d, e, f, g, h, i, j) {n, {a, b, c, d, e}, {f, g, h, i, j}},
#define TI_ISA_END };
#define TI_SPOP_BEGIN static const uint8_t kSpop[] = {
#define TI_SPOP(i, v) v,
#define TI_SPOP_END };
#include "../accel_sim_framework_distributed_b200/csrc/isa_tables.inc"

// TD:100-118 -- binary version -> ISA table column
static inline int arch_column(unsigned v) {
  switch (v) {
    case 35: return 0;
    case 60: case 61: return 1;
    case 70: return 2;
    case 75: return 3;
    case 80: case 86: return 4;
  }
  return -1;
}

struct Cursor {
  const char *p, *end;
  static bool ws(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }
  // next whitespace-delimited token, as `ss >> std::string` would deliver it
  bool token(const char *&b, const char *&e) {
    while (p < end && ws(*p)) ++p;
    if (p >= end) return false;
    b = p;
    while (p < end && !ws(*p)) ++p;
    e = p;
    return true;
  }
};

static inline bool parse_hex(const char *b, const char *e, int max_digits, uint64_t &v) {
  if (e - b > 2 && b[0] == '0' && (b[1] == 'x' || b[1] == 'X')) b += 2;
  if (b >= e) return false;
  while (e - b > 1 && *b == '0') ++b;  // leading zeros never overflow
  if (e - b > max_digits) return false;
  v = 0;
  for (; b < e; ++b) {
    char c = *b;
    unsigned d;
    if (c >= '0' && c <= '9') d = c - '0';
    else if (c >= 'a' && c <= 'f') d = c - 'a' + 10;
    else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
    else return false;
    v = v << 4 | d;
  }
  return true;
}

// decimal with optional '-', magnitude limited to `max_mag` (2^31 for int, 2^63 for long long)
static inline bool parse_dec(const char *b, const char *e, bool allow_neg, uint64_t max_pos, int64_t &v) {
  bool neg = false;
  if (allow_neg && b < e && *b == '-') { neg = true; ++b; }
  if (b >= e || e - b > 19) return false;
  uint64_t m = 0;
  for (; b < e; ++b) {
    if (*b < '0' || *b > '9') return false;
    m = m * 10 + (uint64_t)(*b - '0');
  }
  if (m > max_pos + (neg ? 1 : 0)) return false;
  v = neg ? (int64_t)(0 - m) : (int64_t)m;
  return true;
}

// "R<digits>[anything]" as sscanf("R%d") reads it (TP:154,163); 0..255 only
static inline bool parse_reg(const char *b, const char *e, unsigned &r) {
  if (e - b < 2 || b[0] != 'R' || b[1] < '0' || b[1] > '9') return false;
  unsigned v = 0;
  for (++b; b < e && *b >= '0' && *b <= '9'; ++b) {
    v = v * 10 + (unsigned)(*b - '0');
    if (v > 255) return false;
  }
  r = v;
  return true;
}

struct Line {  // one instruction line, TP:127-214
  uint32_t pc, mask, ndst, nsrc, dst, src[4];
  const char *opc_b, *opc_e;
  bool has_mem;
  uint32_t width;     // from the opcode string, TP:62-76
  uint64_t addrs[32];
  uint32_t flags;     // TI_EF_*
};

// TP:52-60 -- split on '.', dropping empty tokens
static inline void opcode_tokens(const char *b, const char *e, std::vector<std::string> &out) {
  out.clear();
  const char *s = b;
  for (const char *q = b; q <= e; ++q)
    if (q == e || *q == '.') {
      if (q > s) out.emplace_back(s, q);
      s = q + 1;
    }
}

// TP:62-76
static inline bool width_from_tokens(const std::vector<std::string> &t, uint32_t &w) {
  for (const std::string &s : t) {
    size_t k = (s[0] == 'U') ? 1 : 0;
    if (s.size() == k) continue;
    bool digits = true;
    for (size_t i = k; i < s.size(); ++i) digits &= (s[i] >= '0' && s[i] <= '9');
    if (!digits) continue;
    if (s.size() - k > 9) return false;  // std::stoi would throw
    w = (uint32_t)strtoul(s.c_str() + k, NULL, 10) / 8;
    return true;
  }
  w = 4;
  return true;
}

static inline void parse_inst_line(const char *b, const char *e, unsigned trace_version, Line &L) {
  memset(&L, 0, sizeof L);
  Cursor c = {b, e};
  const char *tb, *te;
  int64_t sv;
  uint64_t uv;
  if (trace_version < 3)  // TP:136-143: four ids, read and dropped
    for (int i = 0; i < 4; ++i)
      if (!c.token(tb, te) || !parse_dec(tb, te, false, 0xffffffffull, sv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  if (!c.token(tb, te) || !parse_hex(tb, te, 8, uv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  L.pc = (uint32_t)uv;
  if (!c.token(tb, te) || !parse_hex(tb, te, 8, uv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  L.mask = (uint32_t)uv;
  if (!c.token(tb, te) || !parse_dec(tb, te, false, 0xffffffffull, sv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  if (sv > 1) { L.flags |= TI_EF_TOO_MANY_REGS; return; }  // TP:151
  L.ndst = (uint32_t)sv;
  for (uint32_t i = 0; i < L.ndst; ++i)
    if (!c.token(tb, te) || !parse_reg(tb, te, L.dst)) { L.flags |= TI_EF_BAD_REG; return; }
  if (!c.token(L.opc_b, L.opc_e)) { L.flags |= TI_EF_TRUNCATED; return; }
  if (L.opc_e - L.opc_b > 255) { L.flags |= TI_EF_LONG_OPCODE; return; }
  if (!c.token(tb, te)) { L.flags |= TI_EF_TRUNCATED; return; }
  if (!parse_dec(tb, te, false, 0xffffffffull, sv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  if (sv > 4) { L.flags |= TI_EF_TOO_MANY_REGS; return; }  // TP:160
  L.nsrc = (uint32_t)sv;
  for (uint32_t i = 0; i < L.nsrc; ++i)
    if (!c.token(tb, te) || !parse_reg(tb, te, L.src[i])) { L.flags |= TI_EF_BAD_REG; return; }
  if (!c.token(tb, te)) { L.flags |= TI_EF_TRUNCATED; return; }
  if (!parse_dec(tb, te, false, 0xffffffffull, sv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  if (sv == 0) return;  // TP:172 -- not a memory instruction; the immediate is never read
  L.has_mem = true;
  std::vector<std::string> toks;
  opcode_tokens(L.opc_b, L.opc_e, toks);
  if (!width_from_tokens(toks, L.width)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  if (L.width > 255) { L.flags |= TI_EF_BAD_WIDTH; return; }
  if (!c.token(tb, te)) { L.flags |= TI_EF_TRUNCATED; return; }
  if (!parse_dec(tb, te, false, 0xffffffffull, sv)) { L.flags |= TI_EF_BAD_NUMBER; return; }
  const uint64_t mode = (uint64_t)sv;
  if (mode == 0) {  // TP:181-188 list: one hex per active lane; a missing token reads as 0
    for (int s = 0; s < 32; ++s) {
      if (!(L.mask >> s & 1)) continue;
      if (!c.token(tb, te)) break;
      if (!parse_hex(tb, te, 16, L.addrs[s])) { L.flags |= TI_EF_BAD_NUMBER; return; }
    }
  } else if (mode == 1) {  // TP:189-195 + 86-105
    uint64_t base = 0;
    int64_t stride = 0;
    if (c.token(tb, te)) {
      if (!parse_hex(tb, te, 16, base)) { L.flags |= TI_EF_BAD_NUMBER; return; }
      if (c.token(tb, te) && !parse_dec(tb, te, true, 0x7fffffffull, stride)) { L.flags |= TI_EF_BAD_NUMBER; return; }
    }
    // first active lane gets the base, every lane of the same contiguous run adds the
    // stride, and everything after the first gap stays 0 -- active or not
    int s = 0;
    while (s < 32 && !(L.mask >> s & 1)) ++s;
    uint64_t a = base;
    for (bool first = true; s < 32 && (L.mask >> s & 1); ++s, first = false) {
      if (!first) a += (uint64_t)stride;
      L.addrs[s] = a;
    }
  } else if (mode == 2) {  // TP:196-209 + 107-125
    uint64_t base = 0;
    if (c.token(tb, te) && !parse_hex(tb, te, 16, base)) { L.flags |= TI_EF_BAD_NUMBER; return; }
    uint64_t last = base;
    bool first = true;
    bool exhausted = false;
    for (int s = 0; s < 32; ++s) {
      if (!(L.mask >> s & 1)) continue;
      if (first) { L.addrs[s] = base; first = false; continue; }
      int64_t d = 0;
      if (!exhausted) {
        if (!c.token(tb, te)) exhausted = true;  // failed extraction leaves 0 (C++11)
        else if (!parse_dec(tb, te, true, 0x7fffffffffffffffull, d)) { L.flags |= TI_EF_BAD_NUMBER; return; }
      }
      last += (uint64_t)d;
      L.addrs[s] = last;
    }
  }
  // any other mode: value-initialised addresses stay 0 (TP:174)
}

}  // namespace ti_port
#endif
